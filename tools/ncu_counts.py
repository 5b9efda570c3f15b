#!/usr/bin/env python
"""Turn the ncu raw pages of the bench workload's two kernels into profiles/kernel_counts.json, the per-launch
counters bench.py divides by LIVE kernel times for its roofline (read here, no GPU needed).  The file carries the
hash of the kernel sources it was captured on; bench.py emits `stale: true` when the built sources differ.

    python tools/ncu_counts.py gpurun_out/prof_adj_r12.ncu-rep gpurun_out/prof_fwd_r12.ncu-rep r12 [gpurun_out/bench_r12.json] > profiles/kernel_counts.json
"""
import csv
import io
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

PIPES = (("fmaheavy", "sm__inst_executed_pipe_fmaheavy.sum.pct_of_peak_sustained_active"),
         ("fmalite", "sm__inst_executed_pipe_fmalite.sum.pct_of_peak_sustained_active"),
         ("fma", "sm__inst_executed_pipe_fma.sum.pct_of_peak_sustained_active"),
         ("alu", "sm__inst_executed_pipe_alu.sum.pct_of_peak_sustained_active"),
         ("xu", "sm__inst_executed_pipe_xu.sum.pct_of_peak_sustained_active"),
         ("lsu", "sm__inst_executed_pipe_lsu.sum.pct_of_peak_sustained_active"),
         ("issue_active", "smsp__issue_active.avg.pct_of_peak_sustained_active"))
BYTE_UNITS = {"byte": 1, "bytes": 1, "kbyte": 1e3, "mbyte": 1e6, "gbyte": 1e9, "tbyte": 1e12}


def kernel(rep):
    out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    hdr, units, row = rows[0], rows[1], rows[2]

    def g(k, scale_bytes=False):
        if k not in hdr:
            return None
        i = hdr.index(k)
        try:
            v = float(row[i].replace(",", ""))
        except ValueError:
            return None
        return v * BYTE_UNITS.get(units[i].lower(), 1) if scale_bytes else v

    dr, dw = g("dram__bytes_read.sum", True), g("dram__bytes_write.sum", True)
    return {
        "name": row[hdr.index("Kernel Name")],
        "warp_inst": int(g("smsp__inst_executed.sum")),
        "threads_per_inst": g("smsp__thread_inst_executed_per_inst_executed.ratio"),
        "pipes_pct": {s: g(k) for s, k in PIPES if g(k) is not None},
        "registers": g("launch__registers_per_thread"),
        "ncu_duration_ns": g("gpu__time_duration.sum"),
        "dram_bytes": (dr or 0) + (dw or 0),
        "l1_red_sectors": g("l1tex__t_sectors_pipe_lsu_mem_global_op_red.sum"),
        "l1_ld_sectors": g("l1tex__t_sectors_pipe_lsu_mem_global_op_ld.sum"),
        "lts_sectors": g("lts__t_sectors.sum"),
        "l1_hit_pct": g("l1tex__t_sector_hit_rate.pct"), "l2_hit_pct": g("lts__t_sector_hit_rate.pct"),
    }


def main():
    adj, fwd, tag = sys.argv[1], sys.argv[2], sys.argv[3]
    import bench
    out = {"csrc_sha": bench.csrc_sha(),
           "source": f"profiles/{tag}_adj_full.txt, profiles/{tag}_fwd_full.txt (ncu --set full --clock-control none of `python bench.py "
                     "--steps 2 --warmup 3 --no-cpu --no-configs`, one launch each; tools/ncu_counts.py)",
           "backtrace_kernel": kernel(adj), "trace_kernel": kernel(fwd)}
    if len(sys.argv) > 4:  # the bench line of the same build: ray-steps per launch -> instructions per warp-step
        line = json.loads(open(sys.argv[4]).read().strip().splitlines()[-1])
        cfg = line["config"]
        sf = cfg.get("ray_steps_fwd_rank0", cfg.get("ray_steps_fwd")); sb = cfg.get("ray_steps_adj_rank0", cfg.get("ray_steps_adj"))
        out["trace_kernel"]["ray_steps"] = sf; out["backtrace_kernel"]["ray_steps"] = sb
        out["trace_kernel"]["warp_inst_per_warp_step"] = out["trace_kernel"]["warp_inst"] / (sf / 32.0)
        out["backtrace_kernel"]["warp_inst_per_warp_step"] = out["backtrace_kernel"]["warp_inst"] / (sb / 32.0)
    print(json.dumps(out, indent=1))


if __name__ == "__main__":
    main()
