#!/usr/bin/env python
"""bench.py -- fwd+adjoint ray-steps/s of the eikonal tracer on B200 (BASELINE.json metric).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]

One "step" = one forward trace + one adjoint backtrace (+ gradient all-reduce when
N > 1) over one ray batch.  Workload (config.workload): BASELINE.json configs[3]
"caustic design": 256^3 IOR grid (Gaussian-bump field A of SURVEY.md section 8d), 2048^2 =
4 194 304 collimated rays per GPU, ds = h/2, adjoint seeds dx = dv = 1
(reference src/test.cpp:142-144).  Weak scaling: every rank traces its own 4M-ray batch.

Printed JSON (one line, rank 0): see the task contract; `roofline`, `cpu_baseline`,
`e2e`, `clocks`, `gpu_launches` are included.
"""
from __future__ import annotations

import argparse
import json
import os
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

import numpy as np  # noqa: E402

NG = 256
SENSOR = 2048
METRIC = "fwd+adjoint ray-steps/sec at 256^3 grid"
UNIT = "ray-steps/s"
BYTES_FWD, BYTES_ADJ = 32, 64  # algorithmic bytes per ray-step, SURVEY.md section 8d
WORKLOAD = (f"caustic-256 (BASELINE.json configs[3]): {NG}^3 bump IOR grid, {SENSOR}^2 = {SENSOR * SENSOR} "
            "collimated rays per GPU, ds = h/2, seeds dx=dv=1")


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        with open(p) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    return 6650.0, "fallback (B200_PROFILING.md)"


def workload_host(n_g=None, sensor=None):
    from adjointnonlinearraytracing_b200 import scenes as sc
    n_g = NG if n_g is None else n_g
    sensor = SENSOR if sensor is None else sensor
    rif = sc.field_bump(n_g)
    h = float(sc.grid_spacing(n_g))
    ds = h / 2
    pos, vel = sc.rays_collimated(sensor)
    return rif, h, ds, pos, vel


# --------------------------------------------------------------------------- clocks
class ClockSampler:
    """Samples SM clock and throttle reasons of one GPU during the timed region (NVML)."""

    def __init__(self, index):
        self.index, self.samples, self.reasons, self.max_mhz = index, [], set(), None
        self._stop = threading.Event()
        self._thr = None
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
        except Exception:
            self.nv = None

    def _run(self):
        nv = self.nv
        names = {
            "hw_slowdown": getattr(nv, "nvmlClocksEventReasonHwSlowdown", 0x8),
            "hw_thermal_slowdown": getattr(nv, "nvmlClocksEventReasonHwThermalSlowdown", 0x40),
            "sw_thermal_slowdown": getattr(nv, "nvmlClocksEventReasonSwThermalSlowdown", 0x20),
            "sw_power_cap": getattr(nv, "nvmlClocksEventReasonSwPowerCap", 0x4),
        }
        while not self._stop.is_set():
            try:
                self.samples.append(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM))
                try:
                    r = nv.nvmlDeviceGetCurrentClocksEventReasons(self.h)
                except Exception:
                    r = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                for k, bit in names.items():
                    if r & bit:
                        self.reasons.add(k)
            except Exception:
                pass
            time.sleep(0.01)

    def start(self):
        if self.nv:
            self._thr = threading.Thread(target=self._run, daemon=True)
            self._thr.start()

    def stop(self):
        if self._thr:
            self._stop.set()
            self._thr.join()
        s = sorted(self.samples)
        return {"sm_mhz": (s[len(s) // 2] if s else None), "sm_max_mhz": self.max_mhz,
                "reasons": sorted(self.reasons), "samples": len(s)}


# --------------------------------------------------------------------------- CPU reference arm
def _cpu_lib():
    """The reference's own CPU tracer (oracle/_ref) when its build is present, else the C port."""
    import oracle
    if oracle.ref is not None:
        return oracle.ref, "reference", oracle
    return oracle.port, "port", oracle


def _threads_run(fn, n):
    out, thr = [None] * n, []
    def work(i):
        out[i] = fn(i)
    for i in range(n):
        t = threading.Thread(target=work, args=(i,))
        t.start(); thr.append(t)
    for t in thr:
        t.join()
    return out


class CpuArm:
    """Times the reference CPU path on `threads` host threads: rays are split into one strided
    sub-batch per thread, each thread owns a private gradient grid (the same scheme as the
    multi-GPU path; the reference itself is single-threaded)."""

    def __init__(self, rays_per_thread=None, threads=None, target_s=2.5):
        self.lib, self.kind, self.oracle = _cpu_lib()
        self.rif, self.h, self.ds, pos, vel = workload_host()
        self.threads = threads or max(1, len(os.sched_getaffinity(0)))
        if rays_per_thread is None:  # pilot: size the sample so one step takes ~target_s per thread
            p = pos[:: max(1, len(pos) // 1024)][:1024]
            v = vel[:: max(1, len(vel) // 1024)][:1024]
            t0 = time.perf_counter()
            xt, vt = self.lib.trace(self.rif, p, v, np.float32(self.h), np.float32(self.ds))
            self.lib.backtrace(self.rif, xt, vt, np.ones_like(p), np.ones_like(p), np.float32(self.h), np.float32(self.ds))
            dt = max(time.perf_counter() - t0, 1e-3)
            rays_per_thread = int(min(65536, max(4096, 1024 * target_s / dt)))  # >= 4096 amortises the 64 MB grid copies
        self.rpt = rays_per_thread
        n = self.rpt * self.threads
        stride = max(1, len(pos) // n)
        self.pos = [np.ascontiguousarray(pos[i::stride * self.threads][: self.rpt]) for i in range(self.threads)]
        self.vel = [np.ascontiguousarray(vel[i::stride * self.threads][: self.rpt]) for i in range(self.threads)]
        self.steps = None

    def count_steps(self):
        port = self.oracle.port
        h, ds = np.float32(self.h), np.float32(self.ds)
        import ctypes  # each thread needs its own wrapper state (last_steps)
        def one(i):
            lib = self.oracle._Lib(port.path, "orc_", np.float32, returns_steps=True)
            xt, vt = lib.trace(self.rif, self.pos[i], self.vel[i], h, ds); sf = lib.last_steps
            lib.backtrace(self.rif, xt, vt, np.ones_like(xt), np.ones_like(xt), h, ds); sb = lib.last_steps
            return sf, sb
        r = _threads_run(one, self.threads)
        self.steps = (sum(a for a, _ in r), sum(b for _, b in r))
        return self.steps

    def step(self):
        h, ds = np.float32(self.h), np.float32(self.ds)
        def one(i):
            xt, vt = self.lib.trace(self.rif, self.pos[i], self.vel[i], h, ds)
            return self.lib.backtrace(self.rif, xt, vt, np.ones_like(xt), np.ones_like(xt), h, ds)
        t0 = time.perf_counter()
        _threads_run(one, self.threads)
        return time.perf_counter() - t0

    def describe(self):
        return (f"{self.threads} threads x {self.rpt} rays (strided subsample of the {SENSOR}^2 batch), "
                f"{NG}^3 grid, one trace + one backtrace per thread, private gradient grids")


def run_reference_arm(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    arm = CpuArm()
    sf, sb = arm.count_steps()
    for _ in range(args.warmup):
        arm.step()
    times = [arm.step() for _ in range(args.steps)]
    total = sum(times)
    value = (sf + sb) * args.steps / total
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * total / args.steps,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": WORKLOAD, "rays_per_gpu": SENSOR * SENSOR, "grid": [NG, NG, NG],
                   "note": "CPU reference arm: each step is a bounded strided sample of the batch (cpu_baseline.sample)"},
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": arm.threads, "kind": arm.kind, "sample": arm.describe()},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    _emit(line)


# --------------------------------------------------------------------------- GPU arm
def run_ours(args):
    import torch
    import torch.distributed as dist
    import __graft_entry__ as ge
    ge.build()
    from adjointnonlinearraytracing_b200 import _capi, ops
    from adjointnonlinearraytracing_b200.dist import GradAllReduce

    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device (there is no CPU fallback; use --impl reference for the CPU arm)")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    allreduce = GradAllReduce() if world > 1 else None

    rif_h, h, ds, pos_h, vel_h = workload_host()
    rif = torch.from_numpy(rif_h).to(dev)
    pos_pin = torch.from_numpy(pos_h).pin_memory()
    vel_pin = torch.from_numpy(vel_h).pin_memory()
    pos, vel = pos_pin.to(dev), vel_pin.to(dev)
    n_rays = pos.shape[0]
    dx = torch.ones_like(pos); dv = torch.ones_like(pos)
    grad = torch.empty_like(rif)
    stats = ops.new_stats(dev)
    flags = 0

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    ev = lambda: torch.cuda.Event(enable_timing=True)

    def step(timers=None):
        if timers is not None:
            timers[0].record()
        xt, vt = ops.trace(rif, pos, vel, h, ds, flags=flags, stats=stats)
        if timers is not None:
            timers[1].record()
        ops.backtrace(rif, xt, vt, dx, dv, h, ds, flags=flags, stats=stats, out=grad)
        if timers is not None:
            timers[2].record()
        if allreduce is not None:
            allreduce(grad)
        return xt, vt

    # ---- device-resident timing ------------------------------------------------
    for _ in range(max(args.warmup, 3)):
        step()
    barrier()
    stats.zero_()
    timers = [[ev(), ev(), ev()] for _ in range(args.steps)]
    launches0 = _capi.launch_count()
    sampler = ClockSampler(local)
    t0, t1 = ev(), ev()
    barrier()
    sampler.start()
    t0.record()
    for k in range(args.steps):
        step(timers[k])
    t1.record()
    barrier()
    clocks = sampler.stop()
    launches = _capi.launch_count() - launches0
    ms = t0.elapsed_time(t1)
    steps_total = int(stats[0].item())  # live ray-steps of this rank over the K steps
    fwd_ms = sum(t[0].elapsed_time(t[1]) for t in timers) / args.steps
    adj_ms = sum(t[1].elapsed_time(t[2]) for t in timers) / args.steps

    # ray-step split fwd / adjoint (one extra untimed pass each)
    stats.zero_(); xt, vt = ops.trace(rif, pos, vel, h, ds, flags=flags, stats=stats); s_f = int(stats[0].item())
    stats.zero_(); ops.backtrace(rif, xt, vt, dx, dv, h, ds, flags=flags, stats=stats, out=grad); s_b = int(stats[0].item())
    not_exited = int(stats[1].item())

    # ---- secondary, labelled: opt-in polynomial arithmetic (NOT a parity mode)
    fast = _capi.DRRT_FAST_ARITH
    fe = [ev(), ev()]
    for k in range(4):
        if k == 1:
            fe[0].record()
        xf_, vf_ = ops.trace(rif, pos, vel, h, ds, flags=fast)
        ops.backtrace(rif, xf_, vf_, dx, dv, h, ds, flags=fast, out=grad)
    fe[1].record(); torch.cuda.synchronize()
    fast_ms = fe[0].elapsed_time(fe[1]) / 3
    del xf_, vf_

    # ---- end to end: host ray batch (pinned) -> H2D -> trace -> backtrace -> allreduce -> D2H ----
    grad_pin = torch.empty(rif.shape, dtype=torch.float32).pin_memory()
    stat_pin = torch.empty(2, dtype=torch.int64).pin_memory()

    def e2e_step():
        p = pos_pin.to(dev, non_blocking=True)
        v = vel_pin.to(dev, non_blocking=True)
        stats.zero_()
        xt_, vt_ = ops.trace(rif, p, v, h, ds, flags=flags, stats=stats)
        ops.backtrace(rif, xt_, vt_, dx, dv, h, ds, flags=flags, stats=stats, out=grad)
        if allreduce is not None:
            allreduce(grad)
        grad_pin.copy_(grad, non_blocking=True)
        stat_pin.copy_(stats, non_blocking=True)

    for _ in range(3):
        e2e_step()
    barrier()
    e0, e1 = ev(), ev()
    e0.record()
    for _ in range(args.steps):
        e2e_step()
    e1.record()
    barrier()
    e2e_serial_ms = e0.elapsed_time(e1)

    # the same step through the package's streaming API (pipeline.HostStepPipeline): copies of
    # batch k+1 / gradient k-1 overlap the kernels of batch k; every step's gradient is read on the host
    from adjointnonlinearraytracing_b200.pipeline import HostStepPipeline
    pipe = HostStepPipeline(rif, h, ds, n_rays, flags=flags, allreduce=allreduce)
    for _ in range(3):
        pipe.submit(pos_pin, vel_pin)
    pipe.drain()
    barrier()
    e0, e1 = ev(), ev()
    e0.record()
    checksum = 0
    for k in range(args.steps):
        idx = pipe.submit(pos_pin, vel_pin)
        if k > 0:
            g_host, st_host = pipe.result(idx - 1)
            checksum += int(st_host[0])
    g_host, st_host = pipe.result(idx)
    checksum += int(st_host[0])
    e1.record(pipe.s_out)
    barrier()
    e2e_ms = e0.elapsed_time(e1)
    assert checksum == args.steps * (s_f + s_b), (checksum, s_f, s_b)
    h2d = pos_pin.numel() * 4 + vel_pin.numel() * 4
    d2h = grad_pin.numel() * 4 + stat_pin.numel() * 8

    # ---- reduce over ranks: max time, sum of work ----------------------------------
    red = torch.tensor([ms, e2e_ms, float(steps_total), fwd_ms, adj_ms], dtype=torch.float64, device=dev)
    if world > 1:
        mx = red.clone(); dist.all_reduce(mx, op=dist.ReduceOp.MAX)
        sm = red.clone(); dist.all_reduce(sm, op=dist.ReduceOp.SUM)
        ms, e2e_ms, fwd_ms, adj_ms = float(mx[0]), float(mx[1]), float(mx[3]), float(mx[4])
        steps_all = float(sm[2])
    else:
        steps_all = float(steps_total)
    value = steps_all / (ms * 1e-3)
    e2e_value = steps_all / (e2e_ms * 1e-3)

    if rank == 0:
        peak, peak_src = peaks()
        adj_gbs = BYTES_ADJ * s_b / (adj_ms * 1e-3) / 1e9
        fwd_gbs = BYTES_FWD * s_f / (fwd_ms * 1e-3) / 1e9
        traffic = None
        tp = os.path.join(ROOT, "profiles", "traffic.json")
        if os.path.exists(tp):
            try:
                traffic = json.load(open(tp)).get("backtrace_kernel_dram_bytes_per_launch")
            except Exception:
                traffic = None
        # the adjoint's atomics against the MEASURED L2 sector throughput (tools/l2_probe.cu): the grids are
        # L2-resident at 256^3, so this is the memory roofline the kernel actually lives under
        l2 = None
        lp = os.path.join(ROOT, "profiles", "l2_peaks.json")
        if os.path.exists(lp) and NG == 256:
            try:
                pk = json.load(open(lp))
                red = pk["backtrace_kernel_l1_red_sectors_per_launch"] / (adj_ms * 1e-3) / 1e9
                ld = pk["backtrace_kernel_l1_ld_sectors_per_launch"] / (adj_ms * 1e-3) / 1e9
                l2 = {"red_gsectors_per_s": red, "red_peak_gsectors_per_s": pk["red_scattered_gsectors_per_s"],
                      "red_frac": red / pk["red_scattered_gsectors_per_s"], "gather_gsectors_per_s": ld,
                      "gather_peak_gsectors_per_s": pk["gather_scattered_gsectors_per_s"],
                      "note": "sector counts per launch from the committed ncu capture of this workload, divided by the "
                              "live kernel time; peaks from profiles/r05_l2_probe.log", "source": pk["source"]}
            except Exception:
                l2 = None
        # what actually bounds the kernels at 256^3: instruction issue / the FP32 pipes.
        # warp instructions per launch from the committed ncu capture of this workload over the live kernel
        # times, against 4 schedulers x 1 instruction per clock per SM at the SM clock sampled during the run.
        issue = None
        ip = os.path.join(ROOT, "profiles", "issue_counts.json")
        if os.path.exists(ip) and NG == 256 and clocks and clocks.get("sm_mhz"):
            try:
                ic = json.load(open(ip))
                pk_inst = 148 * 4 * clocks["sm_mhz"] * 1e6 / 1e9
                fi = ic["trace_kernel_warp_inst_per_launch"] / (fwd_ms * 1e-3) / 1e9
                ai = ic["backtrace_kernel_warp_inst_per_launch"] / (adj_ms * 1e-3) / 1e9
                issue = {"unit": "G warp-instructions/s", "peak": pk_inst,
                         "fwd": {"achieved": fi, "frac": fi / pk_inst, "pipes_pct": ic["trace_kernel_pipes_pct"]},
                         "adj": {"achieved": ai, "frac": ai / pk_inst, "pipes_pct": ic["backtrace_kernel_pipes_pct"]},
                         "source": ic["source"]}
            except Exception:
                issue = None
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": max(args.warmup, 3), "ms_per_step": ms / args.steps, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {
                "workload": WORKLOAD,
                "rays_per_gpu": n_rays, "grid": [NG, NG, NG], "parallelism": f"rays-dp{world} + grad allreduce" if world > 1 else "single GPU",
                "l2": (f"no explicit flush: per-step working set (rays in/out {n_rays * 72 / 1e6:.0f} MB + grid {NG ** 3 * 4 / 1e6:.0f} MB + "
                       f"gradient {NG ** 3 * 4 / 1e6:.0f} MB) exceeds the 126 MB L2"
                       + ("; the 67 MB IOR grid staying L2-resident across the step is the design" if NG == 256 else "")),
                "arithmetic": "exact trajectory (default flags=0): ray path bit-identical to the float32 reference formulas", "ray_steps_fwd": s_f, "ray_steps_adj": s_b,
                "rays_not_exited": not_exited,
            },
            "fwd_ms": fwd_ms, "adj_ms": adj_ms,
            "fwd_ray_steps_per_s": s_f / (fwd_ms * 1e-3), "adj_ray_steps_per_s": s_b / (adj_ms * 1e-3),
            "roofline": {
                "bound": "hbm", "kernel": "backtrace_kernel<sdf=0,A_EXACT,AccumFloat> (face-carry flush, no interleave)", "achieved": adj_gbs, "peak": peak,
                "unit": "GB/s", "frac": adj_gbs / peak, "traffic": traffic, "peak_source": peak_src,
                "algorithmic_bytes_per_ray_step": BYTES_ADJ, "l2": l2, "issue": issue,
                "fwd_kernel": {"achieved": fwd_gbs, "frac": fwd_gbs / peak, "algorithmic_bytes_per_ray_step": BYTES_FWD},
                "step": {"achieved": (BYTES_FWD * s_f + BYTES_ADJ * s_b) / ((fwd_ms + adj_ms) * 1e-3) / 1e9,
                         "frac": (BYTES_FWD * s_f + BYTES_ADJ * s_b) / ((fwd_ms + adj_ms) * 1e-3) / 1e9 / peak},
            },
            "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                    "ms_per_step": e2e_ms / args.steps,
                    "path": "pipeline.HostStepPipeline: pinned host rays -> H2D -> ops.trace -> ops.backtrace -> allreduce -> D2H gradient grid + counters into pinned memory, read on the host every step; depth-2 double buffering on copy-in / compute / copy-out streams",
                    "serial_ms_per_step": e2e_serial_ms / args.steps},
            "gpu_launches": launches, "clocks": clocks,
            "secondary": {"fast_arith_ms_per_step": fast_ms, "fast_arith_ray_steps_per_s": (s_f + s_b) / (fast_ms * 1e-3),
                          "note": "DRRT_FAST_ARITH (factored polynomial cell): ray paths differ from the reference at cell faces; not used for value/e2e"},
        }
        if world == 1 and not args.no_cpu:
            arm = CpuArm(target_s=8.0)
            sf, sb = arm.count_steps()
            t = arm.step()
            line["cpu_baseline"] = {"value": (sf + sb) / t, "unit": UNIT, "cores": arm.threads, "kind": arm.kind,
                                    "sample": arm.describe(), "seconds": t}
        _emit(line)
    if allreduce is not None:
        allreduce.close()
    if world > 1:
        dist.destroy_process_group()


_REAL_STDOUT = None


def _guard_stdout():
    """Everything libraries print to fd 1 (e.g. NCCL's version banner) goes to stderr; the one JSON
    line is written to the original stdout by _emit()."""
    global _REAL_STDOUT
    sys.stdout.flush()
    _REAL_STDOUT = os.dup(1)
    os.dup2(2, 1)


def _emit(line: dict):
    data = (json.dumps(line) + "\n").encode()
    if _REAL_STDOUT is None:
        sys.stdout.write(data.decode()); sys.stdout.flush()
    else:
        os.write(_REAL_STDOUT, data)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    ap.add_argument("--workload", default="caustic256", choices=["caustic256", "focal512"],
                    help="caustic256 = BASELINE.json's metric configuration (default, the bench line); focal512 = "
                         "configs[4]: 512^3 grid, 1448^2 = 2.1 M rays per GPU (16.8 M over 8 GPUs), a labelled extra run")
    args = ap.parse_args()
    if args.workload == "focal512":
        global NG, SENSOR, METRIC, WORKLOAD
        NG, SENSOR = 512, 1448
        METRIC = "fwd+adjoint ray-steps/sec at 512^3 grid"
        WORKLOAD = (f"focal-512 (BASELINE.json configs[4]): {NG}^3 bump IOR grid, {SENSOR}^2 = {SENSOR * SENSOR} "
                    "collimated rays per GPU, ds = h/2, seeds dx=dv=1")
    _guard_stdout()
    if args.impl == "reference":
        run_reference_arm(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
