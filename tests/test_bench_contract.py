"""CPU checks of bench.py's contract pieces that do not need a GPU: the reference arm's JSON line (the CPU leg the driver
runs beside ours), the source hash that ties committed ncu counts to the built kernels, and the committed counts file."""
import json
import os
import subprocess
import sys

from conftest import ROOT


def test_reference_arm_prints_one_contract_line():
    out = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--steps", "1", "--warmup", "0"],
                         capture_output=True, text=True, timeout=600)
    assert out.returncode == 0, out.stderr[-2000:]
    lines = [l for l in out.stdout.splitlines() if l.strip()]
    assert len(lines) == 1
    d = json.loads(lines[0])
    assert d["impl"] == "reference" and d["unit"] == "ray-steps/s" and d["higher_is_better"] is True
    assert d["value"] > 0 and d["cpu_baseline"]["value"] == d["value"] and d["cpu_baseline"]["cores"] >= 1
    assert d["cpu_baseline"]["kind"] in ("reference", "port")
    assert d["e2e"] == {"value": d["value"], "unit": d["unit"], "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}
    assert d["scaling"] == "strong" and "2048^2 = 4194304 collimated rays in total" in d["config"]["workload"]


def test_csrc_hash_and_committed_counts():
    sys.path.insert(0, ROOT)
    import bench
    a, b = bench.csrc_sha(), bench.csrc_sha()
    assert a == b and len(a) == 16 and int(a, 16) >= 0
    p = os.path.join(ROOT, "profiles", "kernel_counts.json")
    if os.path.exists(p):
        c = json.load(open(p))
        for k in ("backtrace_kernel", "trace_kernel"):
            assert c[k]["warp_inst"] > 1e9 and 0 < c[k]["threads_per_inst"] <= 32
        assert len(c["csrc_sha"]) == 16
