"""Dev script (GPU box): adjoint time on the bench workload and its neighbours, for A/B builds of the
library (DRRT_B200_LIB=...).  Prints ms and the relative L2 distance to the deterministic mode.

    python tools/gpu_adj_ab.py
"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch

import __graft_entry__ as ge

ge.build()
from adjointnonlinearraytracing_b200 import ops, scenes as sc  # noqa: E402

dev = torch.device("cuda:0")
T = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(dev)


def timeit(fn, iters=7):
    fn(); fn(); torch.cuda.synchronize()
    ts = []
    for _ in range(iters):
        a = torch.cuda.Event(enable_timing=True); b = torch.cuda.Event(enable_timing=True)
        a.record(); fn(); b.record(); torch.cuda.synchronize(); ts.append(a.elapsed_time(b))
    return float(np.median(ts))


def main():
    print("lib:", os.environ.get("DRRT_B200_LIB", "default"), flush=True)
    cases = [(256, 2048, "collimated", 0.0, 2), (256, 2048, "cone", 0.05, 2), (256, 2048, "cone", 0.15, 2),
             (512, 2048, "collimated", 0.0, 2), (65, 1024, "cone", 0.3, 2), (129, 512, "cone", 0.1, 32),
             (32, 256, "collimated", 0.0, 2)]
    for ng, s, kind, spread, sub in cases:
        rif = T(sc.field_bump(ng)); h = float(sc.grid_spacing(ng)); ds = h / sub
        pos, vel = sc.rays_collimated(s) if kind == "collimated" else sc.rays_cone(s, spread=spread)
        pos, vel = T(pos), T(vel)
        xt, vt = ops.trace(rif, pos, vel, h, ds)
        one = torch.ones_like(pos)
        st = ops.new_stats(dev)
        g = ops.backtrace(rif, xt, vt, one, one, h, ds, stats=st); steps = int(st[0])
        gd = ops.backtrace(rif, xt, vt, one, one, h, ds, flags=2)
        err = float((g.double() - gd.double()).norm() / gd.double().norm())
        t = timeit(lambda: ops.backtrace(rif, xt, vt, one, one, h, ds))
        print(f"{ng}^3 {s}^2 {kind} {spread} ds=h/{sub}: adjoint {t:7.3f} ms ({steps/t/1e6:6.1f} G/s)  rel L2 vs deterministic {err:.2e}", flush=True)


if __name__ == "__main__":
    main()
