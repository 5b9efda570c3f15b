"""Dev script (GPU box): time forward/adjoint in every arithmetic / aggregation mode on the
256^3 workload and check the default mode against the literal reference arithmetic.

    python tools/gpu_perf.py [sensor=2048] [rays=collimated,cone]
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
REF, DET, FAST, AGG, NOIL = 1, 2, 8, 16, 32


def timeit(fn, iters=5):
    fn(); torch.cuda.synchronize()
    ts = []
    for _ in range(iters):
        a = torch.cuda.Event(enable_timing=True); b = torch.cuda.Event(enable_timing=True)
        a.record(); fn(); b.record(); torch.cuda.synchronize(); ts.append(a.elapsed_time(b))
    return float(np.median(ts))


def rel_l2(a, b):
    return float((a.double() - b.double()).norm() / b.double().norm())


def main():
    S = int(sys.argv[1]) if len(sys.argv) > 1 else 2048
    kinds = (sys.argv[2] if len(sys.argv) > 2 else "collimated,cone").split(",")
    Ng = int(sys.argv[3]) if len(sys.argv) > 3 else 256
    rif = T(sc.field_bump(Ng)); h = float(sc.grid_spacing(Ng)); ds = h / 2
    for kind in kinds:
        pos, vel = (sc.rays_collimated(S) if kind == "collimated" else sc.rays_cone(S))
        pos, vel = T(pos), T(vel)
        dx = torch.ones_like(pos); dv = torch.ones_like(pos)
        xr, vr = ops.trace(rif, pos, vel, h, ds, flags=REF)
        gref = ops.backtrace(rif, xr, vr, dx, dv, h, ds, flags=REF)
        for nm, fl in [("exact", 0), ("exact-noil", NOIL), ("exact-agg-noil", AGG | NOIL), ("fast", FAST),
                       ("ref", REF), ("exact-det", DET), ("exact-det-noil", DET | NOIL)]:
            st = ops.new_stats(dev)
            xt, vt = ops.trace(rif, pos, vel, h, ds, flags=fl, stats=st); sf = int(st[0])
            tf = timeit(lambda: ops.trace(rif, pos, vel, h, ds, flags=fl))
            st = ops.new_stats(dev)
            g = ops.backtrace(rif, xt, vt, dx, dv, h, ds, flags=fl, stats=st); sb = int(st[0])
            tb = timeit(lambda: ops.backtrace(rif, xt, vt, dx, dv, h, ds, flags=fl))
            same = bool(torch.equal(xt, xr) and torch.equal(vt, vr))
            print(f"{Ng}^3 {S}^2 {kind:10s} {nm:15s} fwd {sf/tf/1e6:6.1f} G/s ({tf:6.2f} ms) adj {sb/tb/1e6:6.1f} G/s ({tb:6.2f} ms) "
                  f"both {(sf+sb)/(tf+tb)/1e6:6.1f} G/s alg {(32*sf+64*sb)/(tf+tb)/1e6:5.0f} GB/s | endpoints==ref {same} "
                  f"grad relL2 vs ref {rel_l2(g, gref):.2e}", flush=True)


if __name__ == "__main__":
    main()
