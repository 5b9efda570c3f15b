"""Dev script (GPU box): forward trace with one ray per thread against two rays per thread
(DRRT_PAIR_RAYS) on the bench workload and its neighbours; checks bit-identity on each.

    python tools/gpu_pair.py
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
PAIR, NOPAIR = 1 << 25, 1 << 26


def timeit(fn, iters=7):
    fn(); fn(); torch.cuda.synchronize()
    ts = []
    for _ in range(iters):
        a = torch.cuda.Event(enable_timing=True); b = torch.cuda.Event(enable_timing=True)
        a.record(); fn(); b.record(); torch.cuda.synchronize(); ts.append(a.elapsed_time(b))
    return float(np.median(ts))


def main():
    print("lib:", os.environ.get("DRRT_B200_LIB", "default"), flush=True)
    cases = [(256, 2048, "collimated", 0.0), (256, 2048, "cone", 0.05), (256, 2048, "cone", 0.15),
             (512, 2048, "collimated", 0.0), (65, 1024, "cone", 0.3), (32, 256, "collimated", 0.0)]
    for ng, s, kind, spread in cases:
        rif = T(sc.field_bump(ng)); h = float(sc.grid_spacing(ng)); ds = h / 2
        pos, vel = sc.rays_collimated(s) if kind == "collimated" else sc.rays_cone(s, spread=spread)
        pos, vel = T(pos), T(vel)
        st = ops.new_stats(dev)
        xa, va = ops.trace(rif, pos, vel, h, ds, flags=NOPAIR, stats=st); steps = int(st[0])
        xb, vb = ops.trace(rif, pos, vel, h, ds, flags=PAIR)
        same = bool(torch.equal(xa.view(torch.int32), xb.view(torch.int32)) and torch.equal(va.view(torch.int32), vb.view(torch.int32)))
        t1 = timeit(lambda: ops.trace(rif, pos, vel, h, ds, flags=NOPAIR))
        t2 = timeit(lambda: ops.trace(rif, pos, vel, h, ds, flags=PAIR))
        print(f"{ng}^3 {s}^2 {kind} {spread}: single {t1:7.3f} ms ({steps/t1/1e6:6.1f} G/s)  pair {t2:7.3f} ms "
              f"({steps/t2/1e6:6.1f} G/s)  x{t1/t2:.3f}  bit-identical {same}", flush=True)


if __name__ == "__main__":
    main()
