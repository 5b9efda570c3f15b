"""Dev script (GPU box): adjoint time vs warp-aggregation policy for several rays-per-cell ratios."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import __graft_entry__ as ge
ge.build()
from adjointnonlinearraytracing_b200 import ops, scenes as sc
from gpu_perf import timeit, T, dev

AGG, NOIL = 16, 32
for Ng, S, kind in [(256, 2048, "collimated"), (256, 1024, "collimated"), (256, 4096, "collimated"), (512, 2048, "collimated"),
                    (256, 2048, "cone05"), (64, 2048, "collimated")]:
    rif = T(sc.field_bump(Ng)); h = float(sc.grid_spacing(Ng)); ds = h / 2
    pos, vel = sc.rays_collimated(S) if kind == "collimated" else sc.rays_cone(S, spread=0.05)
    pos, vel = T(pos), T(vel)
    one = torch.ones_like(pos)
    xt, vt = ops.trace(rif, pos, vel, h, ds)
    st = ops.new_stats(dev); ops.backtrace(rif, xt, vt, one, one, h, ds, stats=st); sb = int(st[0])
    row = []
    for nm, fl in [("plain", NOIL), ("agg-adaptive", AGG | NOIL), ("il-auto", 0), ("il4", 4 << 12), ("il8", 8 << 12), ("il16", 16 << 12), ("il32", 32 << 12), ("il-auto+agg", AGG)]:
        tb = timeit(lambda: ops.backtrace(rif, xt, vt, one, one, h, ds, flags=fl))
        row.append(f"{nm} {tb:6.2f} ms ({sb/tb/1e6:5.1f} G/s)")
    print(f"{Ng}^3 {S}^2 {kind:10s} rays/cell {S/(Ng-1):4.1f}: " + " | ".join(row), flush=True)
    del rif, pos, vel, one, xt, vt
    torch.cuda.empty_cache()
