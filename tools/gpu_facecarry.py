"""Dev script (GPU box): A/B of the adjoint's face-carry flush (DRRT_NO_FACE_CARRY) on the
collimated, cone and multi-view workloads; checks the gradients agree.

    python tools/gpu_facecarry.py
"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch

import __graft_entry__ as ge

ge.build()
from adjointnonlinearraytracing_b200 import ops, scenes as sc  # noqa: E402
from tools.gpu_configs import multiview_rays, timeit  # noqa: E402

dev = torch.device("cuda:0")
T = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(dev)
NOFC, NOIL, DET = 1 << 7, 1 << 5, 2


def case(name, ng, pos, vel):
    rif = T(sc.field_bump(ng)); h = float(sc.grid_spacing(ng)); ds = h / 2
    pos, vel = T(pos), T(vel)
    one = torch.ones_like(pos)
    xt, vt = ops.trace(rif, pos, vel, h, ds)
    st = ops.new_stats(dev); g0 = ops.backtrace(rif, xt, vt, one, one, h, ds, flags=NOFC, stats=st); sb = int(st[0])
    for nm, fl in (("full flush", NOFC), ("face carry", 0), ("full, no interleave", NOFC | NOIL),
                   ("carry, no interleave", NOIL), ("carry, deterministic", DET)):
        g = ops.backtrace(rif, xt, vt, one, one, h, ds, flags=fl)
        tb = timeit(lambda: ops.backtrace(rif, xt, vt, one, one, h, ds, flags=fl))
        err = float((g.double() - g0.double()).norm() / g0.double().norm())
        print(f"{name:30s} {nm:22s} adj {tb:8.3f} ms {sb/tb/1e6:6.1f} G/s  rel L2 vs full flush {err:.2e}", flush=True)


if __name__ == "__main__":
    case("4 caustic 256^3 collimated", 256, *sc.rays_collimated(2048))
    case("4' caustic 256^3 cone .05", 256, *sc.rays_cone(2048, spread=0.05))
    case("3 fuel 65^3 32x64^2x16", 65, *multiview_rays(32, 64, 16))
    case("1 Luneburg 32^3 256^2", 32, *sc.rays_collimated(256))
    case("5 focal 512^3 1448^2", 512, *sc.rays_collimated(1448))
