"""Dev script (GPU box): one forward + N adjoint launches of a named workload with given flags --
the thing to put under ncu.

    python tools/gpu_one.py <caustic|cone|fuel|focal|luneburg|lunyaml> <flags,flags,...> [reps] [step_res]
"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch

import __graft_entry__ as ge

ge.build()
from adjointnonlinearraytracing_b200 import ops, scenes as sc  # noqa: E402
from tools.gpu_configs import multiview_rays  # noqa: E402

dev = torch.device("cuda:0")
T = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(dev)
CASES = {
    "caustic": lambda: (256, *sc.rays_collimated(2048)),
    "cone": lambda: (256, *sc.rays_cone(2048, spread=0.05)),
    "fuel": lambda: (65, *multiview_rays(32, 64, 16)),
    "focal": lambda: (512, *sc.rays_collimated(1448)),
    "luneburg": lambda: (32, *sc.rays_collimated(256)),
    "lunyaml": lambda: (129, *lunyaml()),
    "lunefield": lambda: (32, *sc.rays_collimated(256)),
}
FIELD = {"lunefield": sc.field_luneburg, "lunyaml": sc.field_luneburg}
STEP_RES = {"lunyaml": 32}


def lunyaml():
    from adjointnonlinearraytracing_b200 import source
    torch.manual_seed(0); np.random.seed(0)
    rot = source.random_rotmat().numpy()
    (x, v, planes), nr = source.rand_rays_cube((128, 128), 10, 1.0, circle=True, rotmat=rot)
    return x.cpu().numpy(), v.cpu().numpy()

if __name__ == "__main__":
    name = sys.argv[1]
    flag_list = [int(f, 0) for f in sys.argv[2].split(",")]
    reps = int(sys.argv[3]) if len(sys.argv) > 3 else 1
    ng, pos, vel = CASES[name]()
    step_res = int(sys.argv[4]) if len(sys.argv) > 4 else STEP_RES.get(name, 2)
    rif = T(FIELD.get(name, sc.field_bump)(ng)); h = float(sc.grid_spacing(ng)); ds = h / step_res
    pos, vel = T(pos), T(vel)
    one = torch.ones_like(pos)
    st = ops.new_stats(dev)
    a = torch.cuda.Event(enable_timing=True); b = torch.cuda.Event(enable_timing=True)
    a.record(); xt, vt = ops.trace(rif, pos, vel, h, ds, stats=st); b.record(); torch.cuda.synchronize()
    print(f"{name} ds = h/{step_res} fwd {a.elapsed_time(b):.3f} ms, {int(st[0])} ray-steps, {pos.shape[0]} rays", flush=True)
    g0 = None
    for fl in flag_list:
        for _ in range(reps):
            a = torch.cuda.Event(enable_timing=True); b = torch.cuda.Event(enable_timing=True)
            a.record(); g = ops.backtrace(rif, xt, vt, one, one, h, ds, flags=fl); b.record()
            torch.cuda.synchronize()
            if g0 is None:
                g0 = g.double()
            err = float((g.double() - g0).norm() / g0.norm())
            print(f"{name} flags={fl:#x} adj {a.elapsed_time(b):.3f} ms  rel L2 vs first {err:.1e}", flush=True)
