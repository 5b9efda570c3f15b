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
def tiled(s, tx, ty, fn=sc.rays_collimated, **kw):
    """The s x s sensor batch re-ordered so that 32 consecutive rays form a tx x ty pixel tile (tx * ty = 32)."""
    pos, vel = fn(s, **kw)
    idx = np.arange(s * s).reshape(s // ty, ty, s // tx, tx).transpose(0, 2, 1, 3).reshape(-1)
    return np.ascontiguousarray(pos[idx]), np.ascontiguousarray(vel[idx])


def by_entry_cell(ng, pos, vel):
    """Rays sorted by the (linear index of the) grid cell through which they enter the volume."""
    h = 1.0 / (ng - 1)
    p = pos.astype(np.float64); v = vel.astype(np.float64)
    safe = np.where(v != 0, v, 1.0)
    t0 = np.where(v != 0, (0 - p) / safe, -np.inf); t1 = np.where(v != 0, (1 - p) / safe, np.inf)
    tn = np.maximum(np.minimum(t0, t1).max(axis=1), 0)
    c = np.clip(np.floor((p + v * (tn[:, None] + 1e-6)) / h).astype(np.int64), 0, ng - 2)
    idx = np.argsort((c[:, 2] * ng + c[:, 1]) * ng + c[:, 0], kind="stable")
    return ng, np.ascontiguousarray(pos[idx]), np.ascontiguousarray(vel[idx])


CASES = {
    "fuel_sorted": lambda: by_entry_cell(65, *multiview_rays(32, 64, 16)),
    "cone15": lambda: (256, *sc.rays_cone(2048, spread=0.15)),
    "cone15_sorted": lambda: by_entry_cell(256, *sc.rays_cone(2048, spread=0.15)),
    "lunefield_sorted": lambda: by_entry_cell(32, *sc.rays_collimated(256)),
    "caustic_t8x4": lambda: (256, *tiled(2048, 8, 4)),
    "caustic_t4x8": lambda: (256, *tiled(2048, 4, 8)),
    "caustic_t16x2": lambda: (256, *tiled(2048, 16, 2)),
    "luneburg_t8x4": lambda: (32, *tiled(256, 8, 4)),
    "lunefield_t8x4": lambda: (32, *tiled(256, 8, 4)),
    "caustic": lambda: (256, *sc.rays_collimated(2048)),
    "cone": lambda: (256, *sc.rays_cone(2048, spread=0.05)),
    "fuel": lambda: (65, *multiview_rays(32, 64, 16)),
    "focal": lambda: (512, *sc.rays_collimated(1448)),
    "luneburg": lambda: (32, *sc.rays_collimated(256)),
    "lunyaml": lambda: (129, *lunyaml()),
    "lunefield": lambda: (32, *sc.rays_collimated(256)),
}
FIELD = {"lunefield": sc.field_luneburg, "lunefield_sorted": sc.field_luneburg, "lunefield_t8x4": sc.field_luneburg, "lunyaml": sc.field_luneburg}
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
    for _ in range(2):
        a = torch.cuda.Event(enable_timing=True); b = torch.cuda.Event(enable_timing=True)
        a.record(); ops.trace(rif, pos, vel, h, ds); b.record(); torch.cuda.synchronize()
    print(f"{name} fwd (warm) {a.elapsed_time(b):.3f} ms", flush=True)
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
