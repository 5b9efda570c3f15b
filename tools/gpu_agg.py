"""Dev script (GPU box): adjoint time per aggregation policy on focusing (lens), dense and ordinary batches.

    python tools/gpu_agg.py [small|all|big] [replicas]

(the r09 logs under profiles/ were taken with environment overrides of the thresholds / copy counts that
the library no longer reads)
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
from tools.gpu_one import lunyaml  # noqa: E402

dev = torch.device("cuda:0")
T = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(dev)
AGG = 1 << 4
FLAGS = [("plain", 0), ("agg off", AGG | (1 << 8)), ("agg adaptive", AGG), ("agg run 8", AGG | (8 << 8)), ("agg run warp", AGG | (15 << 8))]
NOREP = 1 << 19
if len(sys.argv) > 2 and sys.argv[2] == "replicas":
    FLAGS = [("one grid", NOREP), ("replicas", 0), ("replicas + agg warp", AGG | (15 << 8))]
elif len(sys.argv) > 2 and sys.argv[2] == "forced":
    FLAGS = [("one grid", NOREP), ("default", 0)] + [(f"{n} copies", n << 20) for n in (2, 4, 8, 15)]
else:
    FLAGS = [(n, f | NOREP) for n, f in FLAGS]
CASES = [("lens 32^3, 256^2", 32, 2, sc.field_luneburg, lambda: sc.rays_collimated(256)),
         ("lens 32^3, 362^2", 32, 2, sc.field_luneburg, lambda: sc.rays_collimated(362)),
         ("bump 32^3, 256^2", 32, 2, sc.field_bump, lambda: sc.rays_collimated(256)),
         ("bump 32^3, 1024^2 (dense)", 32, 2, sc.field_bump, lambda: sc.rays_collimated(1024)),
         ("lens 64^3, 512^2", 64, 2, sc.field_luneburg, lambda: sc.rays_collimated(512)),
         ("bump 64^3, 512^2", 64, 2, sc.field_bump, lambda: sc.rays_collimated(512)),
         ("fuel 65^3, 32 views", 65, 2, sc.field_bump, lambda: multiview_rays(32, 64, 16))]
if len(sys.argv) > 1 and sys.argv[1] == "sparse":  # few rays per voxel: is the 8 N / G bound on the copies right?
    CASES = [("lens 64^3, 256^2", 64, 2, sc.field_luneburg, lambda: sc.rays_collimated(256)),
             ("bump 64^3, 256^2", 64, 2, sc.field_bump, lambda: sc.rays_collimated(256)),
             ("one view 65^3, 64^2 x 16", 65, 2, sc.field_bump, lambda: multiview_rays(1, 64, 16)),
             ("lens 128^3, 512^2", 128, 2, sc.field_luneburg, lambda: sc.rays_collimated(512)),
             ("bump 128^3, 512^2", 128, 2, sc.field_bump, lambda: sc.rays_collimated(512))]
if len(sys.argv) > 1 and sys.argv[1] == "big":
    CASES = []
if len(sys.argv) > 1 and sys.argv[1] in ("all", "big"):
    CASES += [("lens 129^3 yaml, ds=h/32", 129, 32, sc.field_luneburg, lunyaml),
              ("caustic 256^3, 2048^2", 256, 2, sc.field_bump, lambda: sc.rays_collimated(2048)),
              ("lens 256^3, 2048^2", 256, 2, sc.field_luneburg, lambda: sc.rays_collimated(2048)),
              ("cone .05 256^3, 2048^2", 256, 2, sc.field_bump, lambda: sc.rays_cone(2048, spread=0.05))]


def timed(fn, k):
    for _ in range(3):
        fn()
    ts = []
    for _ in range(k):
        a = torch.cuda.Event(enable_timing=True); b = torch.cuda.Event(enable_timing=True)
        a.record(); fn(); b.record(); torch.cuda.synchronize()
        ts.append(a.elapsed_time(b))
    return float(np.median(ts))


for name, ng, sr, field, rays in CASES:
    rif = T(field(ng)); h = float(sc.grid_spacing(ng)); ds = h / sr
    pos, vel = rays()
    pos, vel = T(pos), T(vel)
    one = torch.ones_like(pos)
    st = ops.new_stats(dev)
    xt, vt = ops.trace(rif, pos, vel, h, ds)
    grad = torch.empty_like(rif)
    st.zero_()
    ops.backtrace(rif, xt, vt, one, one, h, ds, out=grad, flags=2, stats=st)  # integer accumulation: the yardstick
    g0 = grad.double().clone(); steps = int(st[0])
    line = f"{name:28s} {pos.shape[0]:8d} rays |"
    for label, fl in FLAGS:
        st.zero_()
        ops.backtrace(rif, xt, vt, one, one, h, ds, out=grad, flags=fl, stats=st)
        err = float((grad.double() - g0).norm() / g0.norm())
        assert int(st[0]) == steps, (label, int(st[0]), steps)
        t = timed(lambda: ops.backtrace(rif, xt, vt, one, one, h, ds, out=grad, flags=fl), 20 if ng <= 128 else 5)
        line += f" {label} {t*1e3:9.1f} us ({steps/t*1e-6:6.1f} G/s, {err:.0e}) |"
    print(line, flush=True)
