"""Dev script (GPU box): the small configurations are bound by the host's launch path, not by the
kernels.  Times one forward + adjoint step three ways: eager ops calls (events around each step),
eager back to back (one event pair around K steps), and a CUDA-graph replay (pipeline.GraphedStep).

    python tools/gpu_small.py [K = 200]
"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch

import __graft_entry__ as ge

ge.build()
from adjointnonlinearraytracing_b200 import ops, scenes as sc  # noqa: E402
from adjointnonlinearraytracing_b200.pipeline import GraphedStep  # noqa: E402

dev = torch.device("cuda:0")
T = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(dev)
K = int(sys.argv[1]) if len(sys.argv) > 1 else 200
CASES = [("Luneburg 32^3, 256^2 rays", 32, 256, sc.field_luneburg), ("fibre (grid form) 64^3, 512^2", 64, 512, sc.field_bump),
         ("fuel-size 65^3, 512^2", 65, 512, sc.field_bump), ("128^3, 1024^2", 128, 1024, sc.field_bump)]


def timed(fn, k):
    a = torch.cuda.Event(enable_timing=True); b = torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize(); a.record()
    for _ in range(k):
        fn()
    b.record(); torch.cuda.synchronize()
    return a.elapsed_time(b) / k


for name, ng, s, field in CASES:
    rif = T(field(ng)); h = float(sc.grid_spacing(ng)); ds = h / 2
    pos, vel = sc.rays_collimated(s)
    pos, vel = T(pos), T(vel)
    one = torch.ones_like(pos)
    st = ops.new_stats(dev)
    xt, vt = ops.trace(rif, pos, vel, h, ds, stats=st)
    ops.backtrace(rif, xt, vt, one, one, h, ds, stats=st)
    steps = int(st[0])
    grad = torch.empty_like(rif)

    def eager():
        x, v = ops.trace(rif, pos, vel, h, ds)
        ops.backtrace(rif, x, v, one, one, h, ds, out=grad)

    for _ in range(10):
        eager()
    per_step = np.median([timed(eager, 1) for _ in range(50)])
    back_to_back = timed(eager, K)
    g = GraphedStep(rif, h, ds, s * s)
    g.pos.copy_(pos); g.vel.copy_(vel)
    for _ in range(10):
        g.step()
    graphed = timed(g.step, K)
    err = float((g.grad - grad).norm() / grad.norm())
    print(f"{name:32s} {steps/1e6:8.2f} M ray-steps | eager (events per step) {per_step*1e3:8.1f} us | eager x{K} {back_to_back*1e3:8.1f} us"
          f" | graph replay {graphed*1e3:8.1f} us = {steps/graphed*1e-6:7.1f} G ray-steps/s | rel L2 {err:.1e}", flush=True)
