"""Dev script (one GPU): forward + adjoint time of ONE rank's share of the 2048^2 batch at world = 8 / 4 under different
block sizes of dist.shard_cyclic (rows dealt singly, in blocks of 8 rows = one cell's worth, contiguous)."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch

import __graft_entry__ as ge

ge.build()
from adjointnonlinearraytracing_b200 import ops, scenes as sc  # noqa: E402
from adjointnonlinearraytracing_b200.dist import shard_bounds, shard_cyclic  # noqa: E402

dev = torch.device("cuda:0")
NG, S = 256, 2048
rif = torch.from_numpy(sc.field_bump(NG)).to(dev)
h = float(sc.grid_spacing(NG)); ds = h / 2
pos_h, vel_h = sc.rays_collimated(S)


def timed(fn, iters=10):
    fn(); fn(); torch.cuda.synchronize()
    ts = []
    for _ in range(iters):
        a = torch.cuda.Event(enable_timing=True); b = torch.cuda.Event(enable_timing=True)
        a.record(); fn(); b.record(); torch.cuda.synchronize(); ts.append(a.elapsed_time(b))
    return float(np.median(ts))


for world in (8, 4):
    for label, block in (("rows", S), ("8-row blocks", 8 * S), ("32-row blocks", 32 * S), ("contiguous", None)):
        worst = [0.0, 0.0]
        for rank in (0, world // 2):  # an edge rank and a middle one
            if block is None:
                a, b = shard_bounds(S * S, rank, world); sel = np.arange(a, b)
            else:
                sel = np.concatenate([np.arange(a, b) for a, b in shard_cyclic(S * S, rank, world, block)])
            pos = torch.from_numpy(np.ascontiguousarray(pos_h[sel])).to(dev)
            vel = torch.from_numpy(np.ascontiguousarray(vel_h[sel])).to(dev)
            one = torch.ones_like(pos); g = torch.empty_like(rif)
            xt, vt = ops.trace(rif, pos, vel, h, ds)
            tf = timed(lambda: ops.trace(rif, pos, vel, h, ds))
            tb = timed(lambda: ops.backtrace(rif, xt, vt, one, one, h, ds, out=g))
            worst = [max(worst[0], tf), max(worst[1], tb)]
        print(f"world {world} {label:14s}: slowest of (rank 0, rank {world // 2}) fwd {worst[0]:.3f} ms adj {worst[1]:.3f} ms "
              f"sum {worst[0] + worst[1]:.3f} ms", flush=True)
