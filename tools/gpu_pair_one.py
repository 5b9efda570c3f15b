"""Dev script (GPU box): one forward trace per mode on the bench workload, for ncu (pipe metrics)."""
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
ng, s = 256, 2048
rif = T(sc.field_bump(ng)); h = float(sc.grid_spacing(ng)); ds = h / 2
pos, vel = sc.rays_collimated(s)
pos, vel = T(pos), T(vel)
for fl in (1 << 26, 1 << 25, 8):  # single, pair, fast arithmetic
    xt, vt = ops.trace(rif, pos, vel, h, ds, flags=fl)
    g = ops.backtrace(rif, xt, vt, torch.ones_like(pos), torch.ones_like(pos), h, ds, flags=fl & 8)
torch.cuda.synchronize()
