"""Dev script (GPU box, under torchrun): one optimiser ITERATION at N ranks -- forward trace, adjoint, gradient sum,
boundary-zero + Adam + clamp -- with the three ways of doing the last two steps:
  (a) NCCL all-reduce, whole-grid Adam on every rank        (dist.GradAllReduce + ops.adam_step_)
  (b) NCCL reduce-scatter, slab Adam, all-gather            (dist.ShardedAdam)
  (c) one kernel over peer memory                           (dist.PeerShardedAdam)

    python -m torch.distributed.run --nproc-per-node N tools/gpu_iter_multi.py [res=512] [sensor=1448]
"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
import torch.distributed as dist

import __graft_entry__ as ge

ge.build()
from adjointnonlinearraytracing_b200 import ops, scenes as sc  # noqa: E402
from adjointnonlinearraytracing_b200.dist import GradAllReduce, PeerShardedAdam, ShardedAdam, ShardedTracer  # noqa: E402

rank = int(os.environ.get("LOCAL_RANK", 0)); world = int(os.environ.get("WORLD_SIZE", 1))
torch.cuda.set_device(rank)
if world > 1:
    dist.init_process_group("nccl", device_id=torch.device("cuda", rank))
res = int(sys.argv[1]) if len(sys.argv) > 1 else 512
S = int(sys.argv[2]) if len(sys.argv) > 2 else 1448
T = lambda a: torch.from_numpy(np.ascontiguousarray(a)).cuda()
rif0 = T(sc.field_bump(res)); h = float(sc.grid_spacing(res)); ds = h / 2
pos, vel = sc.rays_collimated(S)   # every rank traces its own S^2 rays (weak scaling, like bench.py)
pos, vel = T(pos), T(vel)
dx = torch.ones_like(pos); dv = torch.ones_like(pos)
ar = GradAllReduce()
# lr = 0 everywhere: the optimiser pass does all its work but the grid (and with it the ray paths and the step
# counts) stays the same from iteration to iteration, so the variants are timed on identical tracing work


def timeit(fn, iters=8):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    a = torch.cuda.Event(enable_timing=True); b = torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(iters):
        fn()
    b.record(); torch.cuda.synchronize()
    t = torch.tensor([a.elapsed_time(b) / iters], device="cuda")
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t)


out = {}
# (a)
n = rif0.clone(); st = ShardedTracer(n, h, ds, allreduce=ar)
m = torch.zeros_like(n); v = torch.zeros_like(n); k = [0]
def it_a():
    xt, vt = st.trace(pos, vel); g = st.backtrace(xt, vt, dx, dv); k[0] += 1
    ops.adam_step_(n, g, m, v, k[0], 0.0)
out["a"] = timeit(it_a)
del m, v, st, n
# (b)
n = rif0.clone(); st = ShardedTracer(n, h, ds, allreduce=ar, reduce=False); sopt = ShardedAdam(n, 0.0, allreduce=ar)
def it_b():
    xt, vt = st.trace(pos, vel); sopt.step(st.backtrace(xt, vt, dx, dv))
out["b"] = timeit(it_b)
del sopt, st, n
# (c)
n = rif0.clone(); st = ShardedTracer(n, h, ds, allreduce=ar, reduce=False); popt = PeerShardedAdam(n, st.grad, 0.0, allreduce=ar)
def it_c():
    xt, vt = st.trace(pos, vel); st.backtrace(xt, vt, dx, dv); popt.step()
out["c"] = timeit(it_c)
# tracer alone
def it_t():
    xt, vt = st.trace(pos, vel); st.backtrace(xt, vt, dx, dv)
out["t"] = timeit(it_t)
if rank == 0:
    print(f"{res}^3, {S}^2 rays per rank, {world} ranks: trace + adjoint alone {out['t']:.2f} ms; whole iteration with "
          f"all-reduce + whole-grid Adam {out['a']:.2f} ms, reduce-scatter + slab Adam + all-gather {out['b']:.2f} ms, "
          f"peer-memory kernel {out['c']:.2f} ms", flush=True)
popt.close(); ar.close()
if world > 1:
    dist.destroy_process_group()
