"""Dev script (GPU box, under torchrun): the optimiser step after the adjoint at N ranks --
all-reduce + whole-grid Adam on every rank against reduce-scatter + slab Adam + all-gather (dist.ShardedAdam).

    python -m torch.distributed.run --nproc-per-node N tools/gpu_sharded_adam.py [res=512]
"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
import torch.distributed as dist

import __graft_entry__ as ge

ge.build()
from adjointnonlinearraytracing_b200 import ops  # noqa: E402
from adjointnonlinearraytracing_b200.dist import GradAllReduce, PeerShardedAdam, ShardedAdam  # noqa: E402

rank = int(os.environ.get("LOCAL_RANK", 0)); world = int(os.environ.get("WORLD_SIZE", 1))
torch.cuda.set_device(rank)
if world > 1:
    dist.init_process_group("nccl", device_id=torch.device("cuda", rank))
res = int(sys.argv[1]) if len(sys.argv) > 1 else 512
shape = (res, res, res)
gen = torch.Generator(device="cuda").manual_seed(1)
n = 1.0 + 0.1 * torch.rand(shape, device="cuda", generator=gen)
grad = torch.randn(shape, device="cuda", generator=gen)
ar = GradAllReduce()


def timeit(fn, iters=10):
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


m = torch.zeros_like(n); v = torch.zeros_like(n); k = [0]
def full():
    ar(grad); k[0] += 1
    ops.adam_step_(n, grad, m, v, k[0], 1e-3)
t_ar = timeit(lambda: ar(grad))
t_full = timeit(full)
del m, v
n2 = n.clone()
opt = ShardedAdam(n2, 1e-3, allreduce=ar)
t_sh = timeit(lambda: opt.step(grad))
state_mib = 2 * opt.exp_avg.numel() * 4 / 2**20
del opt
n3 = n.clone()
popt = PeerShardedAdam(n3, grad, 1e-3, allreduce=ar)
t_peer = timeit(lambda: popt.step())
t_bar = timeit(lambda: popt._barrier())
if rank == 0:
    print(f"{res}^3, {world} ranks: one kernel over peer memory (two 4-byte barriers + P2P sum + Adam + P2P store) "
          f"{t_peer:.3f} ms; one barrier alone {t_bar:.3f} ms", flush=True)
popt.close()
# the switch does the sum and the broadcast: dist.SwitchShardedAdam (drrt_adam_step_multimem), checked against the peer kernel
if world > 1:
    try:
        from adjointnonlinearraytracing_b200.dist import SwitchAllReduce, SwitchShardedAdam
        sw = SwitchAllReduce()
        n4 = sw.alloc(shape, torch.device("cuda", rank)); g4 = sw.alloc(shape, torch.device("cuda", rank))
        n4.copy_(n); g4.copy_(grad)
        sopt = SwitchShardedAdam(sw, n4, g4, 1e-3)
        n5 = n.clone(); ref = PeerShardedAdam(n5, grad, 1e-3, allreduce=ar)
        sopt.step(); ref.step(); torch.cuda.synchronize()
        err = float((n4 - n5).abs().max())
        t_sw = timeit(lambda: sopt.step())
        t_swar = timeit(lambda: sw(g4))
        if rank == 0:
            print(f"{res}^3, {world} ranks: ONE kernel through the switch (2 barriers + multimem.ld_reduce + Adam + multimem.st) "
                  f"{t_sw:.3f} ms; max |n - n_peer_kernel| after one step {err:.1e}; in-switch all-reduce alone {t_swar:.3f} ms", flush=True)
        ref.close()
    except Exception as e:  # noqa: BLE001
        if rank == 0:
            print("switch path unavailable:", repr(e)[:300], flush=True)
if rank == 0:
    print(f"{res}^3, {world} ranks: all-reduce alone {t_ar:.3f} ms; all-reduce + whole-grid Adam {t_full:.3f} ms; "
          f"reduce-scatter + slab Adam + all-gather {t_sh:.3f} ms; Adam state per rank {2 * n.numel() * 4 / 2**20:.0f} -> "
          f"{state_mib:.0f} MiB", flush=True)
ar.close()
if world > 1:
    dist.destroy_process_group()
