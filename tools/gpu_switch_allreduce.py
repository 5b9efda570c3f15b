"""Dev script (torchrun, N GPUs of one node): the in-switch gradient sum (dist.SwitchAllReduce, our multimem kernel)
against NCCL's all-reduce (dist.GradAllReduce): same result, time per call on the 256^3 and 512^3 gradient grids.

    python -m torch.distributed.run --nproc-per-node N --master-addr 127.0.0.1 tools/gpu_switch_allreduce.py
"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist

import __graft_entry__ as ge

ge.build()
from adjointnonlinearraytracing_b200.dist import GradAllReduce, SwitchAllReduce  # noqa: E402

rank = int(os.environ["RANK"]); local = int(os.environ["LOCAL_RANK"]); world = int(os.environ["WORLD_SIZE"])
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
dist.init_process_group("nccl", device_id=dev)
nccl = GradAllReduce()
try:
    sw = SwitchAllReduce()
except Exception as e:  # noqa: BLE001
    if rank == 0:
        print("SwitchAllReduce unavailable:", repr(e))
    sys.exit(0)


def timed(fn, iters=20):
    for _ in range(3):
        fn()
    dist.barrier(); torch.cuda.synchronize()
    a = torch.cuda.Event(enable_timing=True); b = torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(iters):
        fn()
    b.record(); torch.cuda.synchronize()
    t = torch.tensor([a.elapsed_time(b) / iters], device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t)


for ng in ((256,) if os.environ.get('DRRT_MM_UNROLL') else (256, 512)):
    shape = (ng, ng, ng)
    try:
        buf = sw.alloc(shape, dev)
    except Exception as e:  # noqa: BLE001
        if rank == 0:
            print(f"{ng}^3: alloc failed: {e!r}")
        break
    g = torch.Generator(device=dev).manual_seed(100 + rank)
    src = torch.randn(shape, device=dev, generator=g)
    ref = src.clone(); nccl(ref)
    buf.copy_(src); sw(buf)
    torch.cuda.synchronize()
    err = float((buf.double() - ref.double()).norm() / ref.double().norm())
    same_everywhere = buf.clone(); dist.broadcast(same_everywhere, 0)
    ident = bool(torch.equal(same_everywhere, buf))
    t_n = timed(lambda: nccl(ref))
    t_s = timed(lambda: sw(buf))
    if rank == 0:
        gb = ng ** 3 * 4 / 1e9
        print(f"[unroll {os.environ.get('DRRT_MM_UNROLL', 'default')} ctas/SM {os.environ.get('DRRT_MM_CTAS', 'default')}] {ng}^3 ({gb * 1e3:.0f} MB) on {world} GPUs: NCCL all-reduce {t_n:.3f} ms | in-switch multimem kernel {t_s:.3f} ms "
              f"(incl. 2 barriers) | rel L2 vs NCCL {err:.1e} | identical on all ranks: {ident} | "
              f"algorithmic {gb / (t_s * 1e-3):.0f} GB/s", flush=True)
    del buf, src, ref, same_everywhere
sw.close(); nccl.close()
dist.destroy_process_group()
