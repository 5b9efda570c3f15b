"""Dev script (GPU box): wall time of one optimiser iteration of the Luneburg driver's loop
(core/luneburg_opt.py:92-106) assembled from the drop-in pieces, vs the sum of its kernels.

    python tools/gpu_iter.py [res=33] [nbins=128] [spp=1]
"""
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch

import __graft_entry__ as ge

ge.build()
from adjointnonlinearraytracing_b200 import _capi, ops, optimizer, sensor, source, tracer  # noqa: E402

res = int(sys.argv[1]) if len(sys.argv) > 1 else 33
nbins = int(sys.argv[2]) if len(sys.argv) > 2 else 128
spp = int(sys.argv[3]) if len(sys.argv) > 3 else 1
span = 1.0
h = span / (res - 1); ds = h / 2
torch.manual_seed(0); np.random.seed(0)
n = torch.ones(res, res, res, device="cuda").requires_grad_(True)
opt = optimizer.FusedAdamState(n, 1e-3)


def iteration(sync_loss=True):
    rot = source.random_rotmat().numpy()
    (x, v, planes), rpv = source.rand_rays_cube((nbins, nbins), spp, span, circle=True, rotmat=rot)
    xm, vm = tracer.BackTracerC.apply(n, x, v, h, ds)
    sn = planes[:, 1, :]; sp = planes[:, 0, :]
    xmp, _ = sensor.trace_rays_to_plane((xm, vm), (sp, sn))
    loss = torch.sum((xmp - sp) ** 2) / x.shape[0] / span
    n.grad = None
    loss.backward()
    with torch.no_grad():
        opt.step += 1
        ops.adam_step_(n, n.grad.contiguous(), opt.exp_avg, opt.exp_avg_sq, opt.step, opt.lr, opt.betas, opt.eps,
                       clamp_min=1.0, zero_boundary=True, clamp=True)
    return loss.item() if sync_loss else loss, x.shape[0]


for _ in range(5):
    iteration()
torch.cuda.synchronize()
order = (False, True, False) if os.environ.get("ORDER") == "1" else (True, False)
for sync in order:
    l0 = _capi.launch_count()
    t0 = time.perf_counter()
    K = 50
    stamps = []
    for _ in range(K):
        _, nr = iteration(sync)
        stamps.append(time.perf_counter())
    torch.cuda.synchronize()
    d = np.diff(np.array([t0] + stamps)) * 1e3
    print("   per-iteration host ms: first 5", np.round(d[:5], 2), "median", round(float(np.median(d)), 3), "max", round(float(d.max()), 2))
    dt = (time.perf_counter() - t0) / K
    print(f"res {res}^3, {nr} rays: {dt * 1e3:.3f} ms per iteration ({'loss.item() every step' if sync else 'no host sync'}), "
          f"{(_capi.launch_count() - l0) / K:.1f} library launches per iteration", flush=True)
# kernel-only time of the same work
(x, v, planes), rpv = source.rand_rays_cube((nbins, nbins), spp, span, circle=True)
nd = n.detach()
one = torch.ones_like(x)
def ev(fn, it=20):
    fn(); torch.cuda.synchronize()
    a = torch.cuda.Event(enable_timing=True); b = torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(it): fn()
    b.record(); torch.cuda.synchronize()
    return a.elapsed_time(b) / it
xt, vt = ops.trace(nd, x, v, h, ds)
print(f"  trace {ev(lambda: ops.trace(nd, x, v, h, ds)):.3f} ms, backtrace {ev(lambda: ops.backtrace(nd, xt, vt, one, one, h, ds)):.3f} ms, "
      f"ray generation {ev(lambda: source.rand_rays_cube((nbins, nbins), spp, span, circle=True)):.3f} ms "
      f"(without the circle compaction {ev(lambda: source.rand_rays_cube((nbins, nbins), spp, span, circle=False)):.3f} ms)")

if os.environ.get("PROFILE") == "1":
    import cProfile
    import pstats
    pr = cProfile.Profile()
    pr.enable()
    for _ in range(100):
        iteration(True)
    pr.disable()
    pstats.Stats(pr).sort_stats("cumulative").print_stats(28)
