"""Dev script (GPU box): the kernels either side of the tracer at the bench
configuration's sizes (256^3 grid, 2048^2 = 4 M rays), each against the HBM roofline of its algorithmic
bytes (MEASURED_PEAKS.json hbm_gbs, else 6549.4 GB/s).

    python tools/gpu_rows.py
"""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
import torch

import __graft_entry__ as ge

ge.build()
from adjointnonlinearraytracing_b200 import ops, scenes as sc, sensor, source  # noqa: E402

dev = torch.device("cuda:0")
T = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(dev)
try:
    PEAK = float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"])
except Exception:
    PEAK = 6549.4
FLUSH = torch.empty(256 << 20, dtype=torch.uint8, device=dev)  # > the 126 MB L2


def timeit(fn, iters=7, flush=True):
    fn(); torch.cuda.synchronize()
    ts = []
    for _ in range(iters):
        if flush:
            FLUSH.zero_()
        a = torch.cuda.Event(enable_timing=True); b = torch.cuda.Event(enable_timing=True)
        a.record(); fn(); b.record(); torch.cuda.synchronize(); ts.append(a.elapsed_time(b))
    return float(np.median(ts))


def row(name, ms, nbytes, note=""):
    gbs = nbytes / ms * 1e-6
    print(f"{name:58s} {ms:8.3f} ms | {nbytes/1e6:9.1f} MB algorithmic | {gbs:7.0f} GB/s = {100*gbs/PEAK:5.1f} % of HBM {note}", flush=True)


ng, s = 256, 2048
N, G = s * s, ng ** 3
h = float(sc.grid_spacing(ng)); ds = h / 2
print(f"# HBM peak {PEAK:.1f} GB/s; L2 flushed (256 MB memset) before every timed call; median of 7", flush=True)

# ---- f1: fused Adam step, up-res
n = T(sc.field_bump(ng)); g = torch.randn_like(n) * 1e-3; m = torch.zeros_like(n); v = torch.zeros_like(n)
step = [0]
def adam():
    step[0] += 1
    ops.adam_step_(n, g, m, v, step[0], 1e-3)
row("f1 drrt_adam_step 256^3 (boundary mask + Adam + clamp)", timeit(adam), G * 28, "(4 reads + 3 writes per voxel)")
# what the reference runs: mask multiply, torch.optim.Adam (foreach), clamp
p = torch.nn.Parameter(n.clone()); opt = torch.optim.Adam([p], lr=1e-3)
mask = torch.zeros_like(n); mask[1:-1, 1:-1, 1:-1] = 1
def torch_adam():
    p.grad = g * mask
    opt.step()
    with torch.no_grad():
        p.clamp_(min=1.0)
row("   torch: grad*mask, optim.Adam.step, clamp_ (reference)", timeit(torch_adam), G * 28)
lo = T(sc.field_bump(128))
row("f1 drrt_upres_volume 128^3 -> 256^3 (float64 lerp)", timeit(lambda: ops.upres_volume(lo, [ng] * 3)), G * 4 + 128 ** 3 * 4)

# ---- forward pass of the bench workload: inputs of f2 / f4
rif = T(sc.field_bump(ng))
pos, vel = sc.rays_collimated(s)
pos, vel = T(pos), T(vel)
xt, vt = ops.trace(rif, pos, vel, h, ds)

# ---- f2: sensor epilogue (near field, 512^2 image over the exit face)
plane = (torch.tensor([[0.5, 0.5, 1.0]], device=dev), torch.tensor([[0.0, 0.0, 1.0]], device=dev))
res = 512
x_ = xt.clone().requires_grad_(True); v_ = vt.clone().requires_grad_(True)
img = sensor.generate_sensor((x_, v_), 1.0, plane, res, 1.0)
row("f2 drrt_sensor_splat 4 M rays -> 512^2 image", timeit(lambda: sensor.generate_sensor((xt, vt), 1.0, plane, res, 1.0)),
    N * 24 + res * res * 4, f"(+ 16 REDs per ray into the L2-resident image; image sum {float(img.sum()):.4g})")
dimg = torch.randn_like(img)
def splat_bwd():
    x_.grad = None; v_.grad = None
    i = sensor.generate_sensor((x_, v_), 1.0, plane, res, 1.0)
    i.backward(dimg)
t_fb = timeit(splat_bwd)
t_f = timeit(lambda: sensor.generate_sensor((x_, v_), 1.0, plane, res, 1.0))
row("f2 drrt_sensor_splat_backward (dL/dx, dL/dv)", t_fb - t_f, N * 48 + res * res * 4, "(forward + backward timed together, forward subtracted)")
far = sensor.generate_inf_sensor((xt, vt), 1.0, plane, res, 120)
row("f2 far-field drrt_sensor_splat (generate_inf_sensor)", timeit(lambda: sensor.generate_inf_sensor((xt, vt), 1.0, plane, res, 120)),
    N * 24 + res * res * 4)
tex = torch.rand(res, res, device=dev)
row("f2 drrt_sensor_gather (get_sdf_vals_near)", timeit(lambda: sensor.get_sdf_vals_near((xt, vt), tex, plane, 1.0)),
    N * 28 + res * res * 4)

# ---- f3: ray generation (Philox in the kernel)
row("f3 drrt_source_rand plane 2048^2 x 1 spp", timeit(lambda: source.plane_source3_rand(0.0, (s, s), 1, 1.0, seed=1)),
    N * 60, "(x, v, planes: 60 B out per ray)")
row("f3 32 views x 64^2 x 16 spp (rand_rays_in_sphere)", timeit(lambda: source.rand_rays_in_sphere(32, (64, 64), 16, 1.0, seed=1)),
    32 * 64 * 64 * 16 * 60)
row("f3 six cube faces 128^2 x 10 spp + circle compaction", timeit(lambda: source.rand_rays_cube((128, 128), 10, 1.0, circle=True, seed=1)),
    6 * 128 * 128 * 10 * (61 + 0.785 * 120), "(keep mask + index_select gather of the survivors; one host sync)")

# ---- f4: launch-state gradients next to the plain adjoint
one = torch.ones_like(pos)
grad = torch.empty_like(rif)
t_plain = timeit(lambda: ops.backtrace(rif, xt, vt, one, one, h, ds, out=grad), iters=5)
t_rays = timeit(lambda: ops.backtrace_rays(rif, xt, vt, one, one, h, ds, pos0=pos), iters=5)
print(f"f4 drrt_backtrace_rays (grid gradient + dL/dpos, dL/dvel)      {t_rays:8.3f} ms vs drrt_backtrace {t_plain:8.3f} ms "
      f"(+{100*(t_rays/t_plain-1):.1f} %)", flush=True)
