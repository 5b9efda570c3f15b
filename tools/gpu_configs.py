"""Dev script (GPU box): forward + adjoint timings on the shapes of BASELINE.json's five configs
(synthetic fields; default flags).  Output goes to profiles/ as the per-config table of DESIGN.md.

    python tools/gpu_configs.py
"""
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


def timeit(fn, iters=5):
    fn(); torch.cuda.synchronize()
    ts = []
    for _ in range(iters):
        a = torch.cuda.Event(enable_timing=True); b = torch.cuda.Event(enable_timing=True)
        a.record(); fn(); b.record(); torch.cuda.synchronize(); ts.append(a.elapsed_time(b))
    return float(np.median(ts))


def multiview_rays(n_views, s, spp, seed=0):
    """n_views collimated s x s sensors around the volume centre, spp jittered samples per pixel
    (the shape of path_matrix/run_fuel_injection_2008.py:80-92)."""
    rng = np.random.default_rng(seed)
    u = (np.arange(s) + 0.5) / s
    yy, xx = np.meshgrid(u, u, indexing="ij")
    P, V = [], []
    for k in range(n_views):
        th = np.pi * k / n_views
        d = np.array([np.sin(th), 0.0, np.cos(th)])
        e1 = np.array([np.cos(th), 0.0, -np.sin(th)]); e2 = np.array([0.0, 1.0, 0.0])
        for _ in range(spp):
            jx = (rng.uniform(-0.5, 0.5, xx.shape)) / s; jy = (rng.uniform(-0.5, 0.5, xx.shape)) / s
            a = (xx + jx - 0.5) * 0.7; b = (yy + jy - 0.5) * 0.7
            p = 0.5 + a[..., None] * e1 + b[..., None] * e2 - 0.9 * d
            P.append(p.reshape(-1, 3)); V.append(np.tile(d, (s * s, 1)))
    return np.concatenate(P).astype(np.float32), np.concatenate(V).astype(np.float32)


def grid_case(name, ng, pos, vel, field=sc.field_bump, step_res=2):
    rif = T(field(ng)); h = float(sc.grid_spacing(ng)); ds = h / step_res
    pos, vel = (p if torch.is_tensor(p) else T(p) for p in (pos, vel))
    one = torch.ones_like(pos)
    st = ops.new_stats(dev); xt, vt = ops.trace(rif, pos, vel, h, ds, stats=st); sf, ne = int(st[0]), int(st[1])
    tf = timeit(lambda: ops.trace(rif, pos, vel, h, ds))
    st = ops.new_stats(dev); ops.backtrace(rif, xt, vt, one, one, h, ds, stats=st); sb = int(st[0])
    tb = timeit(lambda: ops.backtrace(rif, xt, vt, one, one, h, ds))
    print(f"{name:34s} grid {ng}^3 rays {pos.shape[0]:9d} | fwd {tf:8.3f} ms {sf/tf/1e6:6.1f} G/s | adj {tb:8.3f} ms "
          f"{sb/tb/1e6:6.1f} G/s | fwd+adj {(sf+sb)/(tf+tb)/1e6:6.1f} G ray-steps/s | not exited {ne}", flush=True)


def cable_case(name, res, s, length=5.0):
    prof = T(sc.cable_profile(res)); radius = 1.0
    ds = radius / res / 2  # core/fiber_opt.py:156
    pos, vel = sc.cable_rays(s, radius)
    pos, vel = T(pos), T(vel)
    tgt = torch.tensor([radius, length, radius], device=dev).repeat(pos.shape[0], 1)
    one = torch.ones_like(pos)
    st = ops.new_stats(dev); xt, vt, d2 = ops.trace_cable(prof, radius, length, pos, vel, tgt, ds, stats=st); sf = int(st[0])
    tf = timeit(lambda: ops.trace_cable(prof, radius, length, pos, vel, tgt, ds), iters=3)
    st = ops.new_stats(dev); ops.backtrace_cable(prof, radius, length, xt, vt, one, one, ds, stats=st); sb = int(st[0])
    tb = timeit(lambda: ops.backtrace_cable(prof, radius, length, xt, vt, one, one, ds), iters=3)
    print(f"{name:34s} profile {res} rays {pos.shape[0]:9d} | fwd {tf:8.3f} ms {sf/tf/1e6:6.1f} G/s | adj {tb:8.3f} ms "
          f"{sb/tb/1e6:6.1f} G/s | fwd+adj {(sf+sb)/(tf+tb)/1e6:6.1f} G ray-steps/s", flush=True)


def luneburg_yaml_rays():
    """data/params-luneburg.yaml: nbins 128, spp 10, six cube faces, circle filter, random rotation --
    generated on the device (source.rand_rays_cube)."""
    from adjointnonlinearraytracing_b200 import source
    import time
    torch.manual_seed(0); np.random.seed(0)
    rot = source.random_rotmat().numpy()
    ts = []
    for _ in range(4):  # the first call pays torch's lazy kernel loading
        torch.cuda.synchronize(); t0 = time.perf_counter()
        (x, v, planes), nr = source.rand_rays_cube((128, 128), 10, 1.0, circle=True, rotmat=rot)
        torch.cuda.synchronize(); ts.append((time.perf_counter() - t0) * 1e3)
    print(f"   (device ray generation: {x.shape[0]} rays in {min(ts[1:]):.2f} ms incl. circle compaction and its one host "
          f"sync; first call {ts[0]:.1f} ms)")
    return x, v


def main():
    grid_case("1 Luneburg 32^3, 256^2 rays", 32, *sc.rays_collimated(256), field=sc.field_luneburg)
    grid_case("1' Luneburg yaml 129^3 ds=h/32", 129, *luneburg_yaml_rays(), field=sc.field_luneburg, step_res=32)
    grid_case("2 fibre (grid form) 64^3, 512^2", 64, *sc.rays_collimated(512))
    cable_case("2 fibre (cable) 129 samples, 512^2", 129, 512)
    grid_case("3 fuel injection 65^3, 32x64^2x16", 65, *multiview_rays(32, 64, 16))
    grid_case("4 caustic 256^3, 2048^2", 256, *sc.rays_collimated(2048))
    grid_case("4' caustic 256^3, 2048^2 cone .05", 256, *sc.rays_cone(2048, spread=0.05))
    grid_case("4'' caustic 256^3, 2048^2 cone .15", 256, *sc.rays_cone(2048, spread=0.15))
    grid_case("5 focal stack 512^3, 2M rays/GPU", 512, *sc.rays_collimated(1448))
    grid_case("5' focal stack 512^3, 4096^2", 512, *sc.rays_collimated(4096))


if __name__ == "__main__":
    main()
