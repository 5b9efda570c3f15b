"""GPU: randomised shapes against the C restatement.  Each seed draws grid dimensions, spacing, step size,
field roughness and a ray batch that starts inside, on and outside the volume; the default kernels must
reproduce the restatement's ray paths bit for bit (all forward variants) and its gradients to float-atomic
rounding, under every accumulation mode."""
import numpy as np
import pytest
import torch

from conftest import rel_l2

pytestmark = pytest.mark.gpu
REF, DET, AGG, NOIL, NOFC, FC, NOREP = 1, 2, 16, 32, 1 << 7, 1 << 24, 1 << 19


def T(a):
    return torch.from_numpy(np.ascontiguousarray(a)).cuda()


def N(t):
    return t.detach().cpu().numpy()


def draw(seed):
    rng = np.random.default_rng(1000 + seed)
    W, H, D = (int(v) for v in rng.integers(5, 34, 3))
    if seed % 4 == 0:  # sometimes large enough for the private gradient copies (G >= 4096)
        W, H, D = max(W, 17), max(H, 17), max(D, 17)
    h = np.float32(rng.uniform(0.02, 0.3))
    step_res = [1, 2, 2, 3, 5, 8, 13][seed % 7]
    ds = np.float32(h / step_res)
    ext = np.array([(W - 1), (H - 1), (D - 1)], np.float64) * float(h)
    # smooth bump + voxel noise, n >= 1
    zz, yy, xx = np.meshgrid(np.arange(D), np.arange(H), np.arange(W), indexing="ij")
    c = ext / 2
    r2 = ((xx * h - c[0]) / ext[0]) ** 2 + ((yy * h - c[1]) / ext[1]) ** 2 + ((zz * h - c[2]) / ext[2]) ** 2
    rif = (1.0 + rng.uniform(0.05, 0.3) * np.exp(-r2 / 0.08) + rng.uniform(0, 0.02) * rng.uniform(0, 1, r2.shape))
    rif = rif.astype(np.float32).reshape(-1)
    n = int(rng.integers(200, 3000))
    kind = rng.integers(0, 3, n)
    pos = rng.uniform(0, 1, (n, 3)) * ext                      # inside
    face = rng.integers(0, 3, n)
    on = kind == 1                                              # on a face (lattice plane)
    pos[on, face[on]] = 0.0
    out = kind == 2                                             # outside, flying in
    pos[out, face[out]] = -rng.uniform(0.1, 3.0, out.sum()) * float(ds)
    vel = rng.standard_normal((n, 3))
    vel[np.arange(n), face] = np.abs(vel[np.arange(n), face]) + 0.5
    vel /= np.linalg.norm(vel, axis=1, keepdims=True)
    return (W, H, D), h, ds, rif, pos.astype(np.float32), vel.astype(np.float32), rng


@pytest.mark.parametrize("seed", range(21))
def test_random_shapes_match_the_restatement(seed, cuda_lib, oracle_mod):
    from adjointnonlinearraytracing_b200 import ops
    res, h, ds, rif, pos, vel, rng = draw(seed)
    n = len(pos)
    port = oracle_mod.port
    xp, vp = port.trace(rif, pos, vel, h, ds, res=res); sf = port.last_steps
    for fl in (0, REF):
        st = ops.new_stats("cuda")
        xt, vt = ops.trace(T(rif), T(pos), T(vel), h, ds, res=res, flags=fl, stats=st)
        assert np.array_equal(N(xt), xp) and np.array_equal(N(vt), vp) and int(st[0]) == sf, (seed, fl)
    if seed % 3 == 0:  # SoA layout and the host-buffer entry points (TracerS)
        from adjointnonlinearraytracing_b200 import drrt
        xs_, vs_ = ops.trace(T(rif), T(pos.T.copy()), T(vel.T.copy()), h, ds, res=res, layout="soa")
        assert np.array_equal(N(xs_).T, xp) and np.array_equal(N(vs_).T, vp)
        ts = drrt.TracerS()
        xh, vh = ts.trace(rif, res, pos, vel, h, ds)
        assert np.array_equal(xh.numpy(), xp) and np.array_equal(vh.numpy(), vp) and ts.last_stats[0] == sf
    # plane and sdf variants
    # planes behind the ray, inside the volume and beyond it; normals along the ray, tilted and facing it
    po = (pos + (rng.uniform(-2, 40, (n, 1)) * float(h)).astype(np.float32) * vel).astype(np.float32)
    pd = (vel + rng.uniform(0, 1.5) * rng.standard_normal((n, 3))).astype(np.float32)
    xq, vq, mq = port.trace_pln(rif, pos, vel, po, pd, h, ds, res=res)
    xg, vg, mg = ops.trace_pln(T(rif), T(pos), T(vel), T(po), T(pd), h, ds, res=res)
    assert np.array_equal(N(xg), xq) and np.array_equal(N(vg), vq) and np.array_equal(N(mg).astype(np.uint8), mq.astype(np.uint8))
    sdf = (rng.uniform(-1, 1, rif.shape) * 0.05 + 0.01).astype(np.float32)
    xs, vs = port.trace_sdf(rif, sdf, pos, vel, h, ds, res=res)
    xsg, vsg = ops.trace_sdf(T(rif), T(sdf), T(pos), T(vel), h, ds, res=res)
    assert np.array_equal(N(xsg), xs) and np.array_equal(N(vsg), vs)
    # adjoint, every accumulation mode
    dx = rng.standard_normal((n, 3)).astype(np.float32); dv = rng.standard_normal((n, 3)).astype(np.float32)
    gp = port.backtrace(rif, xp, vp, dx, dv, h, ds, res=res); sb = port.last_steps
    for fl in (0, REF, DET, AGG, AGG | (15 << 8), NOIL, NOFC, FC, FC | DET, NOREP, 3 << 20):
        st = ops.new_stats("cuda")
        gg = ops.backtrace(T(rif), T(xp), T(vp), T(dx), T(dv), h, ds, res=res, flags=fl, stats=st)
        assert int(st[0]) == sb, (seed, fl)
        assert rel_l2(N(gg), gp) <= 2e-5, (seed, fl, rel_l2(N(gg), gp))
    if seed % 3 == 0:
        gsoa = ops.backtrace(T(rif), T(xp.T.copy()), T(vp.T.copy()), T(dx.T.copy()), T(dv.T.copy()), h, ds, res=res,
                             layout="soa")
        assert rel_l2(N(gsoa), gp) <= 2e-5
        gh = ts.backtrace(rif, res, xp, vp, dx, dv, h, ds)
        assert rel_l2(gh.numpy().reshape(-1), gp.reshape(-1)) <= 2e-5
    gs = port.backtrace_sdf(rif, sdf, xs, vs, dx, dv, h, ds, res=res)
    gsg = ops.backtrace_sdf(T(rif), T(sdf), T(xs), T(vs), T(dx), T(dv), h, ds, res=res)
    assert rel_l2(N(gsg), gs) <= 2e-5 or np.abs(gs).max() == 0
    # launch-state gradients
    gr, dpos, dvel = port.backtrace_rays(rif, xp, vp, dx, dv, h, ds, pos0=pos, res=res)
    g2, dp2, dv2 = ops.backtrace_rays(T(rif), T(xp), T(vp), T(dx), T(dv), h, ds, pos0=T(pos), res=res)
    assert rel_l2(N(g2), gr) <= 2e-5
    assert rel_l2(N(dp2), dpos) <= 1e-4 and rel_l2(N(dv2), dvel) <= 1e-4


@pytest.mark.parametrize("seed", range(8))
def test_random_target_batches_match_the_restatement(seed, cuda_lib, oracle_mod):
    """trace_target is batch-coupled (every ray is tested for a new closest approach until the LAST ray has
    escaped): random batches, targets inside and beyond the volume."""
    from adjointnonlinearraytracing_b200 import ops
    res, h, ds, rif, pos, vel, rng = draw(100 + seed)
    n = len(pos)
    ext = (np.array(res) - 1) * float(h)
    target = (rng.uniform(-0.3, 1.3, (n, 3)) * ext).astype(np.float32)
    xp, vp, dp = oracle_mod.port.trace_target(rif, pos, vel, target, h, ds, res=res)
    for fl in (0, REF):
        xt, vt, d2 = ops.trace_target(T(rif), T(pos), T(vel), T(target), h, ds, res=res, flags=fl)
        assert np.array_equal(N(xt), xp) and np.array_equal(N(vt), vp) and np.array_equal(N(d2), dp), (seed, fl)


@pytest.mark.parametrize("seed", range(10))
def test_random_cables_match_the_restatement(seed, cuda_lib, oracle_mod):
    """Cylinder variant: random profile resolution, radius, length, step and rays entering the face y = 0
    (some grazing the wall, some starting outside the radius)."""
    from adjointnonlinearraytracing_b200 import ops
    rng = np.random.default_rng(2000 + seed)
    res = int(rng.integers(6, 220))
    radius = float(rng.uniform(0.3, 2.0)); length = float(rng.uniform(0.5, 4.0))
    ds = np.float32(radius / res / [0.5, 1, 2, 3][seed % 4])
    r = np.linspace(0, 1, res)
    prof = (rng.uniform(1.2, 1.6) * (1 - rng.uniform(0.01, 0.1) * r * r) + rng.uniform(0, 0.003) * rng.uniform(0, 1, res)).astype(np.float32)
    n = int(rng.integers(100, 1500))
    rr = rng.uniform(0, 1.05, n) ** 0.5 * radius
    th = rng.uniform(0, 2 * np.pi, n)
    pos = np.stack([radius + rr * np.cos(th), rng.uniform(0, 2e-3, n), radius + rr * np.sin(th)], axis=1).astype(np.float32)
    vel = np.array([0, 1, 0]) + rng.uniform(0.01, 0.4) * rng.standard_normal((n, 3))
    vel = (vel / np.linalg.norm(vel, axis=1, keepdims=True)).astype(np.float32)
    target = np.tile(np.array([radius, length, radius], np.float32), (n, 1)) + 0.1 * rng.standard_normal((n, 3)).astype(np.float32)
    port = oracle_mod.port
    xp, vp, dp = port.trace_cable(prof, radius, length, pos, vel, target, ds)
    st = ops.new_stats("cuda")
    xt, vt, d2 = ops.trace_cable(T(prof), radius, length, T(pos), T(vel), T(target), ds, stats=st)
    assert np.array_equal(N(xt), xp) and np.array_equal(N(vt), vp) and np.array_equal(N(d2), dp), seed
    dx = rng.standard_normal((n, 3)).astype(np.float32); dv = rng.standard_normal((n, 3)).astype(np.float32)
    gp = port.backtrace_cable(prof, radius, length, xp, vp, dx, dv, ds)
    for fl in (0, DET, REF, REF | DET):
        gg = ops.backtrace_cable(T(prof), radius, length, T(xp), T(vp), T(dx), T(dv), ds, flags=fl)
        assert rel_l2(N(gg), gp) <= 1e-4, (seed, fl, rel_l2(N(gg), gp))


def _grad_close(a, b, tol=2e-3, frac=0.985):
    scale = max(np.abs(b).max(), 1e-30)
    err = np.abs(a - b).max(axis=1) / scale
    assert np.mean(err <= tol) >= frac, (float(np.mean(err <= tol)), float(err.max()))


@pytest.mark.parametrize("seed", range(10))
def test_random_sensor_planes_match_the_restatement(seed, cuda_lib):
    """Sensor epilogue with oblique planes, explicit and default tangents, odd and even resolutions, per-ray
    energies, rays that miss the sensor; near field, far field and the two texture look-ups."""
    from oracle import sensor_oracle as so
    from adjointnonlinearraytracing_b200 import sensor
    rng = np.random.default_rng(3000 + seed)
    n = int(rng.integers(2000, 20000))
    nn = rng.standard_normal(3); nn[2] = abs(nn[2]) + 0.7; nn /= np.linalg.norm(nn)
    nn = nn[None]
    tg = None
    if seed % 2:
        t = rng.standard_normal(3); t -= (t @ nn[0]) * nn[0]; tg = (t / np.linalg.norm(t))[None]
    p = np.array([[0.5, 0.5, 1.0 + rng.uniform(0, 0.5)]])
    res = int(rng.integers(5, 70)); span = float(rng.uniform(0.6, 2.0))
    x = np.stack([rng.uniform(-0.1, 1.1, n), rng.uniform(-0.1, 1.1, n), rng.uniform(0.6, 1.0, n)], axis=1)
    v = nn[0] + rng.uniform(0.02, 0.3) * rng.standard_normal((n, 3))
    e = rng.uniform(0.2, 2.0, n) if seed % 3 == 0 else float(rng.uniform(0.5, 2.0))
    dimg = rng.standard_normal((res, res))
    for mode in ("near", "far"):
        sp = span if mode == "near" else float(rng.uniform(40, 150))
        img64, dx64, dv64 = so.run(mode, x, v, e, p, nn, tg, res, sp, dimg)
        xg = T(x.astype(np.float32)).requires_grad_(True); vg = T(v.astype(np.float32)).requires_grad_(True)
        eg = T(np.asarray(e, np.float32)) if np.ndim(e) else e
        plane = (T(p.astype(np.float32)), T(nn.astype(np.float32)))
        tt = None if tg is None else T(tg.astype(np.float32))
        if mode == "near":
            img = sensor.generate_sensor((xg, vg), eg, plane, res, sp, tt)
        else:
            img = sensor.generate_inf_sensor((xg, vg), eg, plane, res, angle_span=sp, tangent=tt)
        (img * T(dimg.astype(np.float32))).sum().backward()
        assert np.abs(N(img) - img64).max() <= 5e-5 * max(np.abs(img64).max(), 1e-30), (seed, mode)
        _grad_close(N(vg.grad), dv64)
        if mode == "near":
            _grad_close(N(xg.grad), dx64)
        # texture look-ups at the same coordinates
        tex = rng.uniform(-1, 1, (res, res))
        df = rng.standard_normal(n)
        f64, gx64, gv64 = so.run_gather(mode, x, v, tex, p, nn, tg, sp, df)
        xg = T(x.astype(np.float32)).requires_grad_(True); vg = T(v.astype(np.float32)).requires_grad_(True)
        fn = sensor.get_sdf_vals_near if mode == "near" else sensor.get_sdf_vals_far
        f = fn((xg, vg), T(tex.astype(np.float32)), plane, sp, tt)
        (f * T(df.astype(np.float32))).sum().backward()
        err = np.abs(N(f) - f64)
        assert np.mean(err <= 2e-4) >= 0.99, (seed, mode, float(err.max()))
        _grad_close(N(vg.grad), gv64)
        if mode == "near":
            _grad_close(N(xg.grad), gx64)


@pytest.mark.parametrize("seed", range(10))
def test_random_plane_sources_match_the_restatement(seed, cuda_lib):
    """Ray generation from caller uniforms: random view counts, non-square pixel grids, samples per pixel,
    widths, sensor distances, circle filter, independent jitter, x-axis rotations; the six cube faces with
    the fused rigid rotation."""
    from oracle import source_oracle as so
    from adjointnonlinearraytracing_b200 import source
    rng = np.random.default_rng(4000 + seed)
    nviews = int(rng.integers(1, 8)); P = (int(rng.integers(3, 40)), int(rng.integers(3, 40))); spp = int(rng.integers(1, 6))
    width = float(rng.uniform(0.5, 3.0)); sd = float(rng.uniform(0.0, 1.5)); span = float(rng.choice([90, 180, 270, 360]))
    circle = bool(seed % 2); xaxis = bool((seed // 2) % 2); indep = bool((seed // 4) % 2)
    u = rng.uniform(0, 1, (nviews, 2 * spp, P[0], P[1])).astype(np.float32)
    tol = 3e-6 * max(1.0, width + sd)

    def close(t, ref):
        a = N(t).astype(np.float64)
        assert a.shape == ref.shape, (a.shape, ref.shape)
        assert np.abs(a - ref).max() <= tol, float(np.abs(a - ref).max())

    x64, v64, pl64, nr64 = so.multi_view(so.sphere_views(nviews, span, xaxis), u.astype(np.float64), P, spp, width, circle, sd,
                                         indep)
    (x, v, pl), nr = source.rand_rays_in_sphere(nviews, P, spp, width, angle_span=span, circle=circle, xaxis=xaxis,
                                                sensor_dist=sd, indep=indep, offset=T(u))
    assert nr == nr64
    close(x, x64); close(v, v64); close(pl, pl64)
    # cube faces + rigid rotation about the centre
    q = rng.standard_normal((3, 3)); R, _ = np.linalg.qr(q)
    if np.linalg.det(R) < 0:
        R[:, 0] = -R[:, 0]
    u6 = rng.uniform(0, 1, (6, 2 * spp, P[0], P[1])).astype(np.float32)
    x64, v64, pl64, nr64 = so.multi_view(so.cube_views(), u6.astype(np.float64), P, spp, width, circle, 0.0, False)
    x64, v64, pl64 = so.random_rotate_ic(x64, v64, pl64, width, R)
    (x, v, pl), nr = source.rand_rays_cube(P, spp, width, circle=circle, offset=T(u6), rotmat=R)
    assert nr == nr64
    close(x, x64); close(v, v64); close(pl, pl64)


@pytest.mark.parametrize("seed", range(8))
def test_random_adam_and_upres_shapes_match_torch(seed, cuda_lib):
    """Fused optimiser step on random (non-cubic, odd and multiple-of-4 innermost) shapes against
    torch.optim.Adam + the reference's boundary mask and clamp; up-res against the float64 restatement."""
    from oracle import optim_oracle as oo
    from adjointnonlinearraytracing_b200 import ops
    rng = np.random.default_rng(5000 + seed)
    shape = tuple(int(s) for s in rng.integers(3, 41, 3))
    if seed % 2 == 0:
        shape = shape[:2] + (4 * max(1, shape[2] // 4),)  # the float4 kernel
    lr = float(rng.choice([1e-3, 1e-2, 5e-2])); betas = (float(rng.uniform(0.8, 0.95)), float(rng.uniform(0.99, 0.9999)))
    n0 = (1 + 0.2 * rng.uniform(0, 1, shape)).astype(np.float32)
    p = torch.from_numpy(n0.copy()).cuda().requires_grad_(True)
    opt = torch.optim.Adam([p], lr=lr, betas=betas)
    n = torch.from_numpy(n0.copy()).cuda(); m = torch.zeros_like(n); v = torch.zeros_like(n)
    mask = torch.ones(shape, dtype=torch.bool, device="cuda"); mask[1:-1, 1:-1, 1:-1] = False
    for step in range(1, 6):
        g = torch.from_numpy(rng.standard_normal(shape).astype(np.float32) * 10.0 ** rng.uniform(-4, 1)).cuda()
        p.grad = g.clone(); p.grad[mask] = 0
        opt.step()
        with torch.no_grad():
            p.clamp_(min=1)
        ops.adam_step_(n, g, m, v, step, lr, betas)
        assert float((n - p.detach()).abs().max()) <= 3e-6, (seed, step)
    assert torch.equal(n[mask], torch.from_numpy(n0).cuda()[mask])
    st = opt.state[p]
    assert float((m - st["exp_avg"]).abs().max()) <= 1e-6 * max(1.0, float(st["exp_avg"].abs().max()))
    # up-res of a cubic grid (the reference only up-samples cubes)
    s0 = int(rng.integers(3, 24)); s1 = int(rng.integers(s0, 48))
    a = rng.uniform(1, 2, (s0, s0, s0)).astype(np.float32)
    up = ops.upres_volume(torch.from_numpy(a).cuda(), [s1] * 3)
    ref = oo.upres_volume(a, [s1] * 3)
    assert up.shape == (s1, s1, s1) and np.abs(N(up) - np.asarray(ref)).max() <= 2e-6


@pytest.mark.parametrize("seed", range(8))
def test_random_point_cone_area_sources_match_the_restatement(seed, cuda_lib):
    """The four non-plane source kinds, one random view each, from caller uniforms."""
    from oracle import source_oracle as so
    from adjointnonlinearraytracing_b200 import source
    rng = np.random.default_rng(6000 + seed)
    P = (int(rng.integers(3, 30)), int(rng.integers(3, 30))); spp = int(rng.integers(1, 5))
    n = P[0] * P[1] * spp
    width = float(rng.uniform(0.5, 2.5)); sd = float(rng.uniform(0.1, 1.5)); ang = float(rng.uniform(0, 360))
    xaxis = bool(seed % 2); circle = bool((seed // 2) % 2); cone = float(rng.uniform(10, 170))
    R = so.rotmat(np.float32(ang), xaxis)
    tol = 4e-6 * max(1.0, width + sd)

    def close(t, ref, scale=1.0):
        a = N(t).astype(np.float64)
        assert a.shape == ref.shape, (a.shape, ref.shape)
        assert np.abs(a - ref).max() <= tol * scale, float(np.abs(a - ref).max())

    u = rng.uniform(0, 1, (2 * spp, P[0], P[1])).astype(np.float32)
    x64, v64, pl64, keep = so.point_source3_rand(R, u.astype(np.float64), P, spp, width, circle, sd)
    x, v, pl = source.point_source3_rand(ang, P, spp, width, circle=circle, xaxis=xaxis, sensor_dist=sd, offset=T(u))
    close(x, x64); close(v, v64); close(pl, pl64)
    u = rng.uniform(0, 1, (2, n)).astype(np.float32)
    x64, v64, pl64, _ = so.cone_source3_rand(R, u.astype(np.float64), P, spp, width, sd, cone)
    x, v, pl = source.cone_source3_rand(ang, P, spp, width, xaxis=xaxis, sensor_dist=sd, cone_angle=cone, offset=T(u))
    close(x, x64); close(v, v64, 2.0); close(pl, pl64)
    u = rng.uniform(0, 1, (4, n)).astype(np.float32)
    cone2 = float(rng.uniform(10, 150))  # tpv = sensor_dist / v.y: keep v.y away from 0
    x64, v64, pl64, tpv64 = so.area_source3_cone(R, u.astype(np.float64), P, spp, width, sd, cone2)
    (x, v, pl), tpv = source.area_source3_cone(ang, P, spp, width, xaxis=xaxis, sensor_dist=sd, cone_angle=cone2, offset=T(u))
    close(x, x64); close(v, v64, 2.0); close(pl, pl64)
    assert np.abs(N(tpv) - tpv64).max() <= 1e-5 * np.abs(tpv64).max()
    x64, v64, pl64, xt64, tpv64 = so.area_source3_rand_bias(R, u.astype(np.float64), P, spp, width, sd)
    (x, v, pl), xt, tpv = source.area_source3_rand_bias(ang, P, spp, width, xaxis=xaxis, sensor_dist=sd, offset=T(u))
    close(x, x64, 2.0); close(v, v64); close(pl, pl64); close(xt, xt64)
    assert np.abs(N(tpv) - tpv64).max() <= 1e-5 * np.abs(tpv64).max()


@pytest.mark.parametrize("seed", range(6))
def test_degenerate_rays_match_the_restatement(seed, cuda_lib, oracle_mod):
    """Rays at rest, on the far faces and corners, sliding along faces and edges, axis-aligned on lattice
    lines, and step sizes from far larger than the volume down to h/40 (loop caps computed in float)."""
    from adjointnonlinearraytracing_b200 import ops
    rng = np.random.default_rng(7000 + seed)
    W, H, D = (int(v) for v in rng.integers(3, 20, 3))
    h = np.float32([0.125, 0.1, 1.0 / 3, 0.25, 0.07, 1.0][seed])
    ds = np.float32(h * [0.5, 40.0, 1.0, 0.025, 3.0, 0.5][seed])
    ext = np.array([(W - 1), (H - 1), (D - 1)], np.float64) * float(h)
    rif = (1.0 + 0.3 * rng.uniform(0, 1, W * H * D)).astype(np.float32)
    P, V = [], []
    for _ in range(40):
        c = rng.integers(0, 2, 3) * ext                                    # corners
        P.append(c); V.append(rng.standard_normal(3))
        f = rng.uniform(0, 1, 3) * ext; k = rng.integers(0, 3); f[k] = ext[k] * rng.integers(0, 2)   # on a face
        d = rng.standard_normal(3); d[k] = 0.0                              # sliding along it
        P.append(f); V.append(d)
        lat = np.floor(rng.uniform(0, 1, 3) * (np.array([W, H, D]) - 1)) * float(h)   # lattice point, axis-aligned
        a = np.zeros(3); a[rng.integers(0, 3)] = rng.choice([-1.0, 1.0])
        P.append(lat); V.append(a)
        P.append(rng.uniform(0, 1, 3) * ext); V.append(np.zeros(3))        # at rest
        P.append(rng.uniform(-0.5, 1.5, 3) * ext); V.append(rng.standard_normal(3) * rng.choice([1e-3, 1.0, 30.0]))
    pos = np.array(P, np.float32); vel = np.array(V, np.float32)
    res = (W, H, D)
    port = oracle_mod.port
    xp, vp = port.trace(rif, pos, vel, h, ds, res=res); sf = port.last_steps
    st = ops.new_stats("cuda")
    xt, vt = ops.trace(T(rif), T(pos), T(vel), h, ds, res=res, stats=st)
    assert np.array_equal(N(xt), xp) and np.array_equal(N(vt), vp) and int(st[0]) == sf
    n = len(pos)
    dx = rng.standard_normal((n, 3)).astype(np.float32); dv = rng.standard_normal((n, 3)).astype(np.float32)
    gp = port.backtrace(rif, xp, vp, dx, dv, h, ds, res=res); sb = port.last_steps
    for fl in (0, REF, DET, FC, NOFC, AGG):
        st = ops.new_stats("cuda")
        gg = ops.backtrace(T(rif), T(xp), T(vp), T(dx), T(dv), h, ds, res=res, flags=fl, stats=st)
        assert int(st[0]) == sb, (seed, fl, int(st[0]), sb)
        assert rel_l2(N(gg), gp) <= 2e-5 or np.abs(gp).max() == 0, (seed, fl, rel_l2(N(gg), gp))
    tgt = (rng.uniform(-0.2, 1.2, (n, 3)) * ext).astype(np.float32)
    xq, vq, dq = port.trace_target(rif, pos, vel, tgt, h, ds, res=res)
    xg, vg, dg = ops.trace_target(T(rif), T(pos), T(vel), T(tgt), h, ds, res=res)
    assert np.array_equal(N(xg), xq) and np.array_equal(N(vg), vq) and np.array_equal(N(dg), dq)
