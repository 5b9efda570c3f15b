"""GPU: the two-rays-per-thread forward kernel (DRRT_PAIR_RAYS, trace_pair_kernel) must return exactly the
bits of the one-ray-per-thread kernel -- same float32 operations in the same order, only issued as packed
FP32 -- on every kind of batch: odd ray counts, rays launched outside the box, rays that miss it, rays
that do not exit before the step cap, both layouts, and against the CPU restatement."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu
PAIR, NOPAIR, REF = 1 << 25, 1 << 26, 1


def _bits(t):
    return t.contiguous().view(torch.int32)


def _same(a, b):
    return bool(torch.equal(_bits(a), _bits(b)))


def _both(ops, rif, pos, vel, h, ds, **kw):
    sa, sb = ops.new_stats("cuda"), ops.new_stats("cuda")
    xa, va = ops.trace(rif, pos, vel, h, ds, flags=NOPAIR, stats=sa, **kw)
    xb, vb = ops.trace(rif, pos, vel, h, ds, flags=PAIR, stats=sb, **kw)
    torch.cuda.synchronize()
    assert _same(xa, xb) and _same(va, vb)
    assert int(sa[0]) == int(sb[0]) and int(sa[1]) == int(sb[1])
    return xa, va, int(sa[0]), int(sa[1])


@pytest.mark.parametrize("ng,s,spread", [(17, 32, 0.15), (33, 61, 0.4), (65, 128, 0.05), (129, 200, 0.0)])
def test_pair_matches_single_cone(cuda_lib, ng, s, spread):
    from adjointnonlinearraytracing_b200 import ops, scenes as sc
    rif = torch.from_numpy(sc.field_bump(ng)).cuda()
    h = float(sc.grid_spacing(ng)); ds = h / 2
    if spread:
        pos, vel = sc.rays_cone(s, spread=spread, backoff=0.37 * ds)
    else:
        pos, vel = sc.rays_collimated(s)
    pos, vel = torch.from_numpy(pos).cuda(), torch.from_numpy(vel).cuda()
    for n in (pos.shape[0], pos.shape[0] - 1, 63, 65, 33, 1):  # whole / ragged warps, a lone ray
        _both(ops, rif, pos[:n].contiguous(), vel[:n].contiguous(), h, ds)
    xs, vs = ops.trace(rif, pos.t().contiguous(), vel.t().contiguous(), h, ds, flags=PAIR, layout="soa")
    xa, va = ops.trace(rif, pos, vel, h, ds, flags=NOPAIR)
    assert _same(xs.t(), xa) and _same(vs.t(), va)


def test_pair_random_rays_from_outside_and_missing(cuda_lib):
    from adjointnonlinearraytracing_b200 import ops, scenes as sc
    g = torch.Generator().manual_seed(5)
    for ng in (9, 24, 40):
        rif = torch.from_numpy(sc.field_bump(ng)).cuda()
        h = float(sc.grid_spacing(ng)); span = (ng - 1) * h
        for ds in (h / 2, h / 7, 1.3 * h):
            n = 4099
            # launch points in a shell around (and inside) the box; directions at random: some rays start
            # inside, some enter, some miss, some fly away at once
            pos = (torch.rand(n, 3, generator=g) * 1.6 - 0.3) * span
            vel = torch.randn(n, 3, generator=g)
            vel = vel / vel.norm(dim=1, keepdim=True)
            vel[::17, 0] = 0.0  # axis-aligned components, signed zeros
            vel[5::17, 1] = -0.0
            pos[3::29] = pos[3::29].clamp(0, span)  # exactly on faces
            _, _, steps, ne = _both(ops, rif, pos.cuda(), vel.cuda(), h, ds)
            assert steps > 0


def test_pair_step_cap_not_exited(cuda_lib):
    from adjointnonlinearraytracing_b200 import ops, scenes as sc
    ng = 33
    # a strong lens traps / delays rays: with a long step count some do not exit before the cap
    rif = torch.from_numpy(sc.field_bump(ng)).cuda() * 1.0
    h = float(sc.grid_spacing(ng)); ds = h / 2
    pos, vel = sc.rays_cone(64, spread=0.3, backoff=0.37 * ds)
    pos, vel = torch.from_numpy(pos).cuda(), torch.from_numpy(vel).cuda()
    vel = vel * 0.05  # slow rays: ~20x the steps, far beyond the cap of 4 h max(res) / ds
    _, _, steps, ne = _both(ops, rif, pos, vel, h, ds)
    assert ne > 0


def test_pair_against_cpu_restatement(cuda_lib):
    import oracle
    from adjointnonlinearraytracing_b200 import ops, scenes as sc
    ng, s = 21, 48
    rif = sc.field_bump(ng)
    h = float(sc.grid_spacing(ng)); ds = h / 2
    pos, vel = sc.rays_cone(s, spread=0.2, backoff=0.37 * ds)
    xr, vr = oracle.port.trace(rif, pos, vel, np.float32(h), np.float32(ds))
    xt, vt = ops.trace(torch.from_numpy(rif).cuda(), torch.from_numpy(pos).cuda(), torch.from_numpy(vel).cuda(),
                       h, ds, flags=PAIR)
    assert np.array_equal(xt.cpu().numpy(), xr) and np.array_equal(vt.cpu().numpy(), vr)


def test_pair_full_size_matches(cuda_lib):
    from adjointnonlinearraytracing_b200 import ops, scenes as sc
    ng, s = 256, 512
    rif = torch.from_numpy(sc.field_bump(ng)).cuda()
    h = float(sc.grid_spacing(ng)); ds = h / 2
    pos, vel = sc.rays_cone(s, spread=0.05, backoff=0.37 * ds)
    _both(ops, rif, torch.from_numpy(pos).cuda(), torch.from_numpy(vel).cuda(), h, ds)


def test_pair_guard_bands(cuda_lib):
    """Outputs of a ragged batch (the last thread has one ray or none, the last warp is partial) end exactly
    where they should: sentinel bands around xt / vt stay untouched, in both layouts, through the C ABI."""
    from ctypes import c_void_p
    from adjointnonlinearraytracing_b200 import _capi, ops, scenes as sc
    lib = _capi.load()
    ng = 19
    rif = torch.from_numpy(sc.field_bump(ng)).cuda()
    h = float(sc.grid_spacing(ng)); ds = h / 2
    pos, vel = sc.rays_cone(40, spread=0.3, backoff=0.37 * ds)
    pos, vel = torch.from_numpy(pos).cuda(), torch.from_numpy(vel).cuda()
    P = lambda t, off=0: c_void_p(t.data_ptr() + off)
    pad = 1024
    for n in (1, 31, 32, 33, 63, 64, 65, 97, 1599):
        for lay in (3, 1):
            p = pos[:n].contiguous() if lay == 3 else pos[:n].t().contiguous()
            v = vel[:n].contiguous() if lay == 3 else vel[:n].t().contiguous()
            ref_x, ref_v = ops.trace(rif, p, v, h, ds, flags=NOPAIR, layout="aos" if lay == 3 else "soa")
            xb = torch.full((3 * n + 2 * pad,), 7.5, device="cuda")
            vb = torch.full((3 * n + 2 * pad,), -3.25, device="cuda")
            rc = lib.drrt_trace(P(rif), ng, ng, ng, P(p), P(v), n, lay, h, ds, P(xb, 4 * pad), P(vb, 4 * pad),
                                c_void_p(0), PAIR, c_void_p(torch.cuda.current_stream().cuda_stream))
            assert rc == 0
            torch.cuda.synchronize()
            assert bool((xb[:pad] == 7.5).all()) and bool((xb[-pad:] == 7.5).all())
            assert bool((vb[:pad] == -3.25).all()) and bool((vb[-pad:] == -3.25).all())
            assert _same(xb[pad:-pad].view_as(ref_x), ref_x) and _same(vb[pad:-pad].view_as(ref_v), ref_v)
