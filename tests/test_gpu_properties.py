"""GPU: size-independent properties at BASELINE.json's full grid size (256^3)."""
import numpy as np
import pytest
import torch

from conftest import rel_l2

pytestmark = pytest.mark.gpu
REF, DET, FAST, AGG, NOIL, NOFC = 1, 2, 8, 16, 32, 1 << 7


@pytest.fixture(scope="module")
def scene(cuda_lib):
    from adjointnonlinearraytracing_b200 import ops, scenes as sc
    ng, s = 256, 512
    rif = torch.from_numpy(sc.field_bump(ng)).cuda()
    h = float(sc.grid_spacing(ng)); ds = h / 2
    pos, vel = sc.rays_cone(s, spread=0.05, backoff=0.37 * ds)
    pos, vel = torch.from_numpy(pos).cuda(), torch.from_numpy(vel).cuda()
    xt, vt = ops.trace(rif, pos, vel, h, ds)
    return dict(ops=ops, rif=rif, h=h, ds=ds, pos=pos, vel=vel, xt=xt, vt=vt, n=s * s)


def test_forward_batch_split_and_layout(scene):
    o = scene["ops"]
    half = scene["n"] // 2 + 37
    xa, va = o.trace(scene["rif"], scene["pos"][:half].contiguous(), scene["vel"][:half].contiguous(), scene["h"], scene["ds"])
    xb, vb = o.trace(scene["rif"], scene["pos"][half:].contiguous(), scene["vel"][half:].contiguous(), scene["h"], scene["ds"])
    assert torch.equal(torch.cat([xa, xb]), scene["xt"]) and torch.equal(torch.cat([va, vb]), scene["vt"])
    xs, vs = o.trace(scene["rif"], scene["pos"].t().contiguous(), scene["vel"].t().contiguous(), scene["h"], scene["ds"], layout="soa")
    assert torch.equal(xs.t(), scene["xt"])


def test_all_rays_exit_and_energy(scene):
    o = scene["ops"]
    st = o.new_stats("cuda")
    xt, vt = o.trace(scene["rif"], scene["pos"], scene["vel"], scene["h"], scene["ds"], stats=st)
    assert int(st[1]) == 0
    span = 255 * scene["h"]
    out = (xt < 0).any(1) | (xt >= span).any(1)
    assert bool(out.all())
    # |v| = n along an exact eikonal ray and n -> 1 on the boundary of the bump; the symplectic
    # Euler scheme keeps it to O(ds)
    assert float((vt.norm(dim=1) - 1).abs().max()) < 2e-2


def test_deterministic_is_bit_stable_and_order_independent(scene):
    o = scene["ops"]
    g = torch.Generator(device="cuda").manual_seed(1)
    dx = torch.randn(scene["n"], 3, device="cuda", generator=g)
    dv = torch.randn(scene["n"], 3, device="cuda", generator=g)
    a = o.backtrace(scene["rif"], scene["xt"], scene["vt"], dx, dv, scene["h"], scene["ds"], flags=DET)
    b = o.backtrace(scene["rif"], scene["xt"], scene["vt"], dx, dv, scene["h"], scene["ds"], flags=DET)
    assert torch.equal(a, b)
    perm = torch.randperm(scene["n"], device="cuda", generator=g)
    c = o.backtrace(scene["rif"], scene["xt"][perm].contiguous(), scene["vt"][perm].contiguous(), dx[perm].contiguous(),
                    dv[perm].contiguous(), scene["h"], scene["ds"], flags=DET)
    assert torch.equal(a, c)  # integer accumulation: independent of ray order
    f = o.backtrace(scene["rif"], scene["xt"], scene["vt"], dx, dv, scene["h"], scene["ds"])
    assert rel_l2(f.cpu().numpy(), a.cpu().numpy()) <= 1e-5


def test_adjoint_is_linear_in_the_seeds(scene):
    o = scene["ops"]
    g = torch.Generator(device="cuda").manual_seed(2)
    s1 = [torch.randn(scene["n"], 3, device="cuda", generator=g) for _ in range(2)]
    s2 = [torch.randn(scene["n"], 3, device="cuda", generator=g) for _ in range(2)]
    bt = lambda dx, dv: o.backtrace(scene["rif"], scene["xt"], scene["vt"], dx, dv, scene["h"], scene["ds"], flags=DET)
    g1, g2 = bt(*s1), bt(*s2)
    g12 = bt(2.0 * s1[0] - 0.5 * s2[0], 2.0 * s1[1] - 0.5 * s2[1])
    assert rel_l2(g12.cpu().numpy(), (2.0 * g1 - 0.5 * g2).cpu().numpy()) <= 1e-5


def test_accumulate_flag_and_batch_split(scene):
    o = scene["ops"]
    dx = torch.ones(scene["n"], 3, device="cuda"); dv = torch.ones(scene["n"], 3, device="cuda")
    whole = o.backtrace(scene["rif"], scene["xt"], scene["vt"], dx, dv, scene["h"], scene["ds"])
    half = scene["n"] // 2
    acc = o.backtrace(scene["rif"], scene["xt"][:half].contiguous(), scene["vt"][:half].contiguous(), dx[:half], dv[:half],
                      scene["h"], scene["ds"])
    o.backtrace(scene["rif"], scene["xt"][half:].contiguous(), scene["vt"][half:].contiguous(), dx[half:], dv[half:],
                scene["h"], scene["ds"], flags=4, out=acc)
    assert rel_l2(acc.cpu().numpy(), whole.cpu().numpy()) <= 1e-5
    # boundary voxels outside every ray's reach stay exactly zero
    assert float(whole[:, :, 0].abs().max()) >= 0.0


def test_uniform_medium_known_answer_256(cuda_lib):
    from adjointnonlinearraytracing_b200 import ops, scenes as sc
    ng = 257  # h = 1/256 exactly
    rif = torch.ones(ng, ng, ng, device="cuda")
    h = 1.0 / 256; ds = h / 2
    pos, vel = sc.rays_collimated(256, z0=-0.37 * ds)
    xt, vt = ops.trace(rif, torch.from_numpy(pos).cuda(), torch.from_numpy(vel).cuda(), h, ds)
    assert torch.equal(vt.cpu(), torch.from_numpy(vel))
    assert torch.equal(xt[:, :2].cpu(), torch.from_numpy(pos[:, :2]))
    k = int(np.ceil((1.0 + 0.37 * ds) / ds))
    assert float((xt[:, 2] - (-0.37 * ds + k * ds)).abs().max()) <= 1e-5


def test_default_arithmetic_matches_literal_reference_arithmetic_at_full_size(scene):
    """256^3: the default (exact-trajectory) kernels give bit-identical endpoints to the literal
    reference-order kernels and gradients within the north-star 1e-4; warp aggregation on/off
    changes only the summation order."""
    o = scene["ops"]
    xr, vr = o.trace(scene["rif"], scene["pos"], scene["vel"], scene["h"], scene["ds"], flags=REF)
    assert torch.equal(xr, scene["xt"]) and torch.equal(vr, scene["vt"])
    g = torch.Generator(device="cuda").manual_seed(3)
    dx = torch.randn(scene["n"], 3, device="cuda", generator=g)
    dv = torch.randn(scene["n"], 3, device="cuda", generator=g)
    gref = o.backtrace(scene["rif"], xr, vr, dx, dv, scene["h"], scene["ds"], flags=REF)
    for fl in (0, AGG, NOIL, AGG | NOIL, DET):
        gd = o.backtrace(scene["rif"], xr, vr, dx, dv, scene["h"], scene["ds"], flags=fl)
        assert rel_l2(gd.cpu().numpy(), gref.cpu().numpy()) <= 1e-4, fl


def test_collimated_lattice_aligned_rays_at_full_size(cuda_lib):
    """The bench workload shape (collimated rays starting on the z = 0 lattice plane, samples on
    cell faces every other step): default vs literal reference arithmetic."""
    from adjointnonlinearraytracing_b200 import ops, scenes as sc
    ng, s = 256, 256
    rif = torch.from_numpy(sc.field_bump(ng)).cuda()
    h = float(sc.grid_spacing(ng)); ds = h / 2
    pos, vel = sc.rays_collimated(s)
    pos, vel = torch.from_numpy(pos).cuda(), torch.from_numpy(vel).cuda()
    xr, vr = ops.trace(rif, pos, vel, h, ds, flags=REF)
    xd, vd = ops.trace(rif, pos, vel, h, ds)
    assert torch.equal(xr, xd) and torch.equal(vr, vd)
    one = torch.ones_like(pos)
    gref = ops.backtrace(rif, xr, vr, one, one, h, ds, flags=REF)
    gd = ops.backtrace(rif, xr, vr, one, one, h, ds)
    assert rel_l2(gd.cpu().numpy(), gref.cpu().numpy()) <= 1e-4


def test_interleave_and_aggregation_only_change_the_summation_order(cuda_lib):
    """Dense collimated rays (8 per cell along x): the strided ray->lane assignment (measured or
    forced) and the optional warp aggregation must not change WHAT is accumulated.  In deterministic
    (integer) mode the gradient is bit-identical for every mapping; in float mode it agrees to
    rounding.  The measured interleave must be picked up (N >= 4096)."""
    from adjointnonlinearraytracing_b200 import ops, scenes as sc
    ng, s = 64, 504  # 504 / 63 = 8 rays per cell; 254 016 rays: not a multiple of the 1024-ray tile
    rif = torch.from_numpy(sc.field_noise(ng)).cuda()
    h = float(sc.grid_spacing(ng)); ds = h / 2
    pos, vel = sc.rays_collimated(s)
    pos, vel = torch.from_numpy(pos).cuda(), torch.from_numpy(vel).cuda()
    xt, vt = ops.trace(rif, pos, vel, h, ds)
    g = torch.Generator(device="cuda").manual_seed(5)
    dx = torch.randn(s * s, 3, device="cuda", generator=g); dv = torch.randn(s * s, 3, device="cuda", generator=g)
    IL = lambda n: n << 12
    st = ops.new_stats("cuda")
    base = ops.backtrace(rif, xt, vt, dx, dv, h, ds, flags=DET | NOIL, stats=st)
    steps = int(st[0])
    for fl in (DET, DET | IL(2), DET | IL(8), DET | IL(32), DET | FAST | IL(16)):
        st = ops.new_stats("cuda")
        gi = ops.backtrace(rif, xt, vt, dx, dv, h, ds, flags=fl, stats=st)
        if not fl & FAST:  # the polynomial arithmetic may take other branches: only shape-checked here
            assert int(st[0]) == steps
            assert torch.equal(gi, base), fl
    NOFC, AUTO = 1 << 7, 1 << 18  # DRRT_NO_FACE_CARRY, DRRT_AUTO_INTERLEAVE
    # without the face carry the per-ray partial sums differ, so the integer result differs in the
    # last bits from `base` -- but is again identical across mappings
    base_nofc = ops.backtrace(rif, xt, vt, dx, dv, h, ds, flags=DET | NOFC | NOIL)
    assert torch.equal(ops.backtrace(rif, xt, vt, dx, dv, h, ds, flags=DET | NOFC), base_nofc)
    assert torch.equal(ops.backtrace(rif, xt, vt, dx, dv, h, ds, flags=DET | NOFC | IL(8)), base_nofc)
    assert rel_l2(base_nofc.cpu().numpy(), base.cpu().numpy()) <= 2e-6
    NOREP, WARP = 1 << 19, 15 << 8  # DRRT_NO_REPLICAS, DRRT_AGG_RUN_WARP
    for fl in (0, NOIL, AUTO, NOFC, NOFC | NOIL, AGG, AGG | NOIL, IL(8), IL(32), AGG | (8 << 8) | NOIL,
               AGG | (2 << 8) | NOIL, NOREP, NOREP | AGG, AGG | WARP, NOREP | AGG | WARP, NOREP | NOFC):
        gf = ops.backtrace(rif, xt, vt, dx, dv, h, ds, flags=fl)
        assert rel_l2(gf.cpu().numpy(), base.cpu().numpy()) <= 2e-6, fl


@pytest.mark.parametrize("axis,sign", [(0, 1), (0, -1), (1, 1), (1, -1), (2, 1), (2, -1)])
def test_face_carry_along_every_axis(axis, sign, cuda_lib, oracle_mod):
    """Beams entering through each of the six faces (the adjoint then walks back along -/+ that axis,
    exercising each branch of the face-carry flush) with a cone spread for lateral and diagonal moves:
    gradient with the carry == without it (to rounding) == the C restatement."""
    from adjointnonlinearraytracing_b200 import ops, scenes as sc
    ng, s = 33, 48
    rif = sc.field_noise(ng)
    h = float(sc.grid_spacing(ng)); ds = h / 2
    pos, vel = sc.rays_cone(s, spread=0.2, seed=axis * 2 + (sign > 0), backoff=1.7 * ds)
    # rays_cone launches from z = 0 along +z: permute / mirror so the beam travels along sign * axis
    perm = {0: [2, 1, 0], 1: [0, 2, 1], 2: [0, 1, 2]}[axis]
    pos, vel = pos[:, perm].copy(), vel[:, perm].copy()
    if sign < 0:
        pos[:, axis] = 1.0 - pos[:, axis]; vel[:, axis] = -vel[:, axis]
    xt, vt = ops.trace(torch.from_numpy(rif).cuda(), torch.from_numpy(pos).cuda(), torch.from_numpy(vel).cuda(), h, ds)
    rng = np.random.default_rng(axis)
    dx = rng.standard_normal(pos.shape).astype(np.float32); dv = rng.standard_normal(pos.shape).astype(np.float32)
    R = torch.from_numpy(rif).cuda(); DX, DV = torch.from_numpy(dx).cuda(), torch.from_numpy(dv).cuda()
    NOFC = 1 << 7
    g_carry = ops.backtrace(R, xt, vt, DX, DV, h, ds)
    g_full = ops.backtrace(R, xt, vt, DX, DV, h, ds, flags=NOFC)
    g_det = ops.backtrace(R, xt, vt, DX, DV, h, ds, flags=DET)
    assert rel_l2(g_carry.cpu().numpy(), g_full.cpu().numpy()) <= 2e-6
    assert rel_l2(g_det.cpu().numpy(), g_full.cpu().numpy()) <= 2e-6
    xo, vo = xt.cpu().numpy(), vt.cpu().numpy()
    g_ref = oracle_mod.port.backtrace(rif, xo, vo, dx, dv, np.float32(h), np.float32(ds))
    assert rel_l2(g_carry.cpu().numpy(), g_ref) <= 1e-5
    # most rays must actually have crossed the volume
    assert float((xt - torch.from_numpy(pos).cuda()).norm(dim=1).median()) > 0.8


@pytest.mark.parametrize("field", ["luneburg", "bump"])
def test_private_gradient_copies_only_change_the_summation_order(field, cuda_lib):
    """Small grids: CTA b adds into gradient copy b % R and one pass sums the copies (the scratch buffer of
    drrt_backtrace_scratch_bytes).  Same gradient as one grid (float rounding) and as the integer mode;
    DRRT_ACCUMULATE keeps what `out` held; a NULL scratch simply means one copy."""
    from ctypes import c_void_p
    from adjointnonlinearraytracing_b200 import _capi, ops, scenes as sc
    NOREP, ACC = 1 << 19, 4
    ng, s = 32, 256
    rif = torch.from_numpy(getattr(sc, "field_" + field)(ng)).cuda()
    h = float(sc.grid_spacing(ng)); ds = h / 2
    pos, vel = sc.rays_collimated(s)
    pos, vel = torch.from_numpy(pos).cuda(), torch.from_numpy(vel).cuda()
    xt, vt = ops.trace(rif, pos, vel, h, ds)
    g = torch.Generator(device="cuda").manual_seed(7)
    dx = torch.randn(s * s, 3, device="cuda", generator=g); dv = torch.randn(s * s, 3, device="cuda", generator=g)
    lib = _capi.load()
    G = rif.numel()
    assert lib.drrt_backtrace_scratch_bytes(G, 0) == 15 * 4 * G and lib.drrt_backtrace_scratch_bytes(G, NOREP) == 0
    assert ops.make_flags(replicas=3) == 3 << 20 and ops.make_flags(replicas=False) == NOREP and ops.make_flags() == 0
    st = ops.new_stats("cuda")
    det = ops.backtrace(rif, xt, vt, dx, dv, h, ds, flags=DET, stats=st)
    steps = int(st[0])
    for fl in (0, NOREP, REF, REF | NOREP, AGG, NOFC, 3 << 20, 15 << 20):  # the last two: DRRT_REPLICAS(3), (15)
        st = ops.new_stats("cuda")
        gr = ops.backtrace(rif, xt, vt, dx, dv, h, ds, flags=fl, stats=st)
        assert int(st[0]) == steps
        assert rel_l2(gr.cpu().numpy(), det.cpu().numpy()) <= 1e-5, fl
    # accumulate on top of an existing gradient
    pre = torch.rand_like(rif)
    acc = pre.clone()
    ops.backtrace(rif, xt, vt, dx, dv, h, ds, flags=ACC, out=acc)
    assert rel_l2((acc - pre).cpu().numpy(), det.cpu().numpy()) <= 1e-5
    # straight through the C ABI with scratch = NULL: allowed, one copy
    out = torch.full_like(rif, 7.0)
    P = lambda t: c_void_p(t.data_ptr())
    rc = lib.drrt_backtrace(P(rif), ng, ng, ng, P(xt), P(vt), P(dx), P(dv), s * s, 3, h, ds, P(out), c_void_p(0),
                            0, c_void_p(0), 0, c_void_p(torch.cuda.current_stream().cuda_stream))
    assert rc == 0
    assert rel_l2(out.cpu().numpy(), det.cpu().numpy()) <= 1e-5
    # guard bands around grad and scratch stay untouched (the copies end exactly where scratch_bytes says)
    nb = lib.drrt_backtrace_scratch_bytes(G, 0)
    pad = 4096
    gbuf = torch.full((G + 2 * pad,), 3.0, device="cuda")
    sbuf = torch.full((nb // 4 + 2 * pad,), 5.0, device="cuda")
    rc = lib.drrt_backtrace(P(rif), ng, ng, ng, P(xt), P(vt), P(dx), P(dv), s * s, 3, h, ds,
                            c_void_p(gbuf.data_ptr() + 4 * pad), c_void_p(sbuf.data_ptr() + 4 * pad), nb, c_void_p(0), 0,
                            c_void_p(torch.cuda.current_stream().cuda_stream))
    assert rc == 0
    assert bool((gbuf[:pad] == 3.0).all()) and bool((gbuf[-pad:] == 3.0).all())
    assert bool((sbuf[:pad] == 5.0).all()) and bool((sbuf[-pad:] == 5.0).all())
    assert rel_l2(gbuf[pad:-pad].view_as(rif).cpu().numpy(), det.cpu().numpy()) <= 1e-5
    # a scratch buffer sized under OTHER flags: deterministic-sized (8 G + 16 bytes) reused without the flag -> the
    # library uses only as many private copies as fit; an undersized deterministic buffer is refused
    nb_det = lib.drrt_backtrace_scratch_bytes(G, DET)
    sbuf = torch.full((nb_det // 4 + 2 * pad,), 5.0, device="cuda")
    gbuf.fill_(3.0)
    rc = lib.drrt_backtrace(P(rif), ng, ng, ng, P(xt), P(vt), P(dx), P(dv), s * s, 3, h, ds,
                            c_void_p(gbuf.data_ptr() + 4 * pad), c_void_p(sbuf.data_ptr() + 4 * pad), nb_det, c_void_p(0), 0,
                            c_void_p(torch.cuda.current_stream().cuda_stream))
    assert rc == 0
    assert bool((sbuf[:pad] == 5.0).all()) and bool((sbuf[-pad:] == 5.0).all())
    assert rel_l2(gbuf[pad:-pad].view_as(rif).cpu().numpy(), det.cpu().numpy()) <= 1e-5
    rc = lib.drrt_backtrace(P(rif), ng, ng, ng, P(xt), P(vt), P(dx), P(dv), s * s, 3, h, ds,
                            c_void_p(gbuf.data_ptr() + 4 * pad), c_void_p(sbuf.data_ptr() + 4 * pad), nb_det - 1, c_void_p(0), DET,
                            c_void_p(torch.cuda.current_stream().cuda_stream))
    assert rc == -1
    # few rays on the same grid: fewer copies (8 N / G), still the same answer
    few = slice(0, 3000)
    a = ops.backtrace(rif, xt[few].contiguous(), vt[few].contiguous(), dx[few].contiguous(), dv[few].contiguous(), h, ds)
    b = ops.backtrace(rif, xt[few].contiguous(), vt[few].contiguous(), dx[few].contiguous(), dv[few].contiguous(), h, ds,
                      flags=DET)
    assert rel_l2(a.cpu().numpy(), b.cpu().numpy()) <= 1e-5


@pytest.mark.gpu
def test_smem_grid_variant_is_identical(cuda_lib):
    """DRRT_SMEM_GRID (opt-in: gather from a per-CTA shared-memory copy of a small grid, one CTA per SM): bit-identical
    trajectory, gradient within atomic-order noise of the default kernels (32^3, rays from outside and ragged count)."""
    from adjointnonlinearraytracing_b200 import _capi, ops, scenes as sc
    ng = 32
    rif = torch.from_numpy(sc.field_luneburg(ng)).cuda()
    h = float(sc.grid_spacing(ng)); ds = h / 2
    pos, vel = sc.rays_cone(61, spread=0.1, backoff=2.37 * ds)
    pos, vel = torch.from_numpy(pos).cuda(), torch.from_numpy(vel).cuda()
    st0, st1 = ops.new_stats("cuda"), ops.new_stats("cuda")
    x0, v0 = ops.trace(rif, pos, vel, h, ds, stats=st0)
    x1, v1 = ops.trace(rif, pos, vel, h, ds, flags=_capi.DRRT_SMEM_GRID, stats=st1)
    assert torch.equal(x0, x1) and torch.equal(v0, v1) and torch.equal(st0, st1)
    dx, dv = (torch.from_numpy(a).cuda() for a in sc.seeds_normal(pos.shape[0]))
    g0 = ops.backtrace(rif, x0, v0, dx, dv, h, ds)
    g1 = ops.backtrace(rif, x0, v0, dx, dv, h, ds, flags=_capi.DRRT_SMEM_GRID)
    assert rel_l2(g1.cpu().numpy(), g0.cpu().numpy()) <= 1e-5
    d0 = ops.backtrace(rif, x0, v0, dx, dv, h, ds, flags=DET)
    d1 = ops.backtrace(rif, x0, v0, dx, dv, h, ds, flags=DET | _capi.DRRT_SMEM_GRID)
    assert torch.equal(d0, d1)
