"""GPU parity tests: the CUDA path (through the C ABI) against the golden vectors of
the reference's CPU tracer and against the C restatement on seeded inputs.

Tolerances (unit-span volumes; `span` = 1):
  * default flags (exact-trajectory arithmetic) and DRRT_REFERENCE_ARITH: ray
    endpoints/directions must be BIT-IDENTICAL to the float32 restatement (same
    operations in the same order) and within 1e-6 of the reference's own CPU build;
    gradients: REFERENCE_ARITH differs from the restatement only by the order of the
    float atomics (rel L2 <= 2e-6); the default additionally evaluates Hessian /
    splat / lambda-mu in factored FMA form (rel L2 <= 1e-5 against the restatement,
    <= 1e-4 = the north-star tolerance against the reference's golden gradients).
  * DRRT_FAST_ARITH (opt-in polynomial arithmetic): float32 rounding differs, so a few
    rays per ten thousand take a different branch at a cell face ("coin flips",
    DESIGN.md section numerics): >= 99.5 % of rays within 1e-5 * span (north-star
    tolerance), every ray within one cell-face kick (2e-2), gradients within
    max(1e-4, 3 x the float32-vs-float64 distance of the reference algorithm itself).
"""
import numpy as np
import pytest
import torch

from conftest import rel_l2, rel_max

pytestmark = pytest.mark.gpu

REF = 1  # DRRT_REFERENCE_ARITH
DET = 2  # DRRT_DETERMINISTIC
FAST = 8  # DRRT_FAST_ARITH
AGG = 16  # DRRT_WARP_AGGREGATION
NOIL = 32  # DRRT_NO_INTERLEAVE
NOFC = 1 << 7  # DRRT_NO_FACE_CARRY
FC = 1 << 24  # DRRT_FACE_CARRY (forced; the default decides from h/ds)


@pytest.fixture(scope="module")
def ops(cuda_lib):
    from adjointnonlinearraytracing_b200 import ops as o
    assert torch.cuda.is_available()
    return o


def T(a):
    return torch.from_numpy(np.ascontiguousarray(a)).cuda()


def N(t):
    return t.detach().cpu().numpy()


def endpoint_ok(a, b, frac=0.995, tol=1e-5, worst=2e-2):
    e = np.abs(a - b).reshape(len(a), -1).max(axis=1)
    assert np.mean(e <= tol) >= frac, (float(np.mean(e <= tol)), float(e.max()))
    assert e.max() <= worst, float(e.max())


@pytest.mark.parametrize("case", ["bump9", "lune17", "uniform9"])
def test_golden_trace_backtrace(case, golden, ops, oracle_mod):
    g = golden(case)
    xt, vt = ops.trace(T(g["rif"]), T(g["pos"]), T(g["vel"]), g["h"], g["ds"], flags=REF)
    assert np.abs(N(xt) - g["xt"]).max() <= 1e-6 and np.abs(N(vt) - g["vt"]).max() <= 1e-6
    xp, vp = oracle_mod.port.trace(g["rif"], g["pos"], g["vel"], g["h"], g["ds"])
    assert np.array_equal(N(xt), xp) and np.array_equal(N(vt), vp)  # bit-identical to the restatement
    grad = ops.backtrace(T(g["rif"]), T(g["xt"]), T(g["vt"]), T(g["dx"]), T(g["dv"]), g["h"], g["ds"], flags=REF)
    assert rel_l2(N(grad), g["grad"]) <= 2e-5
    gp = oracle_mod.port.backtrace(g["rif"], g["xt"], g["vt"], g["dx"], g["dv"], g["h"], g["ds"])
    assert rel_l2(N(grad), gp) <= 2e-6
    # default flags: exact-trajectory arithmetic
    xd, vd = ops.trace(T(g["rif"]), T(g["pos"]), T(g["vel"]), g["h"], g["ds"])
    assert np.array_equal(N(xd), xp) and np.array_equal(N(vd), vp)
    for fl in (0, AGG, NOIL, AGG | NOIL, DET):
        gd = ops.backtrace(T(g["rif"]), T(g["xt"]), T(g["vt"]), T(g["dx"]), T(g["dv"]), g["h"], g["ds"], flags=fl)
        assert rel_l2(N(gd), gp) <= 1e-5, (fl, rel_l2(N(gd), gp))
        assert rel_l2(N(gd), g["grad"]) <= 1e-4
    # opt-in polynomial arithmetic
    xf, vf = ops.trace(T(g["rif"]), T(g["pos"]), T(g["vel"]), g["h"], g["ds"], flags=FAST)
    endpoint_ok(N(xf), g["xt"]); endpoint_ok(N(vf), g["vt"], tol=2e-5)
    g64 = oracle_mod.port64.backtrace(g["rif"], g["xt"], g["vt"], g["dx"], g["dv"], g["h"], g["ds"])
    floor = rel_l2(gp, g64)
    gf = ops.backtrace(T(g["rif"]), T(g["xt"]), T(g["vt"]), T(g["dx"]), T(g["dv"]), g["h"], g["ds"], flags=FAST)
    assert rel_l2(N(gf), g["grad"]) <= max(1e-4, 3 * floor), (rel_l2(N(gf), g["grad"]), floor)


@pytest.mark.parametrize("case", ["bump9", "lune17"])
@pytest.mark.parametrize("flags", [REF, 0, FAST])
def test_golden_variants(case, flags, golden, ops):
    g = golden(case)
    rif, pos, vel = T(g["rif"]), T(g["pos"]), T(g["vel"])
    tol = 1e-6 if flags != FAST else None
    p = golden(case + "_plane")
    xt, vt, m = ops.trace_pln(rif, pos, vel, T(p["pln_o"]), T(p["pln_d"]), g["h"], g["ds"], flags=flags)
    t = golden(case + "_target")
    xg, vg, d2 = ops.trace_target(rif, pos, vel, T(t["target"]), g["h"], g["ds"], flags=flags)
    s = golden(case + "_sdf")
    xs, vs = ops.trace_sdf(rif, T(s["sdf"]), T(s["pos"]), T(s["vel"]), g["h"], g["ds"], flags=flags)
    gs = ops.backtrace_sdf(rif, T(s["sdf"]), T(s["xt"]), T(s["vt"]), T(g["dx"]), T(g["dv"]), g["h"], g["ds"],
                           flags=flags)
    if flags != FAST:
        assert np.abs(N(xt) - p["xt"]).max() <= tol and np.abs(N(vt) - p["vt"]).max() <= tol
        assert np.array_equal(N(m).astype(np.uint8), p["mask"])
        assert np.abs(N(xg) - t["xt"]).max() <= tol and np.abs(N(vg) - t["vt"]).max() <= tol
        assert np.abs(N(d2) - t["dist2"]).max() <= tol
        assert np.abs(N(xs) - s["xt"]).max() <= tol and np.abs(N(vs) - s["vt"]).max() <= tol
        assert rel_l2(N(gs), s["grad"]) <= (2e-5 if flags == REF else 1e-4)
    else:
        endpoint_ok(N(xt), p["xt"]); endpoint_ok(N(vt), p["vt"], tol=2e-5)
        assert np.mean(N(m).astype(np.uint8) == p["mask"]) >= 0.995
        endpoint_ok(N(xg), t["xt"], worst=0.1); endpoint_ok(N(vg), t["vt"], tol=2e-5, worst=0.1)
        endpoint_ok(N(d2), t["dist2"])
        endpoint_ok(N(xs), s["xt"], worst=0.1); endpoint_ok(N(vs), s["vt"], tol=2e-5, worst=0.1)
        assert rel_l2(N(gs), s["grad"]) <= 5e-3


def test_golden_cable(golden, ops, oracle_mod):
    g = golden("cable33")
    args = (T(g["rif"]), float(g["radius"]), float(g["length"]))
    xt, vt, d2 = ops.trace_cable(*args, T(g["pos"]), T(g["vel"]), T(g["target"]), g["ds"])
    xp, vp, dp = oracle_mod.port.trace_cable(g["rif"], g["radius"], g["length"], g["pos"], g["vel"], g["target"], g["ds"])
    assert np.array_equal(N(xt), xp) and np.array_equal(N(vt), vp) and np.array_equal(N(d2), dp)
    assert np.abs(N(xt) - g["xt"]).max() <= 2e-5 and np.abs(N(vt) - g["vt"]).max() <= 2e-5
    assert np.abs(N(d2) - g["dist2"]).max() <= 2e-5
    # default = per-ray register accumulation per radial bin; REF = the reference's four adds per step
    for fl in (0, DET, REF, REF | DET):
        grad = ops.backtrace_cable(*args, T(g["xt"]), T(g["vt"]), T(g["dx"]), T(g["dv"]), g["ds"], flags=fl)
        assert rel_l2(N(grad), g["grad"]) <= 1e-4, fl


@pytest.mark.parametrize("case", ["probe3", "probe5"])
def test_volume_probes(case, golden, ops):
    g = golden(case)
    keep = np.ones(len(g["n"]), bool); keep[g["lattice_points"]] = False
    for fl in (REF, 0, FAST):
        n, gr, hs = ops.volume_eval(T(g["rif"]), T(g["pos"]), g["h"], flags=fl)
        assert np.abs(N(n) - g["n"]).max() <= 1e-6
        assert np.abs(N(gr) - g["grad"])[keep].max() <= 1e-5
        assert np.abs(N(hs) - g["hess"])[keep].max() <= 1e-4
    sp = ops.volume_splat(T(np.zeros_like(g["rif"])), T(g["pos"][1:]), T(g["val"][1:]), T(g["sgrad"][1:]), g["h"])
    assert np.abs(N(sp) - g["splat"]).max() <= 1e-5


def test_cylinder_probe(golden, ops):
    g = golden("cylprobe")
    n, gr, hs = ops.cylinder_eval(T(g["rif"]), g["radius"], g["length"], T(g["pos"]))
    assert np.abs(N(n) - g["n"]).max() <= 1e-5 and np.abs(N(gr) - g["grad"]).max() <= 1e-5
    assert np.abs(N(hs) - g["hess"]).max() <= 1e-4
    sp = ops.cylinder_splat(T(np.zeros_like(g["rif"])), g["radius"], g["length"], T(g["pos"]), T(g["val"]), T(g["sgrad"]))
    assert np.abs(N(sp) - g["splat"]).max() <= 1e-4 * max(1.0, np.abs(g["splat"]).max())


@pytest.mark.parametrize("ng,s,field", [(33, 64, "noise"), (64, 128, "bump"), (48, 96, "luneburg")])
def test_seeded_vs_restatement(ng, s, field, ops, oracle_mod):
    from adjointnonlinearraytracing_b200 import scenes as sc
    rif = getattr(sc, "field_" + field)(ng)
    h = sc.grid_spacing(ng); ds = np.float32(h / 2)
    pos, vel = sc.rays_cone(s, backoff=0.37 * float(ds))
    dx, dv = sc.seeds_normal(s * s)
    xp, vp = oracle_mod.port.trace(rif, pos, vel, h, ds); steps_f = oracle_mod.port.last_steps
    st = ops.new_stats("cuda")
    xt, vt = ops.trace(T(rif), T(pos), T(vel), h, ds, flags=REF, stats=st)
    assert np.array_equal(N(xt), xp) and np.array_equal(N(vt), vp)
    assert int(st[0]) == steps_f and int(st[1]) == 0
    gp = oracle_mod.port.backtrace(rif, xp, vp, dx, dv, h, ds); steps_b = oracle_mod.port.last_steps
    st = ops.new_stats("cuda")
    gg = ops.backtrace(T(rif), T(xp), T(vp), T(dx), T(dv), h, ds, flags=REF, stats=st)
    assert int(st[0]) == steps_b
    assert rel_l2(N(gg), gp) <= 2e-6 and rel_max(N(gg), gp) <= 1e-5
    gd = ops.backtrace(T(rif), T(xp), T(vp), T(dx), T(dv), h, ds, flags=REF | DET)
    assert rel_l2(N(gd), gp) <= 2e-6
    # default flags: exact-trajectory arithmetic -> same path, same step counts
    st = ops.new_stats("cuda")
    xd, vd = ops.trace(T(rif), T(pos), T(vel), h, ds, stats=st)
    assert np.array_equal(N(xd), xp) and np.array_equal(N(vd), vp) and int(st[0]) == steps_f
    for fl in (0, AGG, NOIL, AGG | NOIL, DET):
        st = ops.new_stats("cuda")
        gd = ops.backtrace(T(rif), T(xp), T(vp), T(dx), T(dv), h, ds, flags=fl, stats=st)
        assert int(st[0]) == steps_b
        assert rel_l2(N(gd), gp) <= 1e-5 and rel_max(N(gd), gp) <= 1e-4, (fl, rel_l2(N(gd), gp), rel_max(N(gd), gp))
    # opt-in polynomial arithmetic
    xf, vf = ops.trace(T(rif), T(pos), T(vel), h, ds, flags=FAST)
    endpoint_ok(N(xf), xp); endpoint_ok(N(vf), vp, tol=2e-5)
    g64 = oracle_mod.port64.backtrace(rif, xp, vp, dx, dv, h, ds)
    floor = rel_l2(gp, g64)
    gf = ops.backtrace(T(rif), T(xp), T(vp), T(dx), T(dv), h, ds, flags=FAST)
    assert rel_l2(N(gf), gp) <= max(1e-4, 3 * floor), (rel_l2(N(gf), gp), floor)
    if oracle_mod.ref is not None:  # the reference library itself, when it travelled
        xr, vr = oracle_mod.ref.trace(rif, pos, vel, h, ds)
        endpoint_ok(N(xd), xr); endpoint_ok(N(xt), xr)


@pytest.mark.parametrize("step_res", [8, 32])
def test_fine_steps_many_per_cell(step_res, ops, oracle_mod):
    """ds = h/8 and h/32 (the reference's Luneburg yaml runs step_res 32): the register cell cache and
    the face-carry flush (forced here: from 6 steps per cell on the default flushes all eight corners) hold a
    cell for dozens of steps; path and gradient still match the restatement."""
    from adjointnonlinearraytracing_b200 import scenes as sc
    ng, s = 17, 40
    rif = sc.field_luneburg(ng)
    h = sc.grid_spacing(ng); ds = np.float32(h / step_res)
    pos, vel = sc.rays_cone(s, backoff=0.37 * float(ds))
    dx, dv = sc.seeds_normal(s * s)
    xp, vp = oracle_mod.port.trace(rif, pos, vel, h, ds); steps_f = oracle_mod.port.last_steps
    st = ops.new_stats("cuda")
    xd, vd = ops.trace(T(rif), T(pos), T(vel), h, ds, stats=st)
    assert np.array_equal(N(xd), xp) and np.array_equal(N(vd), vp) and int(st[0]) == steps_f
    gp = oracle_mod.port.backtrace(rif, xp, vp, dx, dv, h, ds); steps_b = oracle_mod.port.last_steps
    for fl in (REF, 0, NOFC, FC, FC | AGG, NOIL, AGG, DET, DET | FC):
        st = ops.new_stats("cuda")
        gd = ops.backtrace(T(rif), T(xp), T(vp), T(dx), T(dv), h, ds, flags=fl, stats=st)
        assert int(st[0]) == steps_b
        assert rel_l2(N(gd), gp) <= 2e-5 and rel_max(N(gd), gp) <= 2e-4, (fl, rel_l2(N(gd), gp), rel_max(N(gd), gp))


def test_noncubic_and_ragged(ops, oracle_mod):
    """W != H != D, ray count not a multiple of the warp/CTA size, SoA layout."""
    rng = np.random.default_rng(11)
    W, H, D = 20, 13, 9  # res triple as the reference passes it: idx = (z*H + y)*W + x
    rif = (1.0 + 0.2 * rng.uniform(0, 1, W * H * D)).astype(np.float32)
    h = np.float32(0.0625); ds = np.float32(0.03125)
    n = 1001
    pos = (rng.uniform(0, 1, (n, 3)) * np.array([(W - 1), (H - 1), (D - 1)]) * h).astype(np.float32)
    vel = rng.standard_normal((n, 3)).astype(np.float32); vel /= np.linalg.norm(vel, axis=1, keepdims=True)
    xp, vp = oracle_mod.port.trace(rif, pos, vel, h, ds, res=(W, H, D))
    xt, vt = ops.trace(T(rif), T(pos), T(vel), h, ds, res=(W, H, D), flags=REF)
    assert np.array_equal(N(xt), xp) and np.array_equal(N(vt), vp)
    xs, vs = ops.trace(T(rif), T(pos.T.copy()), T(vel.T.copy()), h, ds, res=(W, H, D), flags=REF, layout="soa")
    assert np.array_equal(N(xs).T, xp) and np.array_equal(N(vs).T, vp)
    dx, dv = rng.standard_normal((n, 3)).astype(np.float32), rng.standard_normal((n, 3)).astype(np.float32)
    gp = oracle_mod.port.backtrace(rif, xp, vp, dx, dv, h, ds, res=(W, H, D))
    xd, vd = ops.trace(T(rif), T(pos), T(vel), h, ds, res=(W, H, D))
    assert np.array_equal(N(xd), xp) and np.array_equal(N(vd), vp)
    for fl in (REF, REF | DET, 0, DET, AGG, NOIL, FAST, FAST | DET):
        gg = ops.backtrace(T(rif), T(xp), T(vp), T(dx), T(dv), h, ds, res=(W, H, D), flags=fl)
        assert rel_l2(N(gg), gp) <= (2e-6 if fl & REF else 5e-3 if fl & FAST else 1e-5), (fl, rel_l2(N(gg), gp))


def test_edge_cases(ops, oracle_mod):
    from adjointnonlinearraytracing_b200 import scenes as sc
    rif = sc.field_bump(9); h = np.float32(0.125); ds = np.float32(0.0625)
    # empty batch
    e = torch.zeros(0, 3, device="cuda")
    xt, vt = ops.trace(T(rif), e, e, h, ds)
    assert xt.shape == (0, 3)
    g0 = ops.backtrace(T(rif), e, e, e, e, h, ds)
    assert float(g0.abs().max()) == 0.0 and g0.shape == rif.shape
    # rays that start outside and never hit, rays that start outside moving away, rays at rest
    pos = np.array([[-0.5, 0.5, 0.5], [0.5, 0.5, -0.2], [2.0, 2.0, 2.0], [0.5, 0.5, 0.5]], np.float32)
    vel = np.array([[0, 1, 0], [0, 0, -1], [1, 1, 1], [0, 0, 0]], np.float32)
    xp, vp = oracle_mod.port.trace(rif, pos, vel, h, ds)
    st = ops.new_stats("cuda")
    xt, vt = ops.trace(T(rif), T(pos), T(vel), h, ds, flags=REF, stats=st)
    assert np.array_equal(N(xt), xp) and np.array_equal(N(vt), vp)
    assert int(st[1]) >= 1  # the ray at rest (and the one flying parallel outside) never exits
    # resolution mismatch: the reference's RuntimeError message
    with pytest.raises(RuntimeError, match="Resolution doesn't match data"):
        ops.trace(T(rif), T(pos), T(vel), h, ds, res=(9, 9, 8))
    with pytest.raises(RuntimeError, match="invalid resolution"):
        ops.trace(T(rif.reshape(-1)[:9 * 9]), T(pos), T(vel), h, ds, res=(1, 9, 9))


def test_autograd_wrappers_match_reference_contract(ops, oracle_mod, golden):
    """BackTracerC.apply(rif, x, v, h, ds) -> (xt, vt); backward returns d/d rif only
    (reference core/tracer.py:294-335)."""
    from adjointnonlinearraytracing_b200 import tracer, drrt
    g = golden("bump9")
    rif = T(g["rif"]).requires_grad_(True)
    x, v = T(g["pos"]).requires_grad_(True), T(g["vel"])
    tracer.FLAGS = REF
    try:
        xt, vt = tracer.BackTracerC.apply(rif, x, v, float(g["h"]), float(g["ds"]))
        (xt * T(g["dx"])).sum().add((vt * T(g["dv"])).sum()).backward()
    finally:
        tracer.FLAGS = 0
    assert x.grad is None
    assert rel_l2(N(rif.grad), g["grad"]) <= 2e-5
    # pybind-style class API, flat rif + res triple
    tc = drrt.TracerC(); tc.flags = REF
    xt2, vt2 = tc.trace(T(g["rif"]).flatten(), g["rif"].shape, T(g["pos"]), T(g["vel"]), g["h"], g["ds"])
    assert np.array_equal(N(xt2), N(xt))
    gflat = tc.backtrace(T(g["rif"]).flatten(), g["rif"].shape, xt2, vt2, T(g["dx"]), T(g["dv"]), g["h"], g["ds"])
    assert gflat.dim() == 1 and rel_l2(N(gflat).reshape(g["rif"].shape), g["grad"]) <= 2e-5
    # host-buffer flavour (TracerS -> drrt_*_host)
    ts = drrt.TracerS(); ts.flags = REF
    xs, vs = ts.trace(torch.from_numpy(g["rif"]).flatten(), g["rif"].shape, torch.from_numpy(g["pos"]),
                      torch.from_numpy(g["vel"]), g["h"], g["ds"])
    assert np.array_equal(xs.numpy(), N(xt))
    gs = ts.backtrace(g["rif"].reshape(-1), g["rif"].shape, g["xt"], g["vt"], g["dx"], g["dv"], g["h"], g["ds"])
    assert rel_l2(gs.numpy().reshape(g["rif"].shape), g["grad"]) <= 2e-5
