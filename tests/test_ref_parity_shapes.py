"""Parity against the REFERENCE LIBRARY ITSELF (oracle/_ref = the reference's Tracer<false,false> compiled from
/root/reference; it travels to the GPU box as a built .so) at the shapes the numbers are quoted on
(VERDICT r1 "next" 1): BASELINE.json configs[0..3] incl. the benchmarked batch -- collimated rays launched ON
the z = 0 lattice plane with ds = h/2 -- a backed-off batch where no lattice excuse applies, the oblique cone
batch, the 32-view tomography batch and the cable.  ENDPOINTS AND GRADIENTS, >= 16 k-ray subsamples.

Stated tolerances (north star): endpoints and directions 1e-5 of the span, gradients 1e-4 rel L2.

Which reference build.  The reference computes the cell of a sample from p * rcp(h) (src/volume.cpp:128) and its
three builds approximate 1/h three different ways: enoki's AVX2 packets (rcp + Newton), AVX-512 packets (rcp14 +
Newton) and the CUDA JIT (rcp.approx.ftz).  For every spacing probed (h = 1/(N-1), N = 9 .. 512) exactly ONE of
the two CPU builds returns the correctly rounded 1/h, which is what the CUDA kernels and the restatement use
(AVX-512 at N = 9, 17, 33, 64, 65, 129, 256, 512; AVX2 at N = 32, 48, 128).  A collimated batch launched on a
lattice plane with ds = h/2 puts every other sample exactly on a cell face, where that last ulp picks the cell;
the reference's own two builds therefore disagree WITH EACH OTHER by 5e-3 .. 3e-2 in the gradient on such batches
(recorded by `test_reference_builds_disagree_on_lattice_batch`).  The tests compare against the build that shares
1/h (`oracle.ref_for_spacing`): there the GPU trajectory is BIT-IDENTICAL (asserted) and the gradient differs only
by the order of the atomics and the FMA-form Hessian / lambda-mu (measured <= 3e-7, asserted <= 1e-4).
"""
import numpy as np
import pytest

from conftest import rel_l2
from adjointnonlinearraytracing_b200 import scenes as sc

N_SUB = 16384
END_TOL, GRAD_TOL = 1e-5, 1e-4


def _sample(case, n, seeds):
    idx = sc.subsample(case["pos"].shape[0], n)
    pos = np.ascontiguousarray(case["pos"][idx]); vel = np.ascontiguousarray(case["vel"][idx])
    if seeds == "ones":
        dx, dv = sc.seeds_ones(len(idx))
    else:
        dx, dv = sc.seeds_normal(len(idx), seed=7)
    return idx, pos, vel, dx, dv


# ---------------------------------------------------------------- CPU: the restatement against the library at 256^3
@pytest.mark.parametrize("name,n", [("caustic256", 1024), ("luneburg32", 8192), ("fuel65", 4096)])
def test_restatement_bit_identical_to_reference_at_config_shapes(name, n, oracle_mod):
    case = sc.config_case(name)
    ref = oracle_mod.ref_for_spacing(case["h"])
    if ref is None:
        pytest.skip("no oracle/_ref build with rcp(h) == 1/h for this spacing on this host")
    _, pos, vel, dx, dv = _sample(case, n, "ones")
    xr, vr = ref.trace(case["rif"], pos, vel, case["h"], case["ds"])
    xp, vp = oracle_mod.port.trace(case["rif"], pos, vel, case["h"], case["ds"])
    assert np.array_equal(xr, xp) and np.array_equal(vr, vp)
    gr = ref.backtrace(case["rif"], xr, vr, dx, dv, case["h"], case["ds"])
    gp = oracle_mod.port.backtrace(case["rif"], xp, vp, dx, dv, case["h"], case["ds"])
    assert np.array_equal(gr, gp)


def test_reference_builds_disagree_on_lattice_batch(oracle_mod):
    """The noise floor of the benchmarked batch: the reference's own AVX2 and AVX-512 builds."""
    b = oracle_mod.ref_builds()
    if len(b) < 2:
        pytest.skip("needs both ISA builds of oracle/_ref on an AVX-512 host")
    v3, v4 = b["v3"], b["v4"]
    assert v3.rcp(sc.grid_spacing(256)) != v4.rcp(sc.grid_spacing(256))
    case = sc.config_case("caustic256")
    _, pos, vel, dx, dv = _sample(case, 1024, "ones")
    out = []
    for lib in (v3, v4):
        xt, vt = lib.trace(case["rif"], pos, vel, case["h"], case["ds"])
        out.append((xt, lib.backtrace(case["rif"], xt, vt, dx, dv, case["h"], case["ds"])))
    e = np.abs(out[0][0] - out[1][0]).max(axis=1)
    floor = rel_l2(out[0][1], out[1][1])
    print(f"reference AVX2 vs AVX-512 on the bench batch: endpoint max {e.max():.2e}, gradient rel L2 {floor:.2e}")
    assert floor > GRAD_TOL  # this is why the bar is stated against the AVX-512 build


# ---------------------------------------------------------------- GPU
def T(a):
    import torch
    return torch.from_numpy(np.ascontiguousarray(a)).cuda()


def _check_against(ref, xt, vt, grad, xr, vr, gr, span, what):
    e = np.abs(xt - xr).max(axis=1) / span
    ev = np.abs(vt - vr).max(axis=1)
    g = rel_l2(grad, gr)
    print(f"{what}: vs oracle/_ref ({ref.isa()}): endpoint max {e.max():.2e}, within 1e-5 {np.mean(e <= END_TOL):.5f}, "
          f"direction max {ev.max():.2e}, gradient rel L2 {g:.2e}")
    assert e.max() <= END_TOL and ev.max() <= 2 * END_TOL, (what, float(e.max()), float(ev.max()))
    assert np.array_equal(xt, xr) and np.array_equal(vt, vr), what  # same float32 operations in the same order
    assert g <= GRAD_TOL, (what, g)


GRID_CASES = [
    ("caustic256", dict(), "ones"),                      # the benchmarked batch: lattice-exact samples
    ("caustic256", dict(backoff=True), "normal"),        # z0 = -0.37 ds: no lattice excuse
    ("caustic256_cone", dict(backoff=True), "normal"),   # oblique batch
    ("fuel65", dict(), "normal"),                        # configs[2]: 32 views through 65^3
    ("luneburg32", dict(), "ones"),                      # configs[0]: all 65 536 rays
    ("fiber64", dict(), "normal"),                       # configs[1], grid form
]


@pytest.mark.gpu
@pytest.mark.parametrize("name,opt,seeds", GRID_CASES, ids=[f"{c[0]}{'-backoff' if c[1] else ''}" for c in GRID_CASES])
@pytest.mark.parametrize("flags", [0, 2], ids=["default", "deterministic"])
def test_gpu_against_reference_library(name, opt, seeds, flags, cuda_lib, oracle_mod):
    from adjointnonlinearraytracing_b200 import ops
    case = sc.config_case(name)
    ref = oracle_mod.ref_for_spacing(case["h"])
    if ref is None:
        pytest.skip("no oracle/_ref build with rcp(h) == 1/h for this spacing on this host")
    n = 65536 if name == "luneburg32" else N_SUB
    _, pos, vel, dx, dv = _sample(case, n, seeds)
    if opt.get("backoff"):
        pos = (pos - np.float32(0.37) * case["ds"] * vel).astype(np.float32)
    rif, h, ds = case["rif"], case["h"], case["ds"]
    xr, vr, gr = oracle_mod.trace_backtrace_mt(ref, rif, pos, vel, dx, dv, h, ds)
    st = ops.new_stats("cuda")
    xt, vt = ops.trace(T(rif), T(pos), T(vel), float(h), float(ds), stats=st)
    assert int(st[1]) == 0  # every ray exits
    grad = ops.backtrace(T(rif), xt, vt, T(dx), T(dv), float(h), float(ds), flags=flags)
    _check_against(ref, xt.cpu().numpy(), vt.cpu().numpy(), grad.cpu().numpy(), xr, vr, gr, 1.0,
                   f"{name}{opt or ''} flags={flags}")


@pytest.mark.gpu
def test_gpu_cable_against_reference_library(cuda_lib, oracle_mod):
    """configs[1], cable form (129-sample profile, 512^2 rays, ds = R/res/2): 4096 rays in chunks, each chunk
    one reference call and one GPU call (trace_cable's closest approach is batch-coupled)."""
    from adjointnonlinearraytracing_b200 import ops
    ref = oracle_mod.ref
    if ref is None:
        pytest.skip("oracle/_ref did not travel")
    case = sc.config_case("cable129")
    idx, pos, vel, dx, dv = _sample(case, 4096, "normal")
    tgt = np.ascontiguousarray(case["target"][idx])
    R, L, ds, prof = case["radius"], case["length"], case["ds"], case["rif"]
    bounds = oracle_mod.chunk_bounds(len(pos))

    def one(i):
        a, b = bounds[i], bounds[i + 1]
        xt, vt, d2 = ref.trace_cable(prof, R, L, pos[a:b], vel[a:b], tgt[a:b], ds)
        return xt, vt, d2, ref.backtrace_cable(prof, R, L, xt, vt, dx[a:b], dv[a:b], ds)
    parts = oracle_mod.run_threads(one, len(bounds) - 1)
    gsum = np.zeros(prof.shape, np.float64); gref = np.zeros(prof.shape, np.float64)
    emax = 0.0
    for i, (xr, vr, d2r, gr) in enumerate(parts):
        a, b = bounds[i], bounds[i + 1]
        xt, vt, d2 = ops.trace_cable(T(prof), R, L, T(pos[a:b]), T(vel[a:b]), T(tgt[a:b]), float(ds))
        g = ops.backtrace_cable(T(prof), R, L, xt, vt, T(dx[a:b]), T(dv[a:b]), float(ds))
        emax = max(emax, float(np.abs(xt.cpu().numpy() - xr).max()) / L, float(np.abs(vt.cpu().numpy() - vr).max()))
        gsum += g.cpu().numpy(); gref += gr
    ge = rel_l2(gsum, gref)
    print(f"cable129: endpoint/direction max {emax:.2e}, gradient rel L2 {ge:.2e}")
    # ~1300 steps per ray through IEEE div / sqrt formulas, which enoki's AVX-512 packets evaluate with
    # rcp14 / rsqrt14 + Newton: endpoints agree to 3e-7; the float32 gradient of the REFERENCE ITSELF is 4e-3 from
    # its float64 evaluation on this batch (oracle.port64), the restatement is 8e-4 from the reference library.
    assert emax <= 1e-5, emax
    assert ge <= 2e-3, ge
