"""dL/dpos, dL/dvel of the launch state from the adjoint's final co-state
(``drrt_backtrace_rays``; contract of the reference's ``ADTracerC.backward``,
core/tracer.py:53-60).

Chain of evidence:
  1. CPU: the C restatement's ``backtrace_rays`` (float64 build) against a float64 TAPE
     differentiation of the forward loop (oracle/ad_oracle.py = what enoki's tape does
     for ADTracerC): <= 1e-10 absolute -- pins the co-state -> launch-state formula.
  2. CPU: float32 restatement against the float64 one: <= 2e-3 relative (float32 paths).
  3. GPU: the CUDA kernel against the float32 restatement: <= 1e-5 relative (same
     trajectory bit for bit; lambda/mu in factored FMA form), and ``ADTracerC`` end to
     end through torch autograd against the tape: <= 2e-3 relative.
Rays are launched outside the volume (backoff) or just inside its first cell layer --
the cases for which the adjoint's stopping rule sees the whole forward path.
"""
import numpy as np
import pytest
import torch

from conftest import rel_max
from adjointnonlinearraytracing_b200 import scenes as sc


def _case(ng, s, backoff_steps, seed=3, sdf=False):
    h = float(sc.grid_spacing(ng)); ds = h / 2
    rif = sc.field_bump(ng)
    pos, vel = sc.rays_cone(s, spread=0.1, backoff=backoff_steps * ds)
    if backoff_steps == 0:
        pos[:, 2] = np.float32(0.3 * ds)
    rng = np.random.default_rng(seed)
    wx = rng.standard_normal(pos.shape).astype(np.float32)
    wv = rng.standard_normal(pos.shape).astype(np.float32)
    return rif, pos, vel, h, ds, wx, wv


@pytest.mark.parametrize("backoff", [0, 3.37])
def test_costate_formula_against_tape_f64(backoff, oracle_mod):
    from oracle import ad_oracle
    rif, pos, vel, h, ds, wx, wv = _case(9, 6, backoff)
    gr, gp, gv, xt, vt = ad_oracle.gradients(rif, pos, vel, h, ds, wx, wv)
    xo, vo = oracle_mod.port64.trace(rif, pos, vel, h, ds)
    assert np.abs(xo - xt).max() <= 1e-12 and np.abs(vo - vt).max() <= 1e-12
    g, dp, dv = oracle_mod.port64.backtrace_rays(rif, xo, vo, wx, wv, h, ds, pos0=pos)
    assert np.abs(dp - gp).max() <= 1e-10, float(np.abs(dp - gp).max())
    assert np.abs(dv - gv).max() <= 1e-10, float(np.abs(dv - gv).max())
    # the grid gradient of the same call is the plain backtrace
    g0 = oracle_mod.port64.backtrace(rif, xo, vo, wx, wv, h, ds)
    assert np.array_equal(g, g0)


def test_costate_f32_against_f64(oracle_mod):
    rif, pos, vel, h, ds, wx, wv = _case(17, 12, 2.6)
    x64, v64 = oracle_mod.port64.trace(rif, pos, vel, h, ds)
    _, dp64, dv64 = oracle_mod.port64.backtrace_rays(rif, x64, v64, wx, wv, h, ds, pos0=pos)
    x32, v32 = oracle_mod.port.trace(rif, pos, vel, np.float32(h), np.float32(ds))
    _, dp32, dv32 = oracle_mod.port.backtrace_rays(rif, x32, v32, wx, wv, np.float32(h), np.float32(ds), pos0=pos)
    assert rel_max(dp32, dp64) <= 2e-3 and rel_max(dv32, dv64) <= 2e-3


def test_no_pos0_means_first_sample_launch(oracle_mod):
    """pos0 omitted: k = 0, dvel = mu - ds*lambda (rays launched on their first in-bounds sample)."""
    rif, pos, vel, h, ds, wx, wv = _case(9, 5, 0)
    xo, vo = oracle_mod.port64.trace(rif, pos, vel, h, ds)
    _, dp_a, dv_a = oracle_mod.port64.backtrace_rays(rif, xo, vo, wx, wv, h, ds, pos0=pos)
    _, dp_b, dv_b = oracle_mod.port64.backtrace_rays(rif, xo, vo, wx, wv, h, ds)
    assert np.allclose(dp_a, dp_b, atol=1e-14) and np.allclose(dv_a, dv_b, atol=1e-14)


# ------------------------------------------------------------------ GPU
def T(a):
    return torch.from_numpy(np.ascontiguousarray(a)).cuda()


@pytest.mark.gpu
@pytest.mark.parametrize("backoff,flags", [(0, 0), (3.37, 0), (3.37, 1), (3.37, 2), (1.2, 8)])
def test_gpu_backtrace_rays_vs_restatement(backoff, flags, cuda_lib, oracle_mod):
    from adjointnonlinearraytracing_b200 import ops
    rif, pos, vel, h, ds, wx, wv = _case(17, 40, backoff)
    xt, vt = ops.trace(T(rif), T(pos), T(vel), h, ds, flags=flags & ~2)
    g, dp, dv = ops.backtrace_rays(T(rif), xt, vt, T(wx), T(wv), h, ds, pos0=T(pos), flags=flags)
    g0 = ops.backtrace(T(rif), xt, vt, T(wx), T(wv), h, ds, flags=flags)
    assert rel_max(g.cpu().numpy(), g0.cpu().numpy()) <= 2e-6
    xo, vo = xt.cpu().numpy(), vt.cpu().numpy()
    _, dpo, dvo = oracle_mod.port.backtrace_rays(rif, xo, vo, wx, wv, np.float32(h), np.float32(ds), pos0=pos)
    tol = 1e-5 if not (flags & 8) else 1e-3
    assert rel_max(dp.cpu().numpy(), dpo) <= tol, rel_max(dp.cpu().numpy(), dpo)
    assert rel_max(dv.cpu().numpy(), dvo) <= tol, rel_max(dv.cpu().numpy(), dvo)


@pytest.mark.gpu
def test_gpu_backtrace_rays_sdf(cuda_lib, oracle_mod):
    from adjointnonlinearraytracing_b200 import ops
    rif, pos, vel, h, ds, wx, wv = _case(17, 24, 2.2)
    sdf = sc.sdf_sphere(17)
    xt, vt = ops.trace_sdf(T(rif), T(sdf), T(pos), T(vel), h, ds)
    g, dp, dv = ops.backtrace_rays(T(rif), xt, vt, T(wx), T(wv), h, ds, pos0=T(pos), sdf=T(sdf))
    g0 = ops.backtrace_sdf(T(rif), T(sdf), xt, vt, T(wx), T(wv), h, ds)
    assert rel_max(g.cpu().numpy(), g0.cpu().numpy()) <= 2e-6
    _, dpo, dvo = oracle_mod.port.backtrace_rays(rif, xt.cpu().numpy(), vt.cpu().numpy(), wx, wv,
                                                 np.float32(h), np.float32(ds), pos0=pos, sdf=sdf)
    assert rel_max(dp.cpu().numpy(), dpo) <= 1e-5 and rel_max(dv.cpu().numpy(), dvo) <= 1e-5


@pytest.mark.gpu
def test_gpu_adtracer_autograd_vs_tape(cuda_lib, oracle_mod):
    """ADTracerC through torch autograd: gradients w.r.t. x and v against the float64 tape."""
    from oracle import ad_oracle
    from adjointnonlinearraytracing_b200.tracer import ADTracerC
    rif, pos, vel, h, ds, wx, wv = _case(9, 8, 2.37)
    gr, gp, gv, _, _ = ad_oracle.gradients(rif, pos, vel, h, ds, wx, wv)
    n = T(rif).requires_grad_(True)
    x = T(pos).requires_grad_(True)
    v = T(vel).requires_grad_(True)
    xt, vt = ADTracerC.apply(n, x, v, h, ds)
    ((xt * T(wx)).sum() + (vt * T(wv)).sum()).backward()
    assert n.grad is not None and n.grad.shape == n.shape
    assert rel_max(x.grad.cpu().numpy(), gp) <= 2e-3, rel_max(x.grad.cpu().numpy(), gp)
    assert rel_max(v.grad.cpu().numpy(), gv) <= 2e-3, rel_max(v.grad.cpu().numpy(), gv)
    # x, v not requiring grad: the plain adjoint path, no ray gradients
    n2 = T(rif).requires_grad_(True)
    xt2, vt2 = ADTracerC.apply(n2, T(pos), T(vel), h, ds)
    ((xt2 * T(wx)).sum() + (vt2 * T(wv)).sum()).backward()
    assert rel_max(n2.grad.cpu().numpy(), n.grad.cpu().numpy()) <= 2e-6
