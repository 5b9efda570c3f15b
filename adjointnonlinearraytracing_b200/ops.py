"""Tensor-level entry points: torch CUDA tensors in, torch CUDA tensors out.

Thin marshalling over the C ABI (``include/drrt_b200.h``): device pointers and the
current torch stream are handed to ``libdrrt_b200.so``; nothing is computed in
Python.  Names, argument order and meaning follow the reference's pybind11
methods (reference ``src/drrt.cpp:47-58``) minus the ``res`` triple, which is
taken from ``rif.shape`` unless given (``core/tracer.py:298-307`` passes
``rif.shape`` and ``rif.flatten()``).

Ray tensors are ``(N, 3)`` float32 (AoS, as in ``core/tracer.py``) or, with
``layout="soa"``, ``(3, N)``.
"""
from __future__ import annotations

from ctypes import c_void_p
from typing import Optional, Sequence, Tuple

import torch

from . import _capi
from ._capi import (DRRT_ACCUMULATE, DRRT_AOS, DRRT_DETERMINISTIC, DRRT_FAST_ARITH, DRRT_INTERLEAVE,
                    DRRT_NO_INTERLEAVE, DRRT_REFERENCE_ARITH, DRRT_SOA, DRRT_WARP_AGGREGATION)

__all__ = [
    "trace", "trace_pln", "trace_target", "trace_sdf", "trace_cable", "backtrace",
    "backtrace_sdf", "backtrace_cable", "backtrace_rays", "volume_eval", "volume_splat", "cylinder_eval",
    "cylinder_splat", "make_flags", "new_stats",
]


def make_flags(reference_arith: bool = False, deterministic: bool = False,
               accumulate: bool = False, fast_arith: bool = False,
               warp_aggregation: bool = False, interleave=None, face_carry=None,
               replicas=True) -> int:
    """Flag word of the C ABI (include/drrt_b200.h).  Default arithmetic = exact trajectory;
    `interleave`: None = the library default (off with the face carry), 0/False = off,
    "auto" = measured on the device, 2^k = forced; `face_carry`: None = chosen from h/ds, True / False = forced;
    `replicas`: True = private gradient copies on small
    grids (the library's choice), False = one grid, n in 2..15 = that many copies on any grid."""
    if reference_arith and fast_arith:
        raise ValueError("reference_arith and fast_arith are mutually exclusive")
    return ((DRRT_REFERENCE_ARITH if reference_arith else 0)
            | (DRRT_DETERMINISTIC if deterministic else 0)
            | (DRRT_ACCUMULATE if accumulate else 0)
            | (DRRT_FAST_ARITH if fast_arith else 0)
            | (DRRT_WARP_AGGREGATION if warp_aggregation else 0)
            | (0 if face_carry is None else _capi.DRRT_FACE_CARRY if face_carry else _capi.DRRT_NO_FACE_CARRY)
            | (0 if replicas is True else _capi.DRRT_NO_REPLICAS if not replicas else _capi.DRRT_REPLICAS(replicas))
            | (0 if interleave is None else _capi.DRRT_AUTO_INTERLEAVE if interleave == "auto"
               else DRRT_NO_INTERLEAVE if not interleave else DRRT_INTERLEAVE(interleave)))


def new_stats(device) -> torch.Tensor:
    """int64[2] device counter: [0] live ray-steps, [1] rays that never exited."""
    return torch.zeros(2, dtype=torch.int64, device=device)


def _ptr(t: Optional[torch.Tensor]):
    return c_void_p(t.data_ptr()) if t is not None else c_void_p(0)


def _stream(t: torch.Tensor):
    return c_void_p(torch.cuda.current_stream(t.device).cuda_stream)


def _dev(t: torch.Tensor, name: str) -> torch.Tensor:
    if not isinstance(t, torch.Tensor):
        raise TypeError(f"{name}: expected a torch.Tensor, got {type(t).__name__}")
    if not t.is_cuda:
        raise RuntimeError(f"{name}: expected a CUDA tensor (there is no CPU path; use TracerS for host buffers)")
    if t.dtype != torch.float32:
        raise RuntimeError(f"{name}: expected float32, got {t.dtype}")
    return t.detach().contiguous()


def _grid(rif: torch.Tensor, res: Optional[Sequence[int]], name="rif"):
    rif = _dev(rif, name)
    if res is None:
        if rif.dim() != 3:
            raise RuntimeError("res must be given when rif is not a 3-D tensor")
        res = tuple(rif.shape)
    W, H, D = (int(r) for r in res)
    if W * H * D != rif.numel():
        raise RuntimeError("Resolution doesn't match data")  # reference src/volume.cpp:37
    return rif, (W, H, D)


def _rays(t: torch.Tensor, name: str, layout: str, n: Optional[int] = None) -> Tuple[torch.Tensor, int]:
    t = _dev(t, name)
    if layout == "aos":
        if t.dim() != 2 or t.shape[1] != 3:
            raise RuntimeError(f"{name}: expected shape (N, 3), got {tuple(t.shape)}")
        cnt = t.shape[0]
    elif layout == "soa":
        if t.dim() != 2 or t.shape[0] != 3:
            raise RuntimeError(f"{name}: expected shape (3, N), got {tuple(t.shape)}")
        cnt = t.shape[1]
    else:
        raise ValueError("layout must be 'aos' or 'soa'")
    if n is not None and cnt != n:
        raise RuntimeError(f"{name}: ray count {cnt} does not match {n}")
    return t, cnt


def _lay(layout: str) -> int:
    return DRRT_AOS if layout == "aos" else DRRT_SOA


def _same_device(*ts):
    dev = None
    for t in ts:
        if t is None:
            continue
        if dev is None:
            dev = t.device
        elif t.device != dev:
            raise RuntimeError("all tensors must be on the same CUDA device")
    return dev


# --------------------------------------------------------------------------- forward
def trace(rif, pos, vel, h, ds, *, res=None, flags=0, stats=None, layout="aos"):
    """``TracerC.trace`` (reference src/drrt.cpp:51, src/tracer.cpp:35-100)."""
    rif, (W, H, D) = _grid(rif, res)
    pos, N = _rays(pos, "pos", layout)
    vel, _ = _rays(vel, "vel", layout, N)
    dev = _same_device(rif, pos, vel, stats)
    xt, vt = torch.empty_like(pos), torch.empty_like(pos)
    with torch.cuda.device(dev):
        _capi.check(_capi.load().drrt_trace(_ptr(rif), W, H, D, _ptr(pos), _ptr(vel), N, _lay(layout),
                                            float(h), float(ds), _ptr(xt), _ptr(vt), _ptr(stats),
                                            int(flags), _stream(rif)))
    return xt, vt


def trace_pln(rif, pos, vel, pln_o, pln_d, h, ds, *, res=None, flags=0, stats=None, layout="aos"):
    """``TracerC.trace_pln`` (src/drrt.cpp:52, src/tracer.cpp:102-172) -> (xt, vt, not_escaped)."""
    rif, (W, H, D) = _grid(rif, res)
    pos, N = _rays(pos, "pos", layout)
    vel, _ = _rays(vel, "vel", layout, N)
    po, _ = _rays(pln_o, "pln_o", layout, N)
    pd, _ = _rays(pln_d, "pln_d", layout, N)
    dev = _same_device(rif, pos, vel, po, pd, stats)
    xt, vt = torch.empty_like(pos), torch.empty_like(pos)
    mask = torch.empty(N, dtype=torch.uint8, device=dev)
    with torch.cuda.device(dev):
        _capi.check(_capi.load().drrt_trace_plane(_ptr(rif), W, H, D, _ptr(pos), _ptr(vel), _ptr(po),
                                                  _ptr(pd), N, _lay(layout), float(h), float(ds),
                                                  _ptr(xt), _ptr(vt), _ptr(mask), _ptr(stats),
                                                  int(flags), _stream(rif)))
    return xt, vt, mask.to(torch.bool)


def trace_target(rif, pos, vel, target, h, ds, *, res=None, flags=0, stats=None, layout="aos"):
    """``TracerC.trace_target`` (src/drrt.cpp:54, src/tracer.cpp:174-242) -> (xt, vt, dist2)."""
    rif, (W, H, D) = _grid(rif, res)
    pos, N = _rays(pos, "pos", layout)
    vel, _ = _rays(vel, "vel", layout, N)
    tgt, _ = _rays(target, "target", layout, N)
    dev = _same_device(rif, pos, vel, tgt, stats)
    xt, vt = torch.empty_like(pos), torch.empty_like(pos)
    d2 = torch.empty(N, dtype=torch.float32, device=dev)
    with torch.cuda.device(dev):
        _capi.check(_capi.load().drrt_trace_target(_ptr(rif), W, H, D, _ptr(pos), _ptr(vel), _ptr(tgt),
                                                   N, _lay(layout), float(h), float(ds), _ptr(xt),
                                                   _ptr(vt), _ptr(d2), _ptr(stats), int(flags),
                                                   _stream(rif)))
    return xt, vt, d2


def trace_sdf(rif, sdf, pos, vel, h, ds, *, res=None, flags=0, stats=None, layout="aos"):
    """``TracerC.trace_sdf`` (src/drrt.cpp:53, src/tracer.cpp:244-310)."""
    rif, (W, H, D) = _grid(rif, res)
    sdf, _ = _grid(sdf, (W, H, D), "sdf")
    pos, N = _rays(pos, "pos", layout)
    vel, _ = _rays(vel, "vel", layout, N)
    dev = _same_device(rif, sdf, pos, vel, stats)
    xt, vt = torch.empty_like(pos), torch.empty_like(pos)
    with torch.cuda.device(dev):
        _capi.check(_capi.load().drrt_trace_sdf(_ptr(rif), _ptr(sdf), W, H, D, _ptr(pos), _ptr(vel), N,
                                                _lay(layout), float(h), float(ds), _ptr(xt), _ptr(vt),
                                                _ptr(stats), int(flags), _stream(rif)))
    return xt, vt


def trace_cable(rif, radius, length, pos, vel, target, ds, *, flags=0, stats=None, layout="aos"):
    """``TracerC.trace_cable`` (src/drrt.cpp:55, src/tracer.cpp:312-382) -> (xt, vt, dist2)."""
    rif = _dev(rif, "rif").reshape(-1)
    pos, N = _rays(pos, "pos", layout)
    vel, _ = _rays(vel, "vel", layout, N)
    tgt, _ = _rays(target, "target", layout, N)
    dev = _same_device(rif, pos, vel, tgt, stats)
    xt, vt = torch.empty_like(pos), torch.empty_like(pos)
    d2 = torch.empty(N, dtype=torch.float32, device=dev)
    with torch.cuda.device(dev):
        _capi.check(_capi.load().drrt_trace_cable(_ptr(rif), rif.numel(), float(radius), float(length),
                                                  _ptr(pos), _ptr(vel), _ptr(tgt), N, _lay(layout),
                                                  float(ds), _ptr(xt), _ptr(vt), _ptr(d2), _ptr(stats),
                                                  int(flags), _stream(rif)))
    return xt, vt, d2


# --------------------------------------------------------------------------- adjoint
def _scratch(G: int, flags: int, dev):
    nbytes = _capi.load().drrt_backtrace_scratch_bytes(G, int(flags))
    return torch.empty(nbytes, dtype=torch.uint8, device=dev) if nbytes else None


def _nbytes(t) -> int:
    return 0 if t is None else int(t.numel() * t.element_size())


def _adj_out(out, rif, flags):
    if out is None:
        if flags & DRRT_ACCUMULATE:
            raise RuntimeError("DRRT_ACCUMULATE needs an existing `out` gradient grid")
        return torch.empty_like(rif)
    if out.shape != rif.shape or out.dtype != torch.float32 or out.device != rif.device or not out.is_contiguous():
        raise RuntimeError("out: must be a contiguous float32 tensor shaped like rif on the same device")
    return out


def backtrace(rif, xt, vt, dx, dv, h, ds, *, res=None, flags=0, stats=None, out=None, layout="aos"):
    """``TracerC.backtrace`` (src/drrt.cpp:56, src/tracer.cpp:384-440) -> dL/d rif, shaped like rif."""
    rif, (W, H, D) = _grid(rif, res)
    xt, N = _rays(xt, "xt", layout)
    vt, _ = _rays(vt, "vt", layout, N)
    dx, _ = _rays(dx, "dx", layout, N)
    dv, _ = _rays(dv, "dv", layout, N)
    dev = _same_device(rif, xt, vt, dx, dv, stats, out)
    grad = _adj_out(out, rif, flags)
    scratch = _scratch(rif.numel(), flags, dev)
    with torch.cuda.device(dev):
        _capi.check(_capi.load().drrt_backtrace(_ptr(rif), W, H, D, _ptr(xt), _ptr(vt), _ptr(dx), _ptr(dv),
                                                N, _lay(layout), float(h), float(ds), _ptr(grad),
                                                _ptr(scratch), _nbytes(scratch), _ptr(stats), int(flags), _stream(rif)))
    return grad


def backtrace_sdf(rif, sdf, xt, vt, dx, dv, h, ds, *, res=None, flags=0, stats=None, out=None,
                  layout="aos"):
    """``TracerC.backtrace_sdf`` (src/drrt.cpp:57, src/tracer.cpp:443-509)."""
    rif, (W, H, D) = _grid(rif, res)
    sdf, _ = _grid(sdf, (W, H, D), "sdf")
    xt, N = _rays(xt, "xt", layout)
    vt, _ = _rays(vt, "vt", layout, N)
    dx, _ = _rays(dx, "dx", layout, N)
    dv, _ = _rays(dv, "dv", layout, N)
    dev = _same_device(rif, sdf, xt, vt, dx, dv, stats, out)
    grad = _adj_out(out, rif, flags)
    scratch = _scratch(rif.numel(), flags, dev)
    with torch.cuda.device(dev):
        _capi.check(_capi.load().drrt_backtrace_sdf(_ptr(rif), _ptr(sdf), W, H, D, _ptr(xt), _ptr(vt),
                                                    _ptr(dx), _ptr(dv), N, _lay(layout), float(h),
                                                    float(ds), _ptr(grad), _ptr(scratch), _nbytes(scratch), _ptr(stats),
                                                    int(flags), _stream(rif)))
    return grad


def backtrace_rays(rif, xt, vt, dx, dv, h, ds, *, pos0=None, sdf=None, res=None, flags=0, stats=None,
                   out=None, layout="aos"):
    """Adjoint pass that also returns the gradient w.r.t. the launch state, the contract of
    ``ADTracerC.backward`` / ``ADSDFTracerC.backward`` (reference core/tracer.py:53-60):
    -> (dL/d rif, dL/d pos, dL/d vel).  ``pos0`` = launch positions (recovers the number of
    straight steps flown before the volume; omit for rays launched on the volume)."""
    rif, (W, H, D) = _grid(rif, res)
    if sdf is not None:
        sdf, _ = _grid(sdf, (W, H, D), "sdf")
    xt, N = _rays(xt, "xt", layout)
    vt, _ = _rays(vt, "vt", layout, N)
    dx, _ = _rays(dx, "dx", layout, N)
    dv, _ = _rays(dv, "dv", layout, N)
    if pos0 is not None:
        pos0, _ = _rays(pos0, "pos0", layout, N)
    dev = _same_device(rif, sdf, xt, vt, dx, dv, pos0, stats, out)
    grad = _adj_out(out, rif, flags)
    scratch = _scratch(rif.numel(), flags, dev)
    dpos, dvel = torch.empty_like(xt), torch.empty_like(xt)
    with torch.cuda.device(dev):
        _capi.check(_capi.load().drrt_backtrace_rays(_ptr(rif), _ptr(sdf), W, H, D, _ptr(xt), _ptr(vt),
                                                     _ptr(dx), _ptr(dv), _ptr(pos0), N, _lay(layout),
                                                     float(h), float(ds), _ptr(grad), _ptr(dpos),
                                                     _ptr(dvel), _ptr(scratch), _nbytes(scratch), _ptr(stats), int(flags),
                                                     _stream(rif)))
    return grad, dpos, dvel


def backtrace_cable(rif, radius, length, xt, vt, dx, dv, ds, *, flags=0, stats=None, out=None,
                    layout="aos"):
    """``TracerC.backtrace_cable`` (src/drrt.cpp:58, src/tracer.cpp:511-567) -> dL/d profile."""
    rif = _dev(rif, "rif").reshape(-1)
    xt, N = _rays(xt, "xt", layout)
    vt, _ = _rays(vt, "vt", layout, N)
    dx, _ = _rays(dx, "dx", layout, N)
    dv, _ = _rays(dv, "dv", layout, N)
    dev = _same_device(rif, xt, vt, dx, dv, stats, out)
    grad = _adj_out(out, rif, flags)
    scratch = _scratch(rif.numel(), flags, dev)
    with torch.cuda.device(dev):
        _capi.check(_capi.load().drrt_backtrace_cable(_ptr(rif), rif.numel(), float(radius), float(length),
                                                      _ptr(xt), _ptr(vt), _ptr(dx), _ptr(dv), N,
                                                      _lay(layout), float(ds), _ptr(grad), _ptr(scratch), _nbytes(scratch),
                                                      _ptr(stats), int(flags), _stream(rif)))
    return grad


# --------------------------------------------------------------------------- probes
def volume_eval(rif, pos, h, *, res=None, flags=0):
    """volume::eval_grad + eval_hess at points -> (n[N], grad[N,3], mixed partials[N,3])."""
    rif, (W, H, D) = _grid(rif, res)
    pos, N = _rays(pos, "pos", "aos")
    n = torch.empty(N, dtype=torch.float32, device=rif.device)
    g, hs = torch.empty_like(pos), torch.empty_like(pos)
    with torch.cuda.device(rif.device):
        _capi.check(_capi.load().drrt_volume_eval(_ptr(rif), W, H, D, _ptr(pos), N, DRRT_AOS, float(h),
                                                  _ptr(n), _ptr(g), _ptr(hs), int(flags), _stream(rif)))
    return n, g, hs


def volume_splat(data, pos, val, grad, h, *, res=None):
    """volume::splat into a copy of ``data``."""
    data, (W, H, D) = _grid(data, res, "data")
    data = data.clone()
    pos, N = _rays(pos, "pos", "aos")
    grad, _ = _rays(grad, "grad", "aos", N)
    val = _dev(val, "val").reshape(-1)
    with torch.cuda.device(data.device):
        _capi.check(_capi.load().drrt_volume_splat(_ptr(data), W, H, D, _ptr(pos), _ptr(val), _ptr(grad),
                                                   N, DRRT_AOS, float(h), _stream(data)))
    return data


def cylinder_eval(rif, radius, length, pos):
    rif = _dev(rif, "rif").reshape(-1)
    pos, N = _rays(pos, "pos", "aos")
    n = torch.empty(N, dtype=torch.float32, device=rif.device)
    g = torch.empty_like(pos)
    hs = torch.empty(N, 3, 3, dtype=torch.float32, device=rif.device)
    with torch.cuda.device(rif.device):
        _capi.check(_capi.load().drrt_cylinder_eval(_ptr(rif), rif.numel(), float(radius), float(length),
                                                    _ptr(pos), N, DRRT_AOS, _ptr(n), _ptr(g), _ptr(hs),
                                                    _stream(rif)))
    return n, g, hs


def cylinder_splat(data, radius, length, pos, val, grad):
    data = _dev(data, "data").reshape(-1).clone()
    pos, N = _rays(pos, "pos", "aos")
    grad, _ = _rays(grad, "grad", "aos", N)
    val = _dev(val, "val").reshape(-1)
    with torch.cuda.device(data.device):
        _capi.check(_capi.load().drrt_cylinder_splat(_ptr(data), data.numel(), float(radius), float(length),
                                                     _ptr(pos), _ptr(val), _ptr(grad), N, DRRT_AOS,
                                                     _stream(data)))
    return data


# --------------------------------------------------------------------------- optimiser step
def adam_step_(n, grad, exp_avg, exp_avg_sq, step, lr, betas=(0.9, 0.999), eps=1e-8, clamp_min=1.0,
               zero_boundary=True, clamp=True):
    """One fused pass = ``n.grad[boundary] = 0; torch.optim.Adam.step(); n.clamp_(min=clamp_min)``
    (reference core/optimizer.py:61-68).  Updates ``n``, ``exp_avg``, ``exp_avg_sq`` in place."""
    for name, t in (("n", n), ("grad", grad), ("exp_avg", exp_avg), ("exp_avg_sq", exp_avg_sq)):
        if not (isinstance(t, torch.Tensor) and t.is_cuda and t.dtype == torch.float32 and t.is_contiguous()):
            raise RuntimeError(f"{name}: expected a contiguous float32 CUDA tensor")
        if t.shape != n.shape:
            raise RuntimeError("Resolution doesn't match data")
    if n.dim() != 3:
        raise RuntimeError("adam_step_: expected a 3-D grid")
    S0, S1, S2 = (int(s) for s in n.shape)
    flags = (0 if zero_boundary else 1) | (0 if clamp else 2)
    with torch.cuda.device(n.device):
        _capi.check(_capi.load().drrt_adam_step(_ptr(n), _ptr(grad), _ptr(exp_avg), _ptr(exp_avg_sq), S0, S1, S2,
                                                int(step), float(lr), float(betas[0]), float(betas[1]), float(eps),
                                                float(clamp_min), flags, _stream(n)))
    return n


def adam_step_slab_(n_slab, grad_slab, exp_avg_slab, exp_avg_sq_slab, shape, i0_begin, step, lr, betas=(0.9, 0.999),
                    eps=1e-8, clamp_min=1.0, zero_boundary=True, clamp=True):
    """``adam_step_`` on planes ``[i0_begin, i0_begin + n_slab.shape[0])`` of a grid of ``shape`` only (one rank's
    share of the sharded optimiser step, ``dist.ShardedAdam``).  The four tensors are (views of) that slab; the
    zeroed boundary is that of the whole grid."""
    S0, S1, S2 = (int(s) for s in shape)
    for name, t in (("n", n_slab), ("grad", grad_slab), ("exp_avg", exp_avg_slab), ("exp_avg_sq", exp_avg_sq_slab)):
        if not (isinstance(t, torch.Tensor) and t.is_cuda and t.dtype == torch.float32 and t.is_contiguous()):
            raise RuntimeError(f"{name}: expected a contiguous float32 CUDA tensor")
        if t.dim() != 3 or tuple(t.shape[1:]) != (S1, S2) or t.shape[0] != n_slab.shape[0]:
            raise RuntimeError("Resolution doesn't match data")
    flags = (0 if zero_boundary else 1) | (0 if clamp else 2)
    with torch.cuda.device(n_slab.device):
        _capi.check(_capi.load().drrt_adam_step_slab(_ptr(n_slab), _ptr(grad_slab), _ptr(exp_avg_slab),
                                                     _ptr(exp_avg_sq_slab), S0, S1, S2, int(i0_begin),
                                                     int(n_slab.shape[0]), int(step), float(lr), float(betas[0]),
                                                     float(betas[1]), float(eps), float(clamp_min), flags,
                                                     _stream(n_slab)))
    return n_slab


def upres_volume(n, new_res):
    """``grid.upres_volume`` / ``optimizer.upres_scene`` (reference core/grid.py:318-330,
    core/optimizer.py:7-10): trilinear resample of a cubic grid to ``new_res``."""
    src = _dev(n, "n")
    if src.dim() != 3:
        raise RuntimeError("upres_volume: expected a 3-D grid")
    d = [int(r) for r in new_res]
    dst = torch.empty(d, dtype=torch.float32, device=src.device)
    with torch.cuda.device(src.device):
        _capi.check(_capi.load().drrt_upres_volume(_ptr(src), *(int(s) for s in src.shape), _ptr(dst), *d, _stream(src)))
    return dst
