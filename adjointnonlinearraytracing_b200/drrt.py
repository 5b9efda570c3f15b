"""Python mirror of the reference's pybind11 module ``drrt`` (reference
``src/drrt.cpp:21-60``): classes ``TracerC`` (GPU), ``TracerD`` (GPU, differentiable)
and ``TracerS`` (host buffers) with the reference's method names and positional
argument order.  Arguments are torch tensors (or anything with a ``.torch()``
method, which is how enoki arrays convert): ``rif`` flat or 3-D, rays ``(N, 3)``.

``import adjointnonlinearraytracing_b200.drrt as drrt`` is the drop-in for
``import drrt`` in ``core/tracer.py``.
"""
from __future__ import annotations

import ctypes
from ctypes import c_void_p

import numpy as np
import torch

from . import _capi, ops


def _t(x):
    return x.torch() if hasattr(x, "torch") and not isinstance(x, torch.Tensor) else x


def _res(res):
    return tuple(int(r) for r in res)


class TracerC:
    """``drrt.TracerC`` (reference src/drrt.cpp:47-58): GPU, no autodiff."""

    flags = 0

    def trace(self, rif, res, pos, vel, h, ds):
        return ops.trace(_t(rif), _t(pos), _t(vel), h, ds, res=_res(res), flags=self.flags)

    def trace_pln(self, rif, res, pos, vel, pln_o, pln_d, h, ds):
        return ops.trace_pln(_t(rif), _t(pos), _t(vel), _t(pln_o), _t(pln_d), h, ds, res=_res(res),
                             flags=self.flags)

    def trace_target(self, rif, res, pos, vel, target, h, ds):
        return ops.trace_target(_t(rif), _t(pos), _t(vel), _t(target), h, ds, res=_res(res), flags=self.flags)

    def trace_sdf(self, rif, sdf, res, pos, vel, h, ds):
        return ops.trace_sdf(_t(rif), _t(sdf), _t(pos), _t(vel), h, ds, res=_res(res), flags=self.flags)

    def trace_cable(self, rif, radius, length, pos, vel, target, ds):
        return ops.trace_cable(_t(rif), radius, length, _t(pos), _t(vel), _t(target), ds, flags=self.flags)

    def backtrace(self, rif, res, xt, vt, dx, dv, h, ds):
        """Returns the flat gradient grid like the reference (FloatC of W*H*D)."""
        return ops.backtrace(_t(rif), _t(xt), _t(vt), _t(dx), _t(dv), h, ds, res=_res(res),
                             flags=self.flags).reshape(-1)

    def backtrace_sdf(self, rif, sdf, res, xt, vt, dx, dv, h, ds):
        return ops.backtrace_sdf(_t(rif), _t(sdf), _t(xt), _t(vt), _t(dx), _t(dv), h, ds, res=_res(res),
                                 flags=self.flags).reshape(-1)

    def backtrace_cable(self, rif, radius, length, xt, vt, dx, dv, ds):
        return ops.backtrace_cable(_t(rif), radius, length, _t(xt), _t(vt), _t(dx), _t(dv), ds,
                                   flags=self.flags)


class TracerD:
    """``drrt.TracerD`` (reference src/drrt.cpp:28-36): the five ``trace*`` methods with
    differentiable outputs.  The reference records an enoki tape; here the outputs are
    attached to torch autograd through the adjoint kernels (gradient w.r.t. ``rif``)."""

    def trace(self, rif, res, pos, vel, h, ds):
        from .tracer import BackTracerC
        rif = _t(rif)
        return BackTracerC.apply(rif.reshape(_res(res)), _t(pos), _t(vel), h, ds)

    def trace_pln(self, rif, res, pos, vel, pln_o, pln_d, h, ds):
        from .tracer import BackPlaneTracerC
        return BackPlaneTracerC.apply(_t(rif).reshape(_res(res)), _t(pos), _t(vel), _t(pln_o), _t(pln_d), h, ds)

    def trace_target(self, rif, res, pos, vel, target, h, ds):
        from .tracer import BackTargetTracerC
        return BackTargetTracerC.apply(_t(rif).reshape(_res(res)), _t(pos), _t(vel), _t(target), h, ds)

    def trace_sdf(self, rif, sdf, res, pos, vel, h, ds):
        from .tracer import BackSDFTracerC
        return BackSDFTracerC.apply(_t(rif).reshape(_res(res)), _t(sdf).reshape(_res(res)), _t(pos), _t(vel), h, ds)

    def trace_cable(self, rif, radius, length, pos, vel, target, ds):
        from .tracer import BackCableTracerC
        return BackCableTracerC.apply(_t(rif), radius, length, _t(pos), _t(vel), _t(target), ds)


class TracerS:
    """``drrt.TracerS`` (reference src/drrt.cpp:38-45): the host-array flavour.  The
    reference runs it on the CPU; here host buffers go through the ``_host`` entry
    points of the C ABI (copy in, run the CUDA kernels, copy out) -- there is no CPU
    compute path.  Accepts CPU torch tensors or numpy arrays; ``trace_sdf`` and
    ``trace_target`` stage through device tensors."""

    flags = 0

    @staticmethod
    def _np(a, shape=None):
        a = _t(a)
        if isinstance(a, torch.Tensor):
            a = a.detach().cpu().numpy()
        a = np.ascontiguousarray(a, dtype=np.float32)
        return a if shape is None else a.reshape(shape)

    @staticmethod
    def _p(a):
        return c_void_p(a.ctypes.data)

    def trace(self, rif, res, pos, vel, h, ds):
        W, H, D = _res(res)
        rif = self._np(rif).reshape(-1)
        if rif.size != W * H * D:
            raise RuntimeError("Resolution doesn't match data")
        pos, vel = self._np(pos), self._np(vel)
        xt, vt = np.empty_like(pos), np.empty_like(pos)
        stats = (ctypes.c_int64 * 2)()
        _capi.check(_capi.load().drrt_trace_host(self._p(rif), W, H, D, self._p(pos), self._p(vel), pos.shape[0],
                                                 _capi.DRRT_AOS, float(h), float(ds), self._p(xt), self._p(vt),
                                                 ctypes.cast(stats, c_void_p), self.flags))
        self.last_stats = (int(stats[0]), int(stats[1]))
        return torch.from_numpy(xt), torch.from_numpy(vt)

    def backtrace(self, rif, res, xt, vt, dx, dv, h, ds):
        W, H, D = _res(res)
        rif = self._np(rif).reshape(-1)
        if rif.size != W * H * D:
            raise RuntimeError("Resolution doesn't match data")
        xt, vt, dx, dv = self._np(xt), self._np(vt), self._np(dx), self._np(dv)
        grad = np.empty_like(rif)
        stats = (ctypes.c_int64 * 2)()
        _capi.check(_capi.load().drrt_backtrace_host(self._p(rif), W, H, D, self._p(xt), self._p(vt), self._p(dx),
                                                     self._p(dv), xt.shape[0], _capi.DRRT_AOS, float(h),
                                                     float(ds), self._p(grad), ctypes.cast(stats, c_void_p),
                                                     self.flags))
        self.last_stats = (int(stats[0]), int(stats[1]))
        return torch.from_numpy(grad)

    def _dev(self, a):
        return torch.from_numpy(self._np(a)).cuda()

    def trace_sdf(self, rif, sdf, res, pos, vel, h, ds):
        xt, vt = ops.trace_sdf(self._dev(rif), self._dev(sdf), self._dev(pos), self._dev(vel), h, ds,
                               res=_res(res), flags=self.flags)
        return xt.cpu(), vt.cpu()

    def trace_target(self, rif, res, pos, vel, target, h, ds):
        xt, vt, d2 = ops.trace_target(self._dev(rif), self._dev(pos), self._dev(vel), self._dev(target), h, ds,
                                      res=_res(res), flags=self.flags)
        return xt.cpu(), vt.cpu(), d2.cpu()


# functional aliases
trace = ops.trace
trace_pln = ops.trace_pln
trace_target = ops.trace_target
trace_sdf = ops.trace_sdf
trace_cable = ops.trace_cable
backtrace = ops.backtrace
backtrace_sdf = ops.backtrace_sdf
backtrace_cable = ops.backtrace_cable
