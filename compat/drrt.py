"""``import drrt`` for an unmodified reference ``core/tracer.py``: the classes
of the reference's pybind11 module (reference src/drrt.cpp:21-60) over ``libdrrt_b200.so``, taking and
returning the enoki stand-in arrays of ``compat/enoki``.  Put ``compat/`` on ``sys.path`` ahead of
the reference's own build directory.
"""
from __future__ import annotations

import os
import sys

_ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if _ROOT not in sys.path:
    sys.path.insert(0, _ROOT)

import enoki  # noqa: E402  (the stand-in next to this file)
from enoki import _tape  # noqa: E402
from enoki.cuda import Float32 as FloatC, Vector3f as Vector3fC  # noqa: E402
from enoki.cuda_autodiff import Float32 as FloatD, Vector3f as Vector3fD  # noqa: E402
from enoki.dynamic import Float32 as FloatS, Vector3f as Vector3fS  # noqa: E402
from adjointnonlinearraytracing_b200 import drrt as _impl  # noqa: E402


def _res(res):
    return tuple(int(r) for r in res)


class TracerC:
    """drrt.TracerC, reference src/drrt.cpp:47-58"""

    def __init__(self):
        self._t = _impl.TracerC()

    def trace(self, rif, res, pos, vel, h, ds):
        xt, vt = self._t.trace(rif, res, pos, vel, h, ds)
        return Vector3fC(xt), Vector3fC(vt)

    def trace_pln(self, rif, res, pos, vel, pln_o, pln_d, h, ds):
        xt, vt, m = self._t.trace_pln(rif, res, pos, vel, pln_o, pln_d, h, ds)
        return Vector3fC(xt), Vector3fC(vt), FloatC(m.to(xt.dtype))

    def trace_target(self, rif, res, pos, vel, target, h, ds):
        xt, vt, d2 = self._t.trace_target(rif, res, pos, vel, target, h, ds)
        return Vector3fC(xt), Vector3fC(vt), FloatC(d2)

    def trace_sdf(self, rif, sdf, res, pos, vel, h, ds):
        xt, vt = self._t.trace_sdf(rif, sdf, res, pos, vel, h, ds)
        return Vector3fC(xt), Vector3fC(vt)

    def trace_cable(self, rif, radius, length, pos, vel, target, ds):
        xt, vt, d2 = self._t.trace_cable(rif, radius, length, pos, vel, target, ds)
        return Vector3fC(xt), Vector3fC(vt), FloatC(d2)

    def backtrace(self, rif, res, xt, vt, dx, dv, h, ds):
        return FloatC(self._t.backtrace(rif, res, xt, vt, dx, dv, h, ds))

    def backtrace_sdf(self, rif, sdf, res, xt, vt, dx, dv, h, ds):
        return FloatC(self._t.backtrace_sdf(rif, sdf, res, xt, vt, dx, dv, h, ds))

    def backtrace_cable(self, rif, radius, length, xt, vt, dx, dv, ds):
        return FloatC(self._t.backtrace_cable(rif, radius, length, xt, vt, dx, dv, ds))


class TracerD:
    """drrt.TracerD, reference src/drrt.cpp:28-36: the forward traces on autodiff arrays.  Each call is
    recorded; ``FloatD.backward()`` differentiates it with the adjoint kernels."""

    def __init__(self):
        self._t = _impl.TracerC()

    def trace(self, rif, res, pos, vel, h, ds):
        xt, vt = self._t.trace(rif, res, pos, vel, h, ds)
        out = (Vector3fD(xt), Vector3fD(vt))
        _tape.record("plain", {"rif": rif, "x": pos, "v": vel}, out, {"res": _res(res), "h": h, "ds": ds})
        return out

    def trace_pln(self, rif, res, pos, vel, pln_o, pln_d, h, ds):
        xt, vt, m = self._t.trace_pln(rif, res, pos, vel, pln_o, pln_d, h, ds)
        out = (Vector3fD(xt), Vector3fD(vt), FloatD(m.to(xt.dtype)))
        _tape.record("plane", {"rif": rif, "x": pos, "v": vel}, out, {"res": _res(res), "h": h, "ds": ds})
        return out

    def trace_target(self, rif, res, pos, vel, target, h, ds):
        xt, vt, d2 = self._t.trace_target(rif, res, pos, vel, target, h, ds)
        out = (Vector3fD(xt), Vector3fD(vt), FloatD(d2))
        _tape.record("target", {"rif": rif, "x": pos, "v": vel}, out, {"res": _res(res), "h": h, "ds": ds})
        return out

    def trace_sdf(self, rif, sdf, res, pos, vel, h, ds):
        xt, vt = self._t.trace_sdf(rif, sdf, res, pos, vel, h, ds)
        out = (Vector3fD(xt), Vector3fD(vt))
        _tape.record("sdf", {"rif": rif, "sdf": sdf, "x": pos, "v": vel}, out, {"res": _res(res), "h": h, "ds": ds})
        return out

    def trace_cable(self, rif, radius, length, pos, vel, target, ds):
        xt, vt, d2 = self._t.trace_cable(rif, radius, length, pos, vel, target, ds)
        out = (Vector3fD(xt), Vector3fD(vt), FloatD(d2))
        _tape.record("cable", {"rif": rif, "x": pos, "v": vel}, out, {"radius": radius, "length": length, "ds": ds})
        return out


class TracerS:
    """drrt.TracerS, reference src/drrt.cpp:38-45: host arrays (through the `_host` entry points)."""

    def __init__(self):
        self._t = _impl.TracerS()

    def trace(self, rif, res, pos, vel, h, ds):
        xt, vt = self._t.trace(rif, res, pos, vel, h, ds)
        return Vector3fS(xt), Vector3fS(vt)

    def trace_sdf(self, rif, sdf, res, pos, vel, h, ds):
        xt, vt = self._t.trace_sdf(rif, sdf, res, pos, vel, h, ds)
        return Vector3fS(xt), Vector3fS(vt)

    def trace_target(self, rif, res, pos, vel, target, h, ds):
        xt, vt, d2 = self._t.trace_target(rif, res, pos, vel, target, h, ds)
        return Vector3fS(xt), Vector3fS(vt), FloatS(d2)

    def backtrace(self, rif, res, xt, vt, dx, dv, h, ds):
        return FloatS(self._t.backtrace(rif, res, xt, vt, dx, dv, h, ds))
