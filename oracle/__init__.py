"""TEST INFRASTRUCTURE ONLY -- CPU oracles for the eikonal trace/backtrace path.

Three checkers, all taking and returning host numpy arrays (ray vectors AoS
``(N,3)``, grids flat or 3-D C-contiguous, index ``(z*H+y)*W+x`` with
``(W,H,D) = res`` as in ``/root/reference/src/volume.cpp:134``):

``port``    the plain-C float32 restatement ``oracle/eikonal_oracle.c``
``port64``  the same source built with ``real = double`` (arbitration only)
``ref``     the reference's own ``Tracer<false,false>`` compiled from
            ``/root/reference`` by ``oracle/build_ref.sh`` (``oracle/_ref/``);
            ``None`` if that library has not been built.

``refad``   the reference's own CPU TAPE-AUTODIFF tracer ``Tracer<true,false>`` on enoki's
            ``DiffArray`` (``oracle/ref_ad_shim.cpp``, ``oracle/_ref/libdrrt_refad_v3.so``):
            what the reference's ``ADTracerC`` / ``ADSDFTracerC`` backward returns
            (``core/tracer.py:45-66``); ``None`` if not built.

Only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s CPU-baseline /
``--impl reference`` legs may import this package.  The product package
``adjointnonlinearraytracing_b200`` never does.
"""
from __future__ import annotations

import ctypes
import os
import subprocess
from ctypes import POINTER, c_float, c_double, c_int, c_int64, c_uint8, c_char_p

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_i64 = c_int64


def build(verbose: bool = False) -> None:
    """Compile the C restatement (and the reference library when its sources exist)."""
    out = subprocess.run(["make", "-C", _HERE], capture_output=True, text=True)
    if out.returncode != 0:
        raise RuntimeError("oracle build failed:\n" + out.stdout + out.stderr)
    if verbose:
        print(out.stdout)


def _cpu_has_avx512() -> bool:
    try:
        with open("/proc/cpuinfo") as f:
            for line in f:
                if line.startswith("flags"):
                    fl = set(line.split(":", 1)[1].split())
                    return {"avx512f", "avx512vl", "avx512dq", "avx512bw", "avx512cd"} <= fl
    except OSError:
        pass
    return False


def _f32(a, shape=None):
    a = np.ascontiguousarray(a, dtype=np.float32)
    return a if shape is None else a.reshape(shape)


class _Lib:
    """ctypes wrapper shared by the three checkers (same call shapes)."""

    def __init__(self, path: str, prefix: str, real=np.float32, returns_steps: bool = True):
        self.path = path
        self.lib = ctypes.CDLL(path)
        self.prefix = prefix
        self.real = np.dtype(real)
        self.creal = c_float if self.real == np.float32 else c_double
        self.returns_steps = returns_steps
        self.last_steps = None

    # -- helpers ---------------------------------------------------------
    def isa(self):
        """ISA of a reference-library build ("avx512" / "avx2"): enoki's rcp(h) differs between the two."""
        f = getattr(self.lib, self.prefix + "isa", None)
        if f is None:
            return "port"
        f.restype = c_char_p
        return f().decode()

    def rcp(self, h):
        """rcp(h) as this library's volume::eval_grad computes it (reference builds; port: IEEE 1/h)."""
        f = getattr(self.lib, self.prefix + "rcp", None)
        if f is None:
            return np.float32(1.0) / np.float32(h)
        f.restype = c_float
        f.argtypes = [c_float]
        return np.float32(f(c_float(float(h))))

    def _fn(self, name, restype=None):
        f = getattr(self.lib, self.prefix + name)
        f.restype = restype if restype is not None else (_i64 if self.returns_steps else c_int)
        return f

    def _a(self, x):
        return np.ascontiguousarray(x, dtype=self.real)

    def _p(self, a):
        return a.ctypes.data_as(POINTER(self.creal))

    def _check(self, rc):
        if self.returns_steps:
            self.last_steps = int(rc)
        elif rc != 0:
            f = self.lib.ref_last_error
            f.restype = c_char_p
            raise RuntimeError(f().decode())

    @staticmethod
    def _res(rif, res):
        if res is None:
            res = tuple(rif.shape)
        W, H, D = (int(r) for r in res)
        if W * H * D != rif.size:
            raise RuntimeError("Resolution doesn't match data")
        return W, H, D

    # -- forward ---------------------------------------------------------
    def trace(self, rif, pos, vel, h, ds, res=None):
        rif = self._a(rif); W, H, D = self._res(rif, res)
        pos = self._a(pos); vel = self._a(vel); N = pos.shape[0]
        xt = np.empty_like(pos); vt = np.empty_like(pos)
        rc = self._fn("trace")(self._p(rif), W, H, D, self._p(pos), self._p(vel), _i64(N),
                               self.creal(h), self.creal(ds), self._p(xt), self._p(vt))
        self._check(rc)
        return xt, vt

    def trace_pln(self, rif, pos, vel, pln_o, pln_d, h, ds, res=None):
        rif = self._a(rif); W, H, D = self._res(rif, res)
        pos = self._a(pos); vel = self._a(vel); po = self._a(pln_o); pd = self._a(pln_d)
        N = pos.shape[0]
        xt = np.empty_like(pos); vt = np.empty_like(pos); m = np.zeros(N, dtype=np.uint8)
        rc = self._fn("trace_plane")(self._p(rif), W, H, D, self._p(pos), self._p(vel), self._p(po),
                                     self._p(pd), _i64(N), self.creal(h), self.creal(ds),
                                     self._p(xt), self._p(vt), m.ctypes.data_as(POINTER(c_uint8)))
        self._check(rc)
        return xt, vt, m.astype(bool)

    def trace_target(self, rif, pos, vel, target, h, ds, res=None):
        rif = self._a(rif); W, H, D = self._res(rif, res)
        pos = self._a(pos); vel = self._a(vel); t = self._a(target); N = pos.shape[0]
        xt = np.empty_like(pos); vt = np.empty_like(pos); d2 = np.empty(N, dtype=self.real)
        rc = self._fn("trace_target")(self._p(rif), W, H, D, self._p(pos), self._p(vel), self._p(t),
                                      _i64(N), self.creal(h), self.creal(ds), self._p(xt),
                                      self._p(vt), self._p(d2))
        self._check(rc)
        return xt, vt, d2

    def trace_sdf(self, rif, sdf, pos, vel, h, ds, res=None):
        rif = self._a(rif); sdf = self._a(sdf); W, H, D = self._res(rif, res)
        pos = self._a(pos); vel = self._a(vel); N = pos.shape[0]
        xt = np.empty_like(pos); vt = np.empty_like(pos)
        rc = self._fn("trace_sdf")(self._p(rif), self._p(sdf), W, H, D, self._p(pos), self._p(vel),
                                   _i64(N), self.creal(h), self.creal(ds), self._p(xt), self._p(vt))
        self._check(rc)
        return xt, vt

    def trace_cable(self, rif, radius, length, pos, vel, target, ds):
        rif = self._a(rif).ravel(); pos = self._a(pos); vel = self._a(vel); t = self._a(target)
        N = pos.shape[0]
        xt = np.empty_like(pos); vt = np.empty_like(pos); d2 = np.empty(N, dtype=self.real)
        rc = self._fn("trace_cable")(self._p(rif), c_int(rif.size), self.creal(radius),
                                     self.creal(length), self._p(pos), self._p(vel), self._p(t),
                                     _i64(N), self.creal(ds), self._p(xt), self._p(vt), self._p(d2))
        self._check(rc)
        return xt, vt, d2

    # -- adjoint ---------------------------------------------------------
    def backtrace(self, rif, xt, vt, dx, dv, h, ds, res=None):
        rif = self._a(rif); W, H, D = self._res(rif, res)
        xt = self._a(xt); vt = self._a(vt); dx = self._a(dx); dv = self._a(dv); N = xt.shape[0]
        g = np.empty(rif.size, dtype=self.real)
        rc = self._fn("backtrace")(self._p(rif), W, H, D, self._p(xt), self._p(vt), self._p(dx),
                                   self._p(dv), _i64(N), self.creal(h), self.creal(ds), self._p(g))
        self._check(rc)
        return g.reshape(rif.shape)

    def backtrace_sdf(self, rif, sdf, xt, vt, dx, dv, h, ds, res=None):
        rif = self._a(rif); sdf = self._a(sdf); W, H, D = self._res(rif, res)
        xt = self._a(xt); vt = self._a(vt); dx = self._a(dx); dv = self._a(dv); N = xt.shape[0]
        g = np.empty(rif.size, dtype=self.real)
        rc = self._fn("backtrace_sdf")(self._p(rif), self._p(sdf), W, H, D, self._p(xt),
                                       self._p(vt), self._p(dx), self._p(dv), _i64(N),
                                       self.creal(h), self.creal(ds), self._p(g))
        self._check(rc)
        return g.reshape(rif.shape)

    def backtrace_rays(self, rif, xt, vt, dx, dv, h, ds, pos0=None, sdf=None, res=None):
        """port / port64 only: (grad, dL/dpos, dL/dvel), see eikonal_oracle.c backtrace_rays."""
        rif = self._a(rif); W, H, D = self._res(rif, res)
        xt = self._a(xt); vt = self._a(vt); dx = self._a(dx); dv = self._a(dv); N = xt.shape[0]
        g = np.empty(rif.size, dtype=self.real)
        dpos = np.empty((N, 3), dtype=self.real); dvel = np.empty((N, 3), dtype=self.real)
        null = POINTER(self.creal)()
        p0 = self._a(pos0) if pos0 is not None else None
        sd = self._a(sdf) if sdf is not None else None
        rc = self._fn("backtrace_rays")(self._p(rif), self._p(sd) if sd is not None else null, W, H, D,
                                        self._p(xt), self._p(vt), self._p(dx), self._p(dv),
                                        self._p(p0) if p0 is not None else null, _i64(N),
                                        self.creal(h), self.creal(ds), self._p(g), self._p(dpos),
                                        self._p(dvel))
        self._check(rc)
        return g.reshape(rif.shape), dpos, dvel

    def backtrace_cable(self, rif, radius, length, xt, vt, dx, dv, ds):
        rif = self._a(rif).ravel()
        xt = self._a(xt); vt = self._a(vt); dx = self._a(dx); dv = self._a(dv); N = xt.shape[0]
        g = np.empty(rif.size, dtype=self.real)
        rc = self._fn("backtrace_cable")(self._p(rif), c_int(rif.size), self.creal(radius),
                                         self.creal(length), self._p(xt), self._p(vt), self._p(dx),
                                         self._p(dv), _i64(N), self.creal(ds), self._p(g))
        self._check(rc)
        return g

    # -- volume probes ---------------------------------------------------
    def volume_eval(self, rif, pos, h, res=None):
        rif = self._a(rif); W, H, D = self._res(rif, res); pos = self._a(pos); N = pos.shape[0]
        n = np.empty(N, dtype=self.real); g = np.empty((N, 3), dtype=self.real)
        hs = np.empty((N, 3), dtype=self.real)
        rc = self._fn("volume_eval", None if not self.returns_steps else c_int)(
            self._p(rif), W, H, D, self._p(pos), _i64(N), self.creal(h), self._p(n), self._p(g),
            self._p(hs))
        if not self.returns_steps:
            self._check(rc)
        return n, g, hs

    def volume_splat(self, data, pos, val, grad, h, res=None):
        data = self._a(data).copy(); W, H, D = self._res(data, res)
        pos = self._a(pos); val = self._a(val); grad = self._a(grad); N = pos.shape[0]
        rc = self._fn("volume_splat", None if not self.returns_steps else c_int)(
            self._p(data), W, H, D, self._p(pos), self._p(val), self._p(grad), _i64(N),
            self.creal(h))
        if not self.returns_steps:
            self._check(rc)
        return data

    def cylinder_eval(self, rif, radius, length, pos):
        rif = self._a(rif).ravel(); pos = self._a(pos); N = pos.shape[0]
        n = np.empty(N, dtype=self.real); g = np.empty((N, 3), dtype=self.real)
        hs = np.empty((N, 3, 3), dtype=self.real)
        rc = self._fn("cylinder_eval", None if not self.returns_steps else c_int)(
            self._p(rif), c_int(rif.size), self.creal(radius), self.creal(length), self._p(pos),
            _i64(N), self._p(n), self._p(g), self._p(hs))
        if not self.returns_steps:
            self._check(rc)
        return n, g, hs

    def cylinder_splat(self, data, radius, length, pos, val, grad):
        data = self._a(data).ravel().copy(); pos = self._a(pos); val = self._a(val)
        grad = self._a(grad); N = pos.shape[0]
        rc = self._fn("cylinder_splat", None if not self.returns_steps else c_int)(
            self._p(data), c_int(data.size), self.creal(radius), self.creal(length), self._p(pos),
            self._p(val), self._p(grad), _i64(N))
        if not self.returns_steps:
            self._check(rc)
        return data


def _load_port(name, prefix, real):
    path = os.path.join(_HERE, "_build", name)
    if not os.path.exists(path):
        build()
    return _Lib(path, prefix, real, returns_steps=True)


def _load_ref():
    cands = ["libdrrt_ref_v4.so", "libdrrt_ref_v3.so"] if _cpu_has_avx512() else ["libdrrt_ref_v3.so"]
    for c in cands:
        path = os.path.join(_HERE, "_ref", c)
        if os.path.exists(path):
            return _Lib(path, "ref_", np.float32, returns_steps=False)
    return None


class _RefAD:
    """ctypes wrapper of oracle/ref_ad_shim.cpp (reference Tracer<true,false> + enoki tape)."""

    def __init__(self, path):
        self.path = path
        self.lib = ctypes.CDLL(path)
        self.lib.refad_trace_backward.restype = c_int
        self.lib.refad_last_error.restype = c_char_p

    def trace_backward(self, rif, pos, vel, h, ds, seed_x, seed_v, sdf=None):
        """(xt, vt, d_rif, d_pos, d_vel) of ADTracerC (sdf=None) / ADSDFTracerC forward + backward."""
        rif = _f32(rif); pos = _f32(pos); vel = _f32(vel); sx = _f32(seed_x); sv = _f32(seed_v)
        W, H, D = (int(r) for r in rif.shape)
        N = pos.shape[0]
        P = lambda a: a.ctypes.data_as(POINTER(c_float))
        sd = _f32(sdf) if sdf is not None else None
        xt = np.empty_like(pos); vt = np.empty_like(pos)
        dr = np.empty(rif.size, dtype=np.float32); dp = np.empty_like(pos); dv = np.empty_like(pos)
        rc = self.lib.refad_trace_backward(P(rif), P(sd) if sd is not None else POINTER(c_float)(), W, H, D,
                                           P(pos), P(vel), _i64(N), c_float(h), c_float(ds), P(sx), P(sv),
                                           P(xt), P(vt), P(dr), P(dp), P(dv))
        if rc != 0:
            raise RuntimeError(self.lib.refad_last_error().decode())
        return xt, vt, dr.reshape(rif.shape), dp, dv


def _load_refad():
    path = os.path.join(_HERE, "_ref", "libdrrt_refad_v3.so")
    return _RefAD(path) if os.path.exists(path) else None


def ref_builds():
    """Every reference-library build that can run on this host: {"v4": _Lib, "v3": _Lib}."""
    out = {}
    for v in (("v4", "v3") if _cpu_has_avx512() else ("v3",)):
        path = os.path.join(_HERE, "_ref", f"libdrrt_ref_{v}.so")
        if os.path.exists(path):
            try:
                out[v] = _Lib(path, "ref_", np.float32, returns_steps=False)
            except OSError:
                pass
    return out


def ref_for_spacing(h):
    """The reference build whose rcp(h) equals the correctly rounded 1/h (what the CUDA kernels and the
    restatement use), or None.  The reference's AVX2 / AVX-512 / CUDA builds approximate 1/h three different
    ways (enoki rcp + Newton, rcp14 + Newton, rcp.approx.ftz); on lattice-exact samples the last ulp of
    p * rcp(h) selects the cell, so two builds of the SAME source differ there by 5e-3 .. 3e-2 in the gradient
    (tests/test_ref_parity_shapes.py records it).  Against the build that shares 1/h the trajectory is
    bit-identical."""
    want = np.float32(1.0) / np.float32(h)
    for v, lib in ref_builds().items():
        try:
            if lib.rcp(h) == want:
                return lib
        except AttributeError:
            pass
    return None


_cache = {}


def __getattr__(name):
    if name == "refad":
        if name not in _cache:
            _cache[name] = _load_refad()
        return _cache[name]
    if name in ("port", "port64", "ref"):
        if name not in _cache:
            if name == "port":
                _cache[name] = _load_port("liborc.so", "orc_", np.float32)
            elif name == "port64":
                _cache[name] = _load_port("liborc64.so", "orc64_", np.float64)
            else:
                _cache[name] = _load_ref()
        return _cache[name]
    raise AttributeError(name)


def run_threads(fn, n):
    """fn(i) on n host threads (ctypes calls release the GIL) -> list of results."""
    import threading
    out, thr = [None] * n, []

    def work(i):
        out[i] = fn(i)
    for i in range(n):
        t = threading.Thread(target=work, args=(i,))
        t.start(); thr.append(t)
    for t in thr:
        t.join()
    return out


def chunk_bounds(n, threads=None):
    threads = max(1, min(threads or len(os.sched_getaffinity(0)), n // 256 or 1))
    return [(n * i) // threads for i in range(threads + 1)]


def trace_backtrace_mt(lib, rif, pos, vel, dx, dv, h, ds, threads=None):
    """Forward trace + adjoint of `lib` (ref / port) with the rays split into contiguous chunks over host
    threads, private gradient grids summed in float64 (the reference itself is single-threaded; for
    trace / backtrace a ray's trajectory and gradient contributions do not depend on the rest of the batch).
    -> (xt, vt, grad float64)."""
    bounds = chunk_bounds(pos.shape[0], threads)

    def one(i):
        a, b = bounds[i], bounds[i + 1]
        xt, vt = lib.trace(rif, pos[a:b], vel[a:b], h, ds)
        return xt, vt, lib.backtrace(rif, xt, vt, dx[a:b], dv[a:b], h, ds)
    parts = run_threads(one, len(bounds) - 1)
    xt = np.concatenate([p[0] for p in parts]); vt = np.concatenate([p[1] for p in parts])
    grad = np.zeros(parts[0][2].shape, dtype=np.float64)
    for p in parts:
        grad += p[2]
    return xt, vt, grad
