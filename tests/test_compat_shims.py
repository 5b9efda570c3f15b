"""``compat/`` makes ``import enoki`` / ``import drrt`` resolve to stand-ins over
libdrrt_b200.so so that the reference's ``core/tracer.py`` runs unmodified.

CPU (build container only, where /root/reference exists): the reference file imports against the
stand-ins and exposes its autograd classes.  GPU: callers written in the reference's calling style
(enoki array constructors, ``drrt.TracerC().trace(...)``, ``.torch()``, the autodiff tape protocol of
core/tracer.py:16-68 and :294-335) give the same numbers as the package's own ``tracer`` module; when
the reference checkout is present the same is done through the reference's own classes.
"""
import importlib.util
import os
import sys

import numpy as np
import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
COMPAT = os.path.join(ROOT, "compat")
REF_TRACER = os.path.join(os.environ.get("DRRT_REFERENCE", "/root/reference"), "core", "tracer.py")


@pytest.fixture()
def compat_path():
    sys.path.insert(0, COMPAT)
    saved = {k: sys.modules.pop(k) for k in list(sys.modules) if k == "drrt" or k == "enoki" or k.startswith("enoki.")}
    yield
    for k in list(sys.modules):
        if k == "drrt" or k == "enoki" or k.startswith("enoki."):
            del sys.modules[k]
    sys.modules.update(saved)
    sys.path.remove(COMPAT)


def _load_reference_tracer():
    spec = importlib.util.spec_from_file_location("reference_core_tracer", REF_TRACER)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


@pytest.mark.skipif(not os.path.exists(REF_TRACER), reason="reference checkout not present (GPU box)")
def test_reference_tracer_py_imports_against_the_stand_ins(cuda_lib, compat_path):
    mod = _load_reference_tracer()
    for name in ("BackTracerC", "BackPlaneTracerC", "BackTargetTracerC", "BackSDFTracerC", "BackCableTracerC",
                 "ADTracerC", "ADSDFTracerC"):
        assert issubclass(getattr(mod, name), torch.autograd.Function)
    import drrt
    assert {"trace", "trace_pln", "trace_target", "trace_sdf", "trace_cable", "backtrace", "backtrace_sdf",
            "backtrace_cable"} <= set(dir(drrt.TracerC))
    assert {"trace", "trace_sdf", "trace_target", "backtrace"} <= set(dir(drrt.TracerS))


def test_stand_in_arrays_cpu(compat_path):
    import enoki
    from enoki.dynamic import Float32 as FloatS, Vector3f as Vector3fS
    f = FloatS(torch.arange(6.0).reshape(2, 3))
    assert f.torch().shape == (6,) and len(f) == 6
    v = Vector3fS(torch.ones(4, 3))
    assert v.torch().shape == (4, 3)
    with pytest.raises(RuntimeError):
        Vector3fS(torch.ones(4, 2))
    enoki.set_requires_gradient(f, True)
    assert enoki.requires_gradient(f) and not enoki.requires_gradient(v)
    assert float(enoki.gradient(f).torch().abs().sum()) == 0.0
    enoki.set_gradient(f, torch.ones(6))
    assert float(enoki.gradient(f).torch().sum()) == 6.0
    enoki.cuda_malloc_trim()


# ------------------------------------------------------------------ GPU
def _reference_style_classes():
    """Callers in the reference's style (restated from core/tracer.py:16-68, :294-335)."""
    import drrt
    import enoki
    from enoki.cuda import Float32 as FloatC, Vector3f as Vector3fC
    from enoki.cuda_autodiff import Float32 as FloatD, Vector3f as Vector3fD

    class Back(torch.autograd.Function):
        @staticmethod
        def forward(ctx, rif, x, v, h, ds):
            ctx.shape, ctx.h, ctx.ds = rif.shape, h, ds
            ctx.rif = FloatC(rif.flatten())
            ctx.outx, ctx.outv = drrt.TracerC().trace(ctx.rif, ctx.shape, Vector3fC(x.detach()), Vector3fC(v.detach()), h, ds)
            enoki.cuda_malloc_trim()
            return ctx.outx.torch(), ctx.outv.torch()

        @staticmethod
        def backward(ctx, gx, gv):
            r = drrt.TracerC().backtrace(ctx.rif, ctx.shape, ctx.outx, ctx.outv, Vector3fC(gx), Vector3fC(gv), ctx.h, ctx.ds)
            return r.torch().reshape(*ctx.shape), None, None, None, None

    class AD(torch.autograd.Function):
        @staticmethod
        def forward(ctx, rif, x, v, h, ds):
            ctx.shape = rif.shape
            ctx.rif, ctx.x, ctx.v = FloatD(rif.flatten()), Vector3fD(x), Vector3fD(v)
            for arr, t in ((ctx.rif, rif), (ctx.x, x), (ctx.v, v)):
                enoki.set_requires_gradient(arr, t.requires_grad)
            ctx.outx, ctx.outv = drrt.TracerD().trace(ctx.rif, ctx.shape, ctx.x, ctx.v, h, ds)
            return ctx.outx.torch(), ctx.outv.torch()

        @staticmethod
        def backward(ctx, gx, gv):
            enoki.set_gradient(ctx.outx, Vector3fC(gx))
            enoki.set_gradient(ctx.outv, Vector3fC(gv))
            FloatD.backward()
            pick = lambda a: enoki.gradient(a).torch() if enoki.requires_gradient(a) else None
            g = pick(ctx.rif)
            return (g.reshape(*ctx.shape) if g is not None else None), pick(ctx.x), pick(ctx.v), None, None

    return Back, AD


@pytest.mark.gpu
def test_gpu_reference_style_callers_match_the_package(cuda_lib, compat_path):
    from adjointnonlinearraytracing_b200 import scenes as sc, tracer
    ng = 17
    h = float(sc.grid_spacing(ng)); ds = h / 2
    rif = sc.field_bump(ng)
    pos, vel = sc.rays_cone(24, spread=0.1, backoff=2.3 * ds)
    rng = np.random.default_rng(2)
    wx = torch.from_numpy(rng.standard_normal(pos.shape).astype(np.float32)).cuda()
    wv = torch.from_numpy(rng.standard_normal(pos.shape).astype(np.float32)).cuda()
    T = lambda a: torch.from_numpy(a).cuda()

    def run(fn, ray_grads):
        n = T(rif).requires_grad_(True)
        x = T(pos).requires_grad_(ray_grads); v = T(vel).requires_grad_(ray_grads)
        xt, vt = fn(n, x, v, h, ds)
        ((xt * wx).sum() + (vt * wv).sum()).backward()
        return xt.detach(), vt.detach(), n.grad, x.grad, v.grad

    classes = [_reference_style_classes()]
    if os.path.exists(REF_TRACER):
        ref = _load_reference_tracer()
        classes.append((ref.BackTracerC, ref.ADTracerC))
    ours_b = run(tracer.BackTracerC.apply, False)
    ours_a = run(tracer.ADTracerC.apply, True)
    for Back, AD in classes:
        b = run(Back.apply, False)
        assert torch.equal(b[0], ours_b[0]) and torch.equal(b[1], ours_b[1])
        assert float((b[2] - ours_b[2]).norm() / ours_b[2].norm()) <= 2e-6
        a = run(AD.apply, True)
        assert torch.equal(a[0], ours_a[0])
        for k in (2, 3, 4):
            assert float((a[k] - ours_a[k]).norm() / ours_a[k].norm()) <= 2e-6, k
