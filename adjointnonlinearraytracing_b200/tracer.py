"""torch.autograd.Function wrappers with the names and call signatures of the
reference's ``core/tracer.py``.

``Back*TracerC`` = forward trace + adjoint backtrace (the paper's method; reference
``core/tracer.py:294-526``).  The ``AD*`` classes of the reference differentiate the
unrolled loop with enoki's tape (``core/tracer.py:16-291``); tape autodiff is the
thing the adjoint replaces, so here they keep their call signature and route to
the same adjoint kernels.  ``ADTracerC`` and ``ADSDFTracerC`` also return
dL/dx and dL/dv of the launch state like the reference (``core/tracer.py:53-60``),
read off the adjoint's final co-state (``ops.backtrace_rays``); the plane, target and
cable AD variants return the ``rif`` gradient only, like their ``Back*`` twins.

Differences from the reference that a caller can observe: none in values beyond
the stated float32 tolerances; ``ctx`` keeps torch tensors instead of enoki arrays (the forward
outputs through ``save_for_backward``, so no reference cycle keeps a batch alive until the cyclic GC)
and there is no ``enoki.cuda_malloc_trim()`` because nothing is pooled outside
torch's allocator.
"""
from __future__ import annotations

import torch

from . import ops

# module-level switches (tests flip them; the reference has no equivalent)
FLAGS = 0


class BackTracerC(torch.autograd.Function):
    """reference core/tracer.py:294-335"""

    @staticmethod
    def forward(ctx, rif, x, v, h, ds):
        ctx.rif = rif.detach()
        ctx.h, ctx.ds = h, ds
        outx, outv = ops.trace(ctx.rif, x, v, h, ds, flags=FLAGS)
        ctx.save_for_backward(outx, outv)  # not ctx attributes: output -> grad_fn -> ctx -> output would be a cycle
        return outx, outv

    @staticmethod
    def backward(ctx, grad_x, grad_v):
        drif = ops.backtrace(ctx.rif, *ctx.saved_tensors, grad_x, grad_v, ctx.h, ctx.ds, flags=FLAGS)
        return drif, None, None, None, None


class BackPlaneTracerC(torch.autograd.Function):
    """reference core/tracer.py:338-386"""

    @staticmethod
    def forward(ctx, rif, x, v, sp, sn, h, ds):
        ctx.rif = rif.detach()
        ctx.h, ctx.ds = h, ds
        outx, outv, outmask = ops.trace_pln(ctx.rif, x, v, sp, sn, h, ds, flags=FLAGS)
        ctx.mark_non_differentiable(outmask)
        ctx.save_for_backward(outx, outv)
        return outx, outv, outmask

    @staticmethod
    def backward(ctx, grad_x, grad_v, outmask):
        # the reference zeroes grad_x on the mask handed to backward when there is one
        # (core/tracer.py:369-370); autograd passes None for the non-differentiable mask
        if outmask is not None and outmask.dtype == torch.bool:
            grad_x = grad_x.clone()
            grad_x[outmask] = 0
        drif = ops.backtrace(ctx.rif, *ctx.saved_tensors, grad_x, grad_v, ctx.h, ctx.ds, flags=FLAGS)
        return drif, None, None, None, None, None, None


class BackTargetTracerC(torch.autograd.Function):
    """reference core/tracer.py:389-432"""

    @staticmethod
    def forward(ctx, rif, x, v, sp, h, ds):
        ctx.rif = rif.detach()
        ctx.h, ctx.ds = h, ds
        outx, outv, outdist2 = ops.trace_target(ctx.rif, x, v, sp, h, ds, flags=FLAGS)
        ctx.save_for_backward(outx, outv)
        return outx, outv, outdist2

    @staticmethod
    def backward(ctx, grad_x, grad_v, outdist):
        drif = ops.backtrace(ctx.rif, *ctx.saved_tensors, grad_x, grad_v, ctx.h, ctx.ds, flags=FLAGS)
        return drif, None, None, None, None, None


class BackSDFTracerC(torch.autograd.Function):
    """reference core/tracer.py:435-479"""

    @staticmethod
    def forward(ctx, rif, sdf, x, v, h, ds):
        ctx.rif = rif.detach()
        ctx.sdf = sdf.detach()
        ctx.h, ctx.ds = h, ds
        outx, outv = ops.trace_sdf(ctx.rif, ctx.sdf, x, v, h, ds, flags=FLAGS)
        ctx.save_for_backward(outx, outv)
        return outx, outv

    @staticmethod
    def backward(ctx, grad_x, grad_v):
        drif = ops.backtrace_sdf(ctx.rif, ctx.sdf, *ctx.saved_tensors, grad_x, grad_v, ctx.h, ctx.ds,
                                 flags=FLAGS)
        return drif, None, None, None, None, None


class BackCableTracerC(torch.autograd.Function):
    """reference core/tracer.py:482-526"""

    @staticmethod
    def forward(ctx, rif, radius, length, x, v, sp, ds):
        ctx.rif = rif.detach()
        ctx.radius, ctx.length, ctx.ds = radius, length, ds
        outx, outv, outdist2 = ops.trace_cable(ctx.rif, radius, length, x, v, sp, ds, flags=FLAGS)
        ctx.save_for_backward(outx, outv)
        return outx, outv, outdist2

    @staticmethod
    def backward(ctx, grad_x, grad_v, outdist):
        drif = ops.backtrace_cable(ctx.rif, ctx.radius, ctx.length, *ctx.saved_tensors, grad_x, grad_v,
                                   ctx.ds, flags=FLAGS)
        return drif.reshape(ctx.rif.shape), None, None, None, None, None, None


class ADTracerC(torch.autograd.Function):
    """reference core/tracer.py:16-68: gradients w.r.t. rif, x and v (each only when it
    requires grad), served by the adjoint instead of enoki's tape."""

    @staticmethod
    def forward(ctx, rif, x, v, h, ds):
        ctx.rif = rif.detach()
        ctx.h, ctx.ds = h, ds
        ctx.x0 = x.detach()
        outx, outv = ops.trace(ctx.rif, x, v, h, ds, flags=FLAGS)
        ctx.save_for_backward(outx, outv)  # not ctx attributes: output -> grad_fn -> ctx -> output would be a cycle
        return outx, outv

    @staticmethod
    def backward(ctx, grad_x, grad_v):
        need_rif, need_x, need_v = ctx.needs_input_grad[:3]
        if not (need_x or need_v):
            drif = ops.backtrace(ctx.rif, *ctx.saved_tensors, grad_x, grad_v, ctx.h, ctx.ds, flags=FLAGS)
            return drif, None, None, None, None
        drif, dx0, dv0 = ops.backtrace_rays(ctx.rif, *ctx.saved_tensors, grad_x, grad_v, ctx.h, ctx.ds,
                                            pos0=ctx.x0, flags=FLAGS)
        return (drif if need_rif else None, dx0 if need_x else None, dv0 if need_v else None, None, None)


class ADSDFTracerC(torch.autograd.Function):
    """reference core/tracer.py:247-291"""

    @staticmethod
    def forward(ctx, rif, sdf, x, v, h, ds):
        ctx.rif = rif.detach()
        ctx.sdf = sdf.detach()
        ctx.h, ctx.ds = h, ds
        ctx.x0 = x.detach()
        outx, outv = ops.trace_sdf(ctx.rif, ctx.sdf, x, v, h, ds, flags=FLAGS)
        ctx.save_for_backward(outx, outv)
        return outx, outv

    @staticmethod
    def backward(ctx, grad_x, grad_v):
        need_rif, _, need_x, need_v = ctx.needs_input_grad[:4]
        if not (need_x or need_v):
            drif = ops.backtrace_sdf(ctx.rif, ctx.sdf, *ctx.saved_tensors, grad_x, grad_v, ctx.h, ctx.ds,
                                     flags=FLAGS)
            return drif, None, None, None, None, None
        drif, dx0, dv0 = ops.backtrace_rays(ctx.rif, *ctx.saved_tensors, grad_x, grad_v, ctx.h, ctx.ds,
                                            pos0=ctx.x0, sdf=ctx.sdf, flags=FLAGS)
        return (drif if need_rif else None, None, dx0 if need_x else None, dv0 if need_v else None,
                None, None)


# Remaining tape-autodiff names of the reference (core/tracer.py:71-244): same signatures,
# served by the adjoint (rif gradient only).
ADPlaneTracerC = BackPlaneTracerC
ADTargetTracerC = BackTargetTracerC
ADCableTracerC = BackCableTracerC
