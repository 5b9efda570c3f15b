"""torch.autograd.Function wrappers with the names and call signatures of the
reference's ``core/tracer.py``.

``Back*TracerC`` = forward trace + adjoint backtrace (the paper's method; reference
``core/tracer.py:294-526``).  The ``AD*`` classes of the reference differentiate the
unrolled loop with enoki's tape (``core/tracer.py:16-291``); tape autodiff is the
thing the adjoint replaces, so here they keep their call signature and route to
the same adjoint kernels (gradient w.r.t. ``rif`` only, like ``Back*``).

Differences from the reference that a caller can observe: none in values beyond
the stated float32 tolerances; ``ctx`` keeps torch tensors instead of enoki arrays
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
        ctx.outx, ctx.outv = ops.trace(ctx.rif, x, v, h, ds, flags=FLAGS)
        return ctx.outx, ctx.outv

    @staticmethod
    def backward(ctx, grad_x, grad_v):
        drif = ops.backtrace(ctx.rif, ctx.outx, ctx.outv, grad_x, grad_v, ctx.h, ctx.ds, flags=FLAGS)
        return drif, None, None, None, None


class BackPlaneTracerC(torch.autograd.Function):
    """reference core/tracer.py:338-386"""

    @staticmethod
    def forward(ctx, rif, x, v, sp, sn, h, ds):
        ctx.rif = rif.detach()
        ctx.h, ctx.ds = h, ds
        ctx.outx, ctx.outv, ctx.outmask = ops.trace_pln(ctx.rif, x, v, sp, sn, h, ds, flags=FLAGS)
        ctx.mark_non_differentiable(ctx.outmask)
        return ctx.outx, ctx.outv, ctx.outmask

    @staticmethod
    def backward(ctx, grad_x, grad_v, outmask):
        # the reference zeroes grad_x on the mask handed to backward when there is one
        # (core/tracer.py:369-370); autograd passes None for the non-differentiable mask
        if outmask is not None and outmask.dtype == torch.bool:
            grad_x = grad_x.clone()
            grad_x[outmask] = 0
        drif = ops.backtrace(ctx.rif, ctx.outx, ctx.outv, grad_x, grad_v, ctx.h, ctx.ds, flags=FLAGS)
        return drif, None, None, None, None, None, None


class BackTargetTracerC(torch.autograd.Function):
    """reference core/tracer.py:389-432"""

    @staticmethod
    def forward(ctx, rif, x, v, sp, h, ds):
        ctx.rif = rif.detach()
        ctx.h, ctx.ds = h, ds
        ctx.outx, ctx.outv, outdist2 = ops.trace_target(ctx.rif, x, v, sp, h, ds, flags=FLAGS)
        return ctx.outx, ctx.outv, outdist2

    @staticmethod
    def backward(ctx, grad_x, grad_v, outdist):
        drif = ops.backtrace(ctx.rif, ctx.outx, ctx.outv, grad_x, grad_v, ctx.h, ctx.ds, flags=FLAGS)
        return drif, None, None, None, None, None


class BackSDFTracerC(torch.autograd.Function):
    """reference core/tracer.py:435-479"""

    @staticmethod
    def forward(ctx, rif, sdf, x, v, h, ds):
        ctx.rif = rif.detach()
        ctx.sdf = sdf.detach()
        ctx.h, ctx.ds = h, ds
        ctx.outx, ctx.outv = ops.trace_sdf(ctx.rif, ctx.sdf, x, v, h, ds, flags=FLAGS)
        return ctx.outx, ctx.outv

    @staticmethod
    def backward(ctx, grad_x, grad_v):
        drif = ops.backtrace_sdf(ctx.rif, ctx.sdf, ctx.outx, ctx.outv, grad_x, grad_v, ctx.h, ctx.ds,
                                 flags=FLAGS)
        return drif, None, None, None, None, None


class BackCableTracerC(torch.autograd.Function):
    """reference core/tracer.py:482-526"""

    @staticmethod
    def forward(ctx, rif, radius, length, x, v, sp, ds):
        ctx.rif = rif.detach()
        ctx.radius, ctx.length, ctx.ds = radius, length, ds
        ctx.outx, ctx.outv, outdist2 = ops.trace_cable(ctx.rif, radius, length, x, v, sp, ds, flags=FLAGS)
        return ctx.outx, ctx.outv, outdist2

    @staticmethod
    def backward(ctx, grad_x, grad_v, outdist):
        drif = ops.backtrace_cable(ctx.rif, ctx.radius, ctx.length, ctx.outx, ctx.outv, grad_x, grad_v,
                                   ctx.ds, flags=FLAGS)
        return drif.reshape(ctx.rif.shape), None, None, None, None, None, None


# Tape-autodiff names of the reference (core/tracer.py:16-291): same signatures,
# served by the adjoint.
ADTracerC = BackTracerC
ADPlaneTracerC = BackPlaneTracerC
ADTargetTracerC = BackTargetTracerC
ADSDFTracerC = BackSDFTracerC
ADCableTracerC = BackCableTracerC
