"""The "tape" of the enoki stand-in: the differentiable drrt.TracerD calls made since the last
``Float32.backward()``.  Each record is differentiated by the adjoint kernels."""
from __future__ import annotations

import torch

RECORDS = []


def record(kind, inputs, outputs, params):
    RECORDS.append((kind, inputs, outputs, params))


def _acc(arr, g):
    from . import ArrayBase
    if arr is None or not isinstance(arr, ArrayBase) or not arr._requires_grad:
        return
    if arr._grad is None:
        arr._grad = type(arr)(g)
    else:
        arr._grad = type(arr)(arr._grad.torch() + g.reshape(arr._grad.torch().shape))


def run_backward():
    from adjointnonlinearraytracing_b200 import ops
    while RECORDS:
        kind, inp, out, prm = RECORDS.pop()
        outx, outv = out[0], out[1]
        gx = outx._grad.torch() if outx._grad is not None else torch.zeros_like(outx.torch())
        gv = outv._grad.torch() if outv._grad is not None else torch.zeros_like(outv.torch())
        if kind == "cable":
            g = ops.backtrace_cable(inp["rif"].torch(), prm["radius"], prm["length"], outx.torch(), outv.torch(), gx,
                                    gv, prm["ds"])
            _acc(inp["rif"], g.reshape(-1))
            continue
        rif = inp["rif"].torch().reshape(prm["res"])
        sdf = inp["sdf"].torch().reshape(prm["res"]) if kind == "sdf" else None
        if kind in ("plain", "sdf"):
            # launch-state gradients as well (core/tracer.py:53-60)
            g, dpos, dvel = ops.backtrace_rays(rif, outx.torch(), outv.torch(), gx, gv, prm["h"], prm["ds"],
                                               pos0=inp["x"].torch(), sdf=sdf)
            _acc(inp["x"], dpos); _acc(inp["v"], dvel)
        else:   # plane / target: the adjoint of the volume path, gradient w.r.t. the grid
            g = ops.backtrace(rif, outx.torch(), outv.torch(), gx, gv, prm["h"], prm["ds"])
        _acc(inp["rif"], g.reshape(-1))
