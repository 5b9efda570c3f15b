"""TEST INFRASTRUCTURE ONLY -- never imported by the product.

Float64 tape differentiation of the forward eikonal loop with torch autograd on the
CPU: the same thing the reference's ``ADTracerC`` does with enoki's tape
(reference core/tracer.py:16-68 over src/tracer.cpp:35-100), restated with torch ops
so that dL/d rif, dL/d pos and dL/d vel of the *launch* state are available as an
independent check of the adjoint's final co-state.

Follows src/tracer.cpp:35-100 and src/volume.cpp:101-181 step for step (masked
gathers, first-outside-sample capture, float-then-truncate loop cap); pure-Python
loop over steps, vectorised over rays -- small cases only.
"""
import numpy as np
import torch


def trace_ad(rif, pos, vel, h, ds, res=None):
    """Differentiable Tracer::trace.  rif (3-D), pos, vel: float64 torch tensors (leaf or not).
    Returns (xt, vt).  NOTE the tape differentiates n(p) and grad n(p) through the trilinear
    weights, i.e. uses the true Jacobian of the piecewise-trilinear field, like enoki's tape."""
    if res is None:
        res = tuple(rif.shape)
    W, H, D = res
    flat = rif.reshape(-1)
    hi = torch.tensor([(W - 1) * h, (H - 1) * h, (D - 1) * h], dtype=torch.float64)

    def inb(p):
        return ((p >= 0) & (p < hi)).all(dim=1)

    def esc(p, v):
        return (((p < 0) & (v < 0)) | ((p >= hi) & (v > 0))).any(dim=1)

    x, v = pos, vel
    xt, vt = pos, vel
    inside = inb(x.detach())
    escaped = torch.zeros(pos.shape[0], dtype=torch.bool)
    max_steps = int(np.float32(4) * np.float32(h) * np.float32(max(res)) / np.float32(ds))
    for _ in range(max_steps):
        m = inside.to(torch.float64)
        n, nx = _eval_grad_full(flat, res, x, h, m)
        v = v + (ds * n)[:, None] * nx
        x = x + ds * v
        cur = inb(x.detach())
        cross = inside & ~cur
        escaped = escaped | cross | esc(x.detach(), v.detach())
        xt = torch.where(cross[:, None], x, xt)
        vt = torch.where(cross[:, None], v, vt)
        if bool(escaped.all()):
            break
        inside = cur
    if not bool(escaped.all()):
        xt = torch.where(escaped[:, None], xt, x)
    return xt, vt


def _eval_grad_full(flat, res, p, h, mask):
    """n and grad n with the tape running through the trilinear weights (true Jacobian)."""
    W, H, D = res
    pm = p / h
    fl = torch.floor(pm.detach())
    w0 = pm - fl
    w1 = 1.0 - w0
    hi = torch.tensor([W - 1, H - 1, D - 1], dtype=torch.long)
    i0 = torch.minimum(torch.clamp(fl.long(), min=0), hi)
    i1 = torch.minimum(torch.clamp(fl.long() + 1, min=0), hi)

    def g(ix, iy, iz):
        return flat[(iz * H + iy) * W + ix] * mask

    x0, y0, z0 = i0[:, 0], i0[:, 1], i0[:, 2]
    x1, y1, z1 = i1[:, 0], i1[:, 1], i1[:, 2]
    v000, v100, v010, v110 = g(x0, y0, z0), g(x1, y0, z0), g(x0, y1, z0), g(x1, y1, z0)
    v001, v101, v011, v111 = g(x0, y0, z1), g(x1, y0, z1), g(x0, y1, z1), g(x1, y1, z1)
    ax, ay, az = w0[:, 0], w0[:, 1], w0[:, 2]
    bx, by, bz = w1[:, 0], w1[:, 1], w1[:, 2]
    n = (bz * (by * (bx * v000 + ax * v100) + ay * (bx * v010 + ax * v110))
         + az * (by * (bx * v001 + ax * v101) + ay * (bx * v011 + ax * v111)))
    gx = (bz * (by * (v100 - v000) + ay * (v110 - v010)) + az * (by * (v101 - v001) + ay * (v111 - v011))) / h
    gy = (bz * (bx * (v010 - v000) + ax * (v110 - v100)) + az * (bx * (v011 - v001) + ax * (v111 - v101))) / h
    gz = (by * (bx * (v001 - v000) + ax * (v101 - v100)) + ay * (bx * (v011 - v010) + ax * (v111 - v110))) / h
    return n, torch.stack([gx, gy, gz], dim=1)


def gradients(rif, pos, vel, h, ds, wx, wv):
    """dL/d rif, dL/d pos, dL/d vel for L = sum(wx * xt) + sum(wv * vt); numpy in, numpy out."""
    r = torch.tensor(np.asarray(rif, dtype=np.float64), requires_grad=True)
    p = torch.tensor(np.asarray(pos, dtype=np.float64), requires_grad=True)
    v = torch.tensor(np.asarray(vel, dtype=np.float64), requires_grad=True)
    xt, vt = trace_ad(r, p, v, float(h), float(ds))
    loss = (xt * torch.tensor(np.asarray(wx, dtype=np.float64))).sum() + \
           (vt * torch.tensor(np.asarray(wv, dtype=np.float64))).sum()
    loss.backward()
    return (r.grad.numpy(), p.grad.numpy(), v.grad.numpy(), xt.detach().numpy(), vt.detach().numpy())
