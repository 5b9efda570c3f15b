"""Minimal stand-in for the `enoki` Python module, exactly as far as the reference's
``core/tracer.py`` uses it (reference core/tracer.py:2-6, :27-29, :47-50, :64, :299-315), so that
file runs unmodified with ``compat/`` on ``sys.path``.

Arrays are thin wrappers over torch tensors (``.torch()`` hands the tensor back); the
``cuda_autodiff`` flavour carries the three things the reference touches -- a requires-gradient
flag, a gradient slot and ``Float32.backward()`` -- and "the tape" is the list of ``drrt.TracerD``
calls since the last backward, each of which is differentiated by the adjoint kernels
(``ops.backtrace_rays``) instead of by replaying a recorded instruction stream.
Nothing here computes on the host: the tensors stay where torch put them.
"""
from __future__ import annotations

import torch

__all__ = ["set_requires_gradient", "requires_gradient", "set_gradient", "gradient", "cuda_malloc_trim",
           "cuda_sync", "cuda_eval"]


def _as_tensor(x, device=None):
    if isinstance(x, ArrayBase):
        return x.torch()
    if isinstance(x, torch.Tensor):
        return x
    return torch.as_tensor(x, dtype=torch.float32, device=device)


class ArrayBase:
    """Common part of Float32 / Vector3f: owns one torch tensor."""
    _device = "cuda"   # overridden by the `dynamic` (host) flavour
    _autodiff = False

    def __init__(self, t):
        self._t = t
        self._requires_grad = False
        self._grad = None

    def torch(self):
        return self._t

    def numpy(self):
        return self._t.detach().cpu().numpy()

    def __len__(self):
        return int(self._t.shape[0])


def make_types(device: str, autodiff: bool):
    """Float32 / Vector3f classes for one enoki flavour (enoki.cuda, enoki.cuda_autodiff, enoki.dynamic)."""

    class Float32(ArrayBase):
        _device, _autodiff = device, autodiff

        def __init__(self, x=0.0):
            t = _as_tensor(x, device).detach()
            super().__init__(t.to(device=device, dtype=torch.float32).reshape(-1).contiguous())

        @classmethod
        def backward(cls):
            """reference core/tracer.py:50 ``FloatD.backward()``: differentiate everything recorded."""
            from . import _tape
            _tape.run_backward()

    class Vector3f(ArrayBase):
        _device, _autodiff = device, autodiff

        def __init__(self, x=None, y=None, z=None):
            if y is not None:   # three component arrays
                t = torch.stack([_as_tensor(c, device).reshape(-1) for c in (x, y, z)], dim=1)
            else:
                t = _as_tensor(x, device).detach()
                if t.dim() != 2 or t.shape[1] != 3:
                    raise RuntimeError("Vector3f: expected an (N, 3) tensor")   # enoki: reference_cast_error
            super().__init__(t.to(device=device, dtype=torch.float32).contiguous())

    Float32.__name__, Vector3f.__name__ = "Float32", "Vector3f"
    return Float32, Vector3f


def set_requires_gradient(arr: ArrayBase, value: bool = True):
    arr._requires_grad = bool(value)


def requires_gradient(arr: ArrayBase) -> bool:
    return bool(arr._requires_grad)


def set_gradient(arr: ArrayBase, grad):
    arr._grad = grad if isinstance(grad, ArrayBase) else type(arr)(grad)


def gradient(arr: ArrayBase):
    if arr._grad is None:
        return type(arr)(torch.zeros_like(arr.torch()))
    return arr._grad


def cuda_malloc_trim():
    """enoki returns its pooled device memory; here nothing is pooled outside torch's allocator."""


def cuda_sync():
    torch.cuda.synchronize()


def cuda_eval():
    """enoki flushes its JIT queue; the kernels here are launched eagerly."""
