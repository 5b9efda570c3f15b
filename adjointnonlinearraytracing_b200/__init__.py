"""B200-native eikonal ray tracer and adjoint (drop-in for the reference's `drrt` module).

Layout of the package:
  csrc/      hand-written sm_100a CUDA kernels + the C ABI (include/drrt_b200.h)
  _capi.py   ctypes loader of lib/libdrrt_b200.so (no fallback)
  ops.py     torch-tensor marshalling over the C ABI
  drrt.py    TracerC / TracerD / TracerS mirroring reference src/drrt.cpp
  tracer.py  torch.autograd.Function wrappers mirroring reference core/tracer.py
  dist.py    ray sharding + NCCL all-reduce of the gradient grid
  scenes.py  seeded synthetic grids / ray batches for tests and bench
"""
__version__ = "0.1.0"
