"""Multi-GPU path: rays shard across ranks, each rank accumulates a private
gradient grid, the grids are summed with ONE all-reduce over NVLink
(SURVEY.md section 8e).  One process per GPU; ``torch.distributed`` supplies
rendezvous only -- the collective itself is ``drrt_allreduce_grad`` on a NCCL
communicator this library creates (``drrt_comm_init_rank``), enqueued on the same
stream as the adjoint kernel so no host synchronisation separates them.

The reference is single-GPU (no collective anywhere in it); this module is new.
"""
from __future__ import annotations

import ctypes
from ctypes import c_void_p
from typing import Optional, Tuple

import torch
import torch.distributed as dist

from . import _capi, ops


def shard_bounds(n: int, rank: int, world: int) -> Tuple[int, int]:
    """Contiguous, balanced [start, end) block of `n` rays for `rank` (earlier ranks take the
    remainder).  Contiguous blocks keep whole sensor rows/views together, which keeps a warp's
    rays spatially coherent."""
    if world <= 0 or not (0 <= rank < world):
        raise ValueError("bad rank/world")
    base, rem = divmod(n, world)
    start = rank * base + min(rank, rem)
    return start, start + base + (1 if rank < rem else 0)


def shard_cyclic(n: int, rank: int, world: int, block: int):
    """Block-cyclic shard: rank r takes blocks r, r + world, r + 2 world, ... of `block` consecutive rays (the last block
    may be short).  Returns a list of (start, end) ranges.  Contiguous blocks give every rank one REGION of the sensor,
    and regions differ in cost (rays through the middle of the volume bend more and change cells on more axes: measured
    5.73 ms for the slowest of 4 ranks against 5.15 ms on average on the 256^3 batch); cyclic blocks of whole sensor rows
    give every rank the same mix while a warp still holds 32 neighbouring pixels."""
    if world <= 0 or not (0 <= rank < world) or block <= 0:
        raise ValueError("bad rank/world/block")
    nblocks = (n + block - 1) // block
    return [(b * block, min(n, (b + 1) * block)) for b in range(rank, nblocks, world)]


def shard_rays(rank: int, world: int, *tensors):
    """Slice every (N, 3) / (N,) tensor to this rank's block."""
    n = tensors[0].shape[0]
    a, b = shard_bounds(n, rank, world)
    return tuple(t[a:b].contiguous() for t in tensors)


class GradAllReduce:
    """Sum-all-reduce of the private gradient grids.

    backend="nccl-native": a NCCL communicator owned by libdrrt_b200.so; the unique id is
    broadcast through the existing torch.distributed group.  backend="torch": plain
    ``torch.distributed.all_reduce`` (used by the gloo CPU tests of the host logic).
    """

    def __init__(self, backend: Optional[str] = None, group=None):
        self.group = group
        self.world = dist.get_world_size(group) if dist.is_initialized() else 1
        self.rank = dist.get_rank(group) if dist.is_initialized() else 0
        if backend is None:
            backend = "nccl-native" if (self.world > 1 and dist.get_backend(group) == "nccl") else "torch"
        self.backend = backend
        self.comm = None
        if self.world > 1 and backend == "nccl-native":
            lib = _capi.load()
            buf = ctypes.create_string_buffer(_capi.DRRT_UNIQUE_ID_BYTES)
            if self.rank == 0:
                _capi.check(lib.drrt_comm_unique_id(buf))
            box = [bytes(buf.raw)]
            dist.broadcast_object_list(box, src=0, group=group)
            uid = ctypes.create_string_buffer(box[0], _capi.DRRT_UNIQUE_ID_BYTES)
            comm = c_void_p()
            _capi.check(lib.drrt_comm_init_rank(ctypes.byref(comm), self.world, uid, self.rank))
            self.comm = comm

    def __call__(self, grad: torch.Tensor) -> torch.Tensor:
        if self.world == 1:
            return grad
        if self.backend == "nccl-native":
            if not grad.is_cuda or not grad.is_contiguous() or grad.dtype != torch.float32:
                raise RuntimeError("grad must be a contiguous float32 CUDA tensor")
            stream = c_void_p(torch.cuda.current_stream(grad.device).cuda_stream)
            _capi.check(_capi.load().drrt_allreduce_grad(c_void_p(grad.data_ptr()), grad.numel(), self.comm, stream))
        else:
            dist.all_reduce(grad, op=dist.ReduceOp.SUM, group=self.group)
        return grad

    def alloc(self, shape, device) -> torch.Tensor:
        """A gradient grid this collective can reduce (any float32 CUDA tensor; SwitchAllReduce needs its own)."""
        return torch.empty(tuple(shape), dtype=torch.float32, device=device)

    def close(self):
        if self.comm is not None:
            _capi.load().drrt_comm_destroy(self.comm)
            self.comm = None


class SwitchAllReduce:
    """The gradient sum as THIS library's kernel over NVLink SHARP instead of a NCCL call: the gradient grid lives in
    a buffer that every rank maps at one MULTICAST address, rank r sums slab r of it with ``multimem.ld_reduce`` (the
    switch fetches the P copies and returns their sum) and broadcasts the result with ``multimem.st``
    (``drrt_allreduce_multimem``, csrc/drrt_misc_kernels.cu).  Per GPU the links carry 1/P of the grid each way.

    PyTorch's symmetric memory supplies the plumbing only -- the allocation, the multicast mapping and the two
    stream-ordered barriers around the kernel; single node, needs a multicast-capable NVSwitch fabric.
    Gradient grids must come from ``alloc()`` (``ops.backtrace(..., out=buf)`` writes straight into them)."""

    def __init__(self, group=None):
        if not dist.is_initialized():
            raise RuntimeError("SwitchAllReduce needs an initialised torch.distributed process group")
        import torch.distributed._symmetric_memory as symm
        self._symm = symm
        self.group = group if group is not None else dist.group.WORLD
        self.world, self.rank = dist.get_world_size(self.group), dist.get_rank(self.group)
        self.backend = "switch-multimem"
        self._handles = {}
        enable = getattr(symm, "enable_symm_mem_for_group", None)
        if enable is not None:
            try:
                enable(self.group.group_name)
            except Exception:
                pass  # newer torch enables groups implicitly

    def alloc(self, shape, device) -> torch.Tensor:
        n = 1
        for d in shape:
            n *= int(d)
        if n % 4:
            raise RuntimeError("SwitchAllReduce: the grid must hold a multiple of 4 floats")
        if self.world == 1:  # nothing to sum: any buffer will do
            return torch.empty(tuple(int(d) for d in shape), dtype=torch.float32, device=device)
        with torch.cuda.device(device):
            buf = self._symm.empty(n, dtype=torch.float32, device=torch.device(device))
            hdl = self._symm.rendezvous(buf, self.group.group_name)
        mc = int(getattr(hdl, "multicast_ptr", 0) or 0)
        if mc == 0:
            raise RuntimeError("SwitchAllReduce: this fabric / driver exposes no multicast mapping")
        view = buf.view(*shape)
        self._handles[view.data_ptr()] = (hdl, mc, buf)
        return view

    def __call__(self, grad: torch.Tensor) -> torch.Tensor:
        if self.world == 1:
            return grad
        ent = self._handles.get(grad.data_ptr())
        if ent is None:
            raise RuntimeError("SwitchAllReduce: grad must be a buffer returned by alloc()")
        hdl, mc, buf = ent
        if grad.numel() != buf.numel() or not grad.is_contiguous():
            raise RuntimeError("SwitchAllReduce: grad must be the whole contiguous buffer returned by alloc()")
        with torch.cuda.device(grad.device):
            stream = c_void_p(torch.cuda.current_stream(grad.device).cuda_stream)
            hdl.barrier(channel=0)  # every rank's adjoint has finished writing its copy
            _capi.check(_capi.load().drrt_allreduce_multimem(c_void_p(mc), grad.numel(), self.rank, self.world, stream))
            hdl.barrier(channel=1)  # every slab has landed in every copy
        return grad

    def close(self):
        self._handles.clear()


class ShardedAdam:
    """The optimiser step of ``multires_opt`` (reference core/optimizer.py:61-68) split over the ranks by
    slabs of the slowest grid axis (the 512^3 configuration):

        reduce-scatter of the private gradient grids  ->  fused boundary-zero + Adam + clamp on this rank's slab
        ->  all-gather of the updated slabs into every rank's replica of ``n``.

    Same bytes on the wire as the all-reduce of ``GradAllReduce``, 1/world of the optimiser pass and of the Adam
    state per rank.  Needs ``n.shape[0] % world == 0`` (equal slabs); ``step`` leaves every rank with the same grid:
    the one ``ops.adam_step_`` produces from the summed gradient (the update is per voxel, so splitting it changes
    nothing; only NCCL's summation order inside reduce-scatter vs all-reduce can differ in the last bit).
    ``step_fn(n_slab, grad_slab, m_slab, v_slab, shape, i0_begin, step)`` replaces the CUDA kernel in the CPU
    (gloo) tests of this host logic.
    """

    def __init__(self, n: torch.Tensor, lr: float, betas=(0.9, 0.999), eps: float = 1e-8, clamp_min: float = 1.0,
                 allreduce: Optional[GradAllReduce] = None, step_fn=None):
        if n.dim() != 3 or not n.is_contiguous():
            raise RuntimeError("ShardedAdam: expected a contiguous 3-D grid")
        self.ar = allreduce if allreduce is not None else GradAllReduce()
        self.world, self.rank = self.ar.world, self.ar.rank
        if n.shape[0] % self.world:
            raise RuntimeError(f"ShardedAdam: {n.shape[0]} planes do not split into {self.world} equal slabs")
        self.n = n
        self.planes = n.shape[0] // self.world
        self.i0 = self.rank * self.planes
        self.lr, self.betas, self.eps, self.clamp_min = float(lr), tuple(betas), float(eps), float(clamp_min)
        self.step_count = 0
        slab = n[self.i0:self.i0 + self.planes]
        self.exp_avg = torch.zeros_like(slab)
        self.exp_avg_sq = torch.zeros_like(slab)
        self.step_fn = step_fn

    def _slab(self, t: torch.Tensor) -> torch.Tensor:
        return t[self.i0:self.i0 + self.planes]

    def step(self, grad: torch.Tensor) -> torch.Tensor:
        """``grad``: this rank's private gradient grid (overwritten: scratch after the call).  Returns ``n``."""
        if grad.shape != self.n.shape or not grad.is_contiguous():
            raise RuntimeError("Resolution doesn't match data")
        n = self.n
        count = self.planes * n.shape[1] * n.shape[2]
        native = self.world > 1 and self.ar.backend == "nccl-native"
        if native:
            stream = c_void_p(torch.cuda.current_stream(grad.device).cuda_stream)
            _capi.check(_capi.load().drrt_reduce_scatter_grad(c_void_p(grad.data_ptr()), count, self.rank,
                                                              self.ar.comm, stream))
        elif self.world > 1:  # gloo has no reduce-scatter: sum everything, keep the slab
            dist.all_reduce(grad, op=dist.ReduceOp.SUM, group=self.ar.group)
        self.step_count += 1
        with torch.no_grad():
            args = (self._slab(n.detach()), self._slab(grad), self.exp_avg, self.exp_avg_sq, tuple(n.shape), self.i0,
                    self.step_count)
            if self.step_fn is not None:
                self.step_fn(*args)
            else:
                ops.adam_step_slab_(*args, self.lr, self.betas, self.eps, clamp_min=self.clamp_min)
            if native:
                _capi.check(_capi.load().drrt_allgather_grid(c_void_p(n.data_ptr()), count, self.rank, self.ar.comm,
                                                             stream))
            elif self.world > 1:
                parts = [torch.empty_like(self._slab(n.detach())) for _ in range(self.world)]
                dist.all_gather(parts, self._slab(n.detach()).contiguous(), group=self.ar.group)
                n.detach().copy_(torch.cat(parts, dim=0))
        return n


class PeerShardedAdam:
    """``ShardedAdam`` as ONE kernel over peer memory (NVLink P2P) instead of three passes: for every voxel of this
    rank's slab ``drrt_adam_step_peers`` loads the ``world`` private gradients straight from the peers' grids, applies
    boundary-zero + Adam + clamp and stores the new value into all ``world`` replicas of ``n``.

    ``n`` and ``grad`` are this rank's replica of the grid and its private gradient buffer; both must stay the same
    buffers for the lifetime of the object (the peers map them through CUDA IPC once, here) and must come from
    ``cudaMalloc`` (PyTorch's default allocator; not expandable segments / cudaMallocAsync).  ``step()`` is ordered
    after every rank's adjoint and before every rank's next read of ``n`` by two stream-ordered barriers (a 4-byte
    NCCL all-reduce each), so nothing synchronises with the host.  Single node only.
    """

    def __init__(self, n: torch.Tensor, grad: torch.Tensor, lr: float, betas=(0.9, 0.999), eps: float = 1e-8,
                 clamp_min: float = 1.0, allreduce: Optional[GradAllReduce] = None):
        for name, t in (("n", n), ("grad", grad)):
            if not (t.is_cuda and t.dtype == torch.float32 and t.is_contiguous() and t.dim() == 3):
                raise RuntimeError(f"{name}: expected a contiguous float32 3-D CUDA tensor")
        if grad.shape != n.shape:
            raise RuntimeError("Resolution doesn't match data")
        self.ar = allreduce if allreduce is not None else GradAllReduce()
        self.world, self.rank = self.ar.world, self.ar.rank
        if self.world > 1 and self.ar.backend != "nccl-native":
            raise RuntimeError("PeerShardedAdam needs the NCCL communicator of GradAllReduce (GPU ranks of one node)")
        if self.world > _capi.DRRT_MAX_PEERS:
            raise RuntimeError(f"PeerShardedAdam: at most {_capi.DRRT_MAX_PEERS} ranks")
        if n.shape[0] % self.world:
            raise RuntimeError(f"PeerShardedAdam: {n.shape[0]} planes do not split into {self.world} equal slabs")
        self.n, self.grad = n, grad
        self.planes = n.shape[0] // self.world
        self.i0 = self.rank * self.planes
        self.lr, self.betas, self.eps, self.clamp_min = float(lr), tuple(betas), float(eps), float(clamp_min)
        self.step_count = 0
        slab_shape = (self.planes,) + tuple(n.shape[1:])
        self.exp_avg = torch.zeros(slab_shape, device=n.device)
        self.exp_avg_sq = torch.zeros(slab_shape, device=n.device)
        self._token = torch.zeros(1, device=n.device)
        lib = _capi.load()
        mine = []
        with torch.cuda.device(n.device):
            for t in ((n, grad) if self.world > 1 else ()):  # one rank: nothing to map, any allocator will do
                h = ctypes.create_string_buffer(_capi.DRRT_IPC_HANDLE_BYTES)
                off = ctypes.c_uint64(0)
                _capi.check(lib.drrt_ipc_export(c_void_p(t.data_ptr()), h, ctypes.byref(off)))
                mine.append((bytes(h.raw), int(off.value)))
            everyone = [None] * self.world
            if self.world > 1:
                dist.all_gather_object(everyone, mine, group=self.ar.group)
            else:
                everyone[0] = mine
            ptrs = [[0] * self.world, [0] * self.world]
            for r in range(self.world):
                for k in range(2):
                    if r == self.rank:
                        ptrs[k][r] = (n, grad)[k].data_ptr()
                    else:
                        hb, off = everyone[r][k]
                        out = c_void_p()
                        _capi.check(lib.drrt_ipc_open(ctypes.create_string_buffer(hb, _capi.DRRT_IPC_HANDLE_BYTES), off,
                                                      ctypes.byref(out)))
                        ptrs[k][r] = out.value
        self._n_ptrs = (c_void_p * self.world)(*ptrs[0])
        self._g_ptrs = (c_void_p * self.world)(*ptrs[1])
        if self.world > 1:
            dist.barrier(group=self.ar.group)  # every peer has mapped every buffer

    def _barrier(self):
        if self.world > 1:
            self.ar(self._token)

    def step(self) -> torch.Tensor:
        """Uses the gradient every rank has accumulated in its ``grad`` buffer.  Returns ``n`` (updated everywhere)."""
        n = self.n
        self._barrier()  # every rank's adjoint has finished writing its gradient grid
        self.step_count += 1
        with torch.cuda.device(n.device):
            stream = c_void_p(torch.cuda.current_stream(n.device).cuda_stream)
            _capi.check(_capi.load().drrt_adam_step_peers(
                self._n_ptrs, self._g_ptrs, self.world, self.rank, c_void_p(self.exp_avg.data_ptr()),
                c_void_p(self.exp_avg_sq.data_ptr()), int(n.shape[0]), int(n.shape[1]), int(n.shape[2]), self.i0,
                self.planes, self.step_count, self.lr, self.betas[0], self.betas[1], self.eps, self.clamp_min, 0, stream))
        self._barrier()  # every rank's slab has landed in every replica
        return n

    def close(self):
        """Unmap the peers' buffers (all PeerShardedAdam objects of this process)."""
        if self.world > 1:
            torch.cuda.synchronize(self.n.device)
            dist.barrier(group=self.ar.group)
            _capi.load().drrt_ipc_close_all()


class SwitchShardedAdam:
    """``PeerShardedAdam`` with the NVSwitch doing the gradient sum and the broadcast of the new slab
    (``drrt_adam_step_multimem``): one ``multimem.ld_reduce`` and one ``multimem.st`` per float4 of this rank's slab.
    Per GPU the links carry the grid once out (its gradient copy, pulled by the switch) and once in (the updated slabs).

    ``n`` and ``grad`` must be buffers of ``switch.alloc()`` (a ``SwitchAllReduce``): create the grid with
    ``n = switch.alloc(shape, device); n.copy_(n0)`` and pass ``out=grad`` to ``ops.backtrace``.  Same ordering contract
    as ``PeerShardedAdam`` (two stream barriers, here the symmetric-memory ones).  Rows of ``n.shape[2] % 4 == 0``."""

    def __init__(self, switch: "SwitchAllReduce", n: torch.Tensor, grad: torch.Tensor, lr: float, betas=(0.9, 0.999),
                 eps: float = 1e-8, clamp_min: float = 1.0):
        if grad.shape != n.shape or n.dim() != 3:
            raise RuntimeError("Resolution doesn't match data")
        if n.shape[2] % 4:
            raise RuntimeError("SwitchShardedAdam: rows must hold a multiple of 4 floats (use PeerShardedAdam otherwise)")
        en, eg = switch._handles.get(n.data_ptr()), switch._handles.get(grad.data_ptr())
        if en is None or eg is None:
            raise RuntimeError("SwitchShardedAdam: n and grad must be buffers of SwitchAllReduce.alloc()")
        self.sw, self.n, self.grad = switch, n, grad
        self.world, self.rank = switch.world, switch.rank
        if n.shape[0] % self.world:
            raise RuntimeError(f"SwitchShardedAdam: {n.shape[0]} planes do not split into {self.world} equal slabs")
        self._hdl, self._n_mc, self._g_mc = eg[0], en[1], eg[1]
        self.planes = n.shape[0] // self.world
        self.i0 = self.rank * self.planes
        self.lr, self.betas, self.eps, self.clamp_min = float(lr), tuple(betas), float(eps), float(clamp_min)
        self.step_count = 0
        slab_shape = (self.planes,) + tuple(n.shape[1:])
        self.exp_avg = torch.zeros(slab_shape, device=n.device)
        self.exp_avg_sq = torch.zeros(slab_shape, device=n.device)

    def step(self) -> torch.Tensor:
        n = self.n
        self.step_count += 1
        with torch.cuda.device(n.device):
            stream = c_void_p(torch.cuda.current_stream(n.device).cuda_stream)
            if self.world > 1:
                self._hdl.barrier(channel=0)  # every rank's adjoint has finished writing its gradient copy
            _capi.check(_capi.load().drrt_adam_step_multimem(
                c_void_p(self._n_mc), c_void_p(self._g_mc), c_void_p(n.data_ptr()), c_void_p(self.exp_avg.data_ptr()),
                c_void_p(self.exp_avg_sq.data_ptr()), int(n.shape[0]), int(n.shape[1]), int(n.shape[2]), self.i0,
                self.planes, self.step_count, self.lr, self.betas[0], self.betas[1], self.eps, self.clamp_min, 0, stream))
            if self.world > 1:
                self._hdl.barrier(channel=1)  # every slab has landed in every replica
        return n


class ShardedTracer:
    """Forward + adjoint over this rank's ray block with the all-reduced gradient.

    >>> st = ShardedTracer(rif, h, ds)            # rif replicated on every rank
    >>> xt, vt = st.trace(pos_local, vel_local)   # outputs stay sharded (per-ray loss)
    >>> grad = st.backtrace(xt, vt, dx, dv)       # identical on every rank after the sum
    """

    def __init__(self, rif: torch.Tensor, h: float, ds: float, flags: int = 0,
                 allreduce: Optional[GradAllReduce] = None, reduce: bool = True):
        # reduce=False: ``backtrace`` leaves this rank's PRIVATE gradient in ``self.grad`` -- the input of
        # ``PeerShardedAdam(rif, st.grad, lr)`` / ``ShardedAdam.step(st.grad)``, which do the sum themselves
        self.reduce = bool(reduce)
        self.rif, self.h, self.ds, self.flags = rif, float(h), float(ds), int(flags)
        self.allreduce = allreduce if allreduce is not None else GradAllReduce()
        self.grad = torch.empty_like(rif)
        self.stats = ops.new_stats(rif.device) if rif.is_cuda else None

    def trace(self, pos, vel):
        return ops.trace(self.rif, pos, vel, self.h, self.ds, flags=self.flags, stats=self.stats)

    def backtrace(self, xt, vt, dx, dv):
        g = ops.backtrace(self.rif, xt, vt, dx, dv, self.h, self.ds, flags=self.flags, stats=self.stats,
                          out=self.grad)
        return self.allreduce(g) if self.reduce else g
