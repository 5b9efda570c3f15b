"""Streaming forward + adjoint over HOST-resident ray batches.

The reference regenerates its ray batch on the CPU every iteration and ships it to the GPU
(``core/luneburg_opt.py:53-59``: ``torch.rand`` + ``.to(device)``), then reads the gradient back
for the optimiser.  At 4 M rays that is 100 MB in and 64 MiB out per iteration -- about 3 ms of
PCIe time next to ~31 ms of kernels.  ``HostStepPipeline`` double-buffers the step so the copies
of batch k+1 (host->device) and of gradient k-1 (device->host) overlap the kernels of batch k:

    copy-in stream :  H2D rays[k+1]        |  H2D rays[k+2]      | ...
    compute stream :  trace+backtrace[k]   |  trace+backtrace[k+1]
    copy-out stream:  D2H grad[k-1]        |  D2H grad[k]

Everything on the device is still the C ABI (``ops.trace`` / ``ops.backtrace`` / the NCCL
all-reduce); this module only sequences streams and events.

``GraphedStep`` is the other end of the size range: at the reference's small configurations
(32^3 / 64^3 grids, 65 k - 262 k rays) the two kernels run for 30-300 us, and the step is captured
once as a CUDA graph and replayed so that the host's launch path (Python marshalling + allocator +
~6 stream operations) cannot become the bound when the host is busy.  Measured on an idle host the
gain is 1-2 % (profiles/r09_small_graph.log): even these steps are bound by the kernels.
"""
from __future__ import annotations

from typing import Callable, Optional, Tuple

import torch

from . import ops


class HostStepPipeline:
    """Depth-2 software pipeline: ``submit(pos_pinned, vel_pinned)`` enqueues one forward + adjoint
    step and returns its index k; ``result(k)`` blocks until the gradient grid of step k has landed
    in pinned host memory and returns ``(grad_host, stats_host)`` (views valid until step k+2 is
    submitted).

    ``seed_fn(xt, vt) -> (dx, dv)`` produces the adjoint seeds on the device from the forward
    outputs (the loss gradient); the default is ``dx = dv = 1`` (reference ``src/test.cpp:142-144``).
    """

    DEPTH = 2

    def __init__(self, rif: torch.Tensor, h: float, ds: float, n_rays: int, flags: int = 0,
                 allreduce: Optional[Callable[[torch.Tensor], torch.Tensor]] = None,
                 seed_fn: Optional[Callable[[torch.Tensor, torch.Tensor], Tuple[torch.Tensor, torch.Tensor]]] = None,
                 host_planes: Optional[Tuple[int, int]] = None):
        """``host_planes = (i0, count)``: copy only planes ``[i0, i0 + count)`` of the gradient grid to the host.  After
        the all-reduce every rank holds the SAME summed gradient, so with one slab per rank (``dist.shard_bounds`` over
        ``rif.shape[0]``) the job as a whole still lands the full gradient in host memory every step while each rank's
        PCIe link carries 1/world of it; ``result()`` then returns that slab."""
        if not rif.is_cuda:
            raise RuntimeError("rif must be a CUDA tensor (there is no CPU path)")
        self.rif, self.h, self.ds, self.flags = rif, float(h), float(ds), int(flags)
        self.allreduce, self.seed_fn = allreduce, seed_fn
        dev = rif.device
        self.dev = dev
        self.s_in, self.s_out = torch.cuda.Stream(dev), torch.cuda.Stream(dev)
        D = self.DEPTH
        self.pos = [torch.empty(n_rays, 3, device=dev) for _ in range(D)]
        self.vel = [torch.empty(n_rays, 3, device=dev) for _ in range(D)]
        alloc = getattr(allreduce, "alloc", None)  # a collective may need the gradient grid in its own buffers
        self.grad = [alloc(rif.shape, dev) if alloc is not None else torch.empty_like(rif) for _ in range(D)]
        self.stats = [ops.new_stats(dev) for _ in range(D)]
        self.host_planes = (0, int(rif.shape[0])) if host_planes is None else (int(host_planes[0]), int(host_planes[1]))
        if not (0 <= self.host_planes[0] and self.host_planes[0] + self.host_planes[1] <= rif.shape[0]):
            raise RuntimeError("host_planes: slab outside the grid")
        hshape = (self.host_planes[1],) + tuple(rif.shape[1:])
        self.grad_host = [torch.empty(hshape, dtype=torch.float32).pin_memory() for _ in range(D)]
        self.stats_host = [torch.empty(2, dtype=torch.int64).pin_memory() for _ in range(D)]
        self.ones = torch.ones(n_rays, 3, device=dev) if seed_fn is None else None
        self.ev_in = [torch.cuda.Event() for _ in range(D)]
        self.ev_compute = [torch.cuda.Event() for _ in range(D)]
        self.ev_out = [torch.cuda.Event() for _ in range(D)]
        self.k = 0
        self.h2d_bytes = 2 * n_rays * 3 * 4
        self.d2h_bytes = self.grad_host[0].numel() * 4 + 16

    def submit(self, pos_host: torch.Tensor, vel_host: torch.Tensor) -> int:
        k, b = self.k, self.k % self.DEPTH
        compute = torch.cuda.current_stream(self.dev)
        # copy-in: the device ray buffers of slot b are free once step k-2's kernels are done
        with torch.cuda.stream(self.s_in):
            if k >= self.DEPTH:
                self.s_in.wait_event(self.ev_compute[b])
            self.pos[b].copy_(pos_host, non_blocking=True)
            self.vel[b].copy_(vel_host, non_blocking=True)
            self.ev_in[b].record(self.s_in)
        # compute: needs the rays of this step and slot b's gradient buffer drained (step k-2 copied out)
        compute.wait_event(self.ev_in[b])
        if k >= self.DEPTH:
            compute.wait_event(self.ev_out[b])
        self.stats[b].zero_()
        xt, vt = ops.trace(self.rif, self.pos[b], self.vel[b], self.h, self.ds, flags=self.flags, stats=self.stats[b])
        dx, dv = (self.ones, self.ones) if self.seed_fn is None else self.seed_fn(xt, vt)
        ops.backtrace(self.rif, xt, vt, dx, dv, self.h, self.ds, flags=self.flags, stats=self.stats[b], out=self.grad[b])
        if self.allreduce is not None:
            self.allreduce(self.grad[b])
        self.ev_compute[b].record(compute)
        # copy-out
        with torch.cuda.stream(self.s_out):
            self.s_out.wait_event(self.ev_compute[b])
            i0, cnt = self.host_planes
            self.grad_host[b].copy_(self.grad[b][i0:i0 + cnt], non_blocking=True)
            self.stats_host[b].copy_(self.stats[b], non_blocking=True)
            self.ev_out[b].record(self.s_out)
        # keep xt/vt alive until the kernels have consumed them (caching allocator is stream-aware
        # for the compute stream they were allocated on, nothing else touches them)
        self.k += 1
        return k

    def result(self, k: int):
        if not (self.k - self.DEPTH <= k < self.k):
            raise RuntimeError(f"step {k} is not in flight (submitted {self.k}, depth {self.DEPTH})")
        b = k % self.DEPTH
        self.ev_out[b].synchronize()
        return self.grad_host[b], self.stats_host[b]

    def drain(self):
        for b in range(min(self.k, self.DEPTH)):
            self.ev_out[b].synchronize()


class GraphedStep:
    """``trace -> seed_fn -> backtrace`` over fixed device buffers, captured once as a CUDA graph.

    ``step(pos, vel)`` copies the rays into the captured input buffers (device or pinned host tensors;
    ``None`` keeps the previous batch), replays the graph and returns ``(xt, vt, grad, stats)`` --
    views of the captured output buffers, overwritten by the next replay.  ``rif`` is captured by
    address: in-place updates (``ops.adam_step_``) are seen by the next replay, a re-allocated grid
    (an up-res level) needs a new ``GraphedStep``.  ``seed_fn`` must be capturable (device-side torch
    ops only, no ``.item()``); the library's device-pointer entry points never synchronise.
    """

    def __init__(self, rif: torch.Tensor, h: float, ds: float, n_rays: int, flags: int = 0,
                 seed_fn: Optional[Callable[[torch.Tensor, torch.Tensor], Tuple[torch.Tensor, torch.Tensor]]] = None,
                 warmup: int = 2):
        if not rif.is_cuda or rif.dtype != torch.float32 or not rif.is_contiguous():
            raise RuntimeError("rif must be a contiguous CUDA float32 tensor (there is no CPU path)")
        self.rif, self.h, self.ds, self.flags, self.seed_fn = rif, float(h), float(ds), int(flags), seed_fn
        dev = rif.device
        self.dev = dev
        self.pos = torch.zeros(n_rays, 3, device=dev)
        self.vel = torch.zeros(n_rays, 3, device=dev)
        self.grad = torch.empty_like(rif)
        self.stats = ops.new_stats(dev)
        self.ones = torch.ones(n_rays, 3, device=dev) if seed_fn is None else None
        self.launches = 0
        # warm up on a side stream (module load, kernel attributes, allocator pools), then capture
        side = torch.cuda.Stream(dev)
        side.wait_stream(torch.cuda.current_stream(dev))
        with torch.cuda.stream(side):
            for _ in range(max(1, warmup)):
                self._body()
        torch.cuda.current_stream(dev).wait_stream(side)
        self.graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(self.graph):
            self.xt, self.vt = self._body()

    def _body(self):
        self.stats.zero_()
        xt, vt = ops.trace(self.rif, self.pos, self.vel, self.h, self.ds, flags=self.flags, stats=self.stats)
        dx, dv = (self.ones, self.ones) if self.seed_fn is None else self.seed_fn(xt, vt)
        ops.backtrace(self.rif, xt, vt, dx, dv, self.h, self.ds, flags=self.flags, stats=self.stats, out=self.grad)
        return xt, vt

    def step(self, pos: Optional[torch.Tensor] = None, vel: Optional[torch.Tensor] = None):
        if pos is not None:
            self.pos.copy_(pos, non_blocking=True)
        if vel is not None:
            self.vel.copy_(vel, non_blocking=True)
        self.graph.replay()
        self.launches += 1
        return self.xt, self.vt, self.grad, self.stats
