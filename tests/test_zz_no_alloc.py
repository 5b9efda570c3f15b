"""drrt_backtrace_noalloc (include/drrt_b200.h): the adjoint for callers that cannot allocate -- its scratch is a static
device array inside the library.  No counterpart in the reference, so the
oracle here is the existing entry point drrt_backtrace, called through the same C ABI with a caller-owned scratch buffer.

Collected last on purpose: a failure here hides no other test.  Passes alone in a
fresh process and after the other tests: it shares no state with them except the loaded library.

Tolerances are the ones the other GPU tests use: deterministic mode bit-identical, float atomics rel L2 <= 1e-5.
"""
import ctypes
from ctypes import c_void_p

import numpy as np
import pytest

from conftest import rel_l2

DET = 2          # DRRT_DETERMINISTIC
NOIL = 1 << 5    # DRRT_NO_INTERLEAVE (what the no-alloc path forces)
MAX_VOXELS = 2146689


def test_noalloc_rejects_grids_beyond_the_builtin_scratch(cuda_lib):
    """No GPU needed: the size check comes before anything touches the device."""
    lib = cuda_lib
    buf = (ctypes.c_float * 16)()
    p = ctypes.cast(buf, c_void_p)
    assert lib.drrt_backtrace_noalloc(p, 130, 130, 130, p, p, p, p, 4, 3, 0.1, 0.05, p, None, 0, None) == -1
    msg = lib.drrt_last_error().decode()
    assert "DRRT_NOALLOC_MAX_VOXELS" in msg and str(MAX_VOXELS) in msg and "2197000" in msg
    assert lib.drrt_backtrace_noalloc(p, 0, 4, 4, p, p, p, p, 4, 3, 0.1, 0.05, p, None, 0, None) == -2
    from adjointnonlinearraytracing_b200 import _capi
    assert _capi.DRRT_NOALLOC_MAX_VOXELS == MAX_VOXELS == 129 ** 3


@pytest.mark.gpu
@pytest.mark.parametrize("ng,s", [(33, 96), (65, 64)])
def test_noalloc_matches_the_existing_entry_point_twice_in_a_row(ng, s, cuda_lib):
    import torch
    from adjointnonlinearraytracing_b200 import ops, scenes as sc
    lib = cuda_lib
    T = lambda a: torch.from_numpy(np.ascontiguousarray(a)).cuda()
    P = lambda t: c_void_p(t.data_ptr())
    rif = T(sc.field_noise(ng, seed=2))
    h = float(sc.grid_spacing(ng)); ds = h / 2
    pos, vel = sc.rays_cone(s, spread=0.12, backoff=1.37 * ds)
    xt, vt = ops.trace(rif, T(pos), T(vel), h, ds)
    n = xt.shape[0]
    G = ng ** 3
    seeds = [tuple(T(a) for a in sc.seeds_normal(n, seed=k)) for k in (1, 2)]
    stream = c_void_p(torch.cuda.current_stream().cuda_stream)

    def existing(dx, dv, flags):
        nb = lib.drrt_backtrace_scratch_bytes(G, flags)
        scratch = torch.empty(max(nb, 1), dtype=torch.uint8, device="cuda")
        out = torch.full((ng, ng, ng), 7.0, device="cuda")
        rc = lib.drrt_backtrace(P(rif), ng, ng, ng, P(xt), P(vt), P(dx), P(dv), n, 3, h, ds, P(out),
                                P(scratch) if nb else c_void_p(0), nb, c_void_p(0), flags, stream)
        assert rc == 0, lib.drrt_last_error()
        return out

    def noalloc(dx, dv, flags, out):
        rc = lib.drrt_backtrace_noalloc(P(rif), ng, ng, ng, P(xt), P(vt), P(dx), P(dv), n, 3, h, ds, P(out),
                                        c_void_p(0), flags, stream)
        assert rc == 0, lib.drrt_last_error()
        return out

    for flags in (DET, 0):
        refs = [existing(dx, dv, flags | NOIL) for dx, dv in seeds]
        out = torch.full((ng, ng, ng), -3.0, device="cuda")
        # two calls in a row, DIFFERENT seeds: the second finds the first call's sums in the shared scratch array
        for k, (dx, dv) in enumerate(seeds):
            if k == 1:  # the call itself must not allocate: free device memory is the same before and after
                torch.cuda.synchronize()
                free0 = torch.cuda.mem_get_info()[0]
            got = noalloc(dx, dv, flags, out)
            torch.cuda.synchronize()
            if k == 1:
                assert torch.cuda.mem_get_info()[0] == free0
            if flags & DET:
                assert torch.equal(got, refs[k]), (flags, k)
            else:
                assert rel_l2(got.cpu().numpy(), refs[k].cpu().numpy()) <= 1e-5, (flags, k)
        # and the same call again gives the same answer (nothing left over from the call before)
        again = noalloc(*seeds[1], flags, torch.empty_like(out))
        torch.cuda.synchronize()
        if flags & DET:
            assert torch.equal(again, refs[1])
        else:
            assert rel_l2(again.cpu().numpy(), refs[1].cpu().numpy()) <= 1e-5
    # interleave requests are ignored on this path (their estimate allocates): same result as without
    il = noalloc(*seeds[0], DET | (1 << 18) | (8 << 12), torch.empty_like(out))
    torch.cuda.synchronize()
    assert torch.equal(il, existing(*seeds[0], DET | NOIL))
