"""CPU: the C-ABI library builds, loads and exports every symbol include/drrt_b200.h declares."""
import ctypes
import os
import re

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_symbols():
    src = open(os.path.join(ROOT, "include", "drrt_b200.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(drrt_[a-z0-9_]+)\s*\(", src)))


def test_header_symbols_exported(cuda_lib):
    names = declared_symbols()
    assert len(names) >= 24
    missing = [n for n in names if not hasattr(cuda_lib, n)]
    assert not missing, missing


def test_python_prototypes_cover_header(cuda_lib):
    from adjointnonlinearraytracing_b200 import _capi
    assert set(declared_symbols()) == set(_capi.PROTOTYPES)
    assert cuda_lib.drrt_abi_version() == 2


def test_library_is_sm100a_only():
    from adjointnonlinearraytracing_b200 import _capi
    import subprocess
    out = subprocess.run(["cuobjdump", "-lelf", _capi.LIB_PATH], capture_output=True, text=True).stdout
    archs = set(re.findall(r"sm_\d+a?", out))
    assert archs == {"sm_100a"}, archs


def test_no_compute_without_gpu_is_loud():
    """ops refuse CPU tensors instead of silently computing on the host."""
    import pytest
    import torch
    from adjointnonlinearraytracing_b200 import ops
    rif = torch.ones(4, 4, 4)
    rays = torch.zeros(8, 3)
    with pytest.raises(RuntimeError, match="CUDA"):
        ops.trace(rif, rays, rays, 0.25, 0.125)


def test_argument_validation_of_the_epilogue_entry_points(cuda_lib):
    """Bad arguments are rejected before any CUDA call (so this runs without a GPU), with a message."""
    from ctypes import c_void_p
    import ctypes as C
    v3 = (C.c_float * 3)(0, 0, 1)
    null = c_void_p(0)
    buf = c_void_p(0x1000)  # never dereferenced: validation fails first
    lib = cuda_lib

    def err():
        return lib.drrt_last_error().decode()

    assert lib.drrt_sensor_splat(buf, buf, null, 1.0, 8, 3, v3, v3, v3, v3, 0, 1.0, buf, 0, null) == -1
    assert "res and span" in err()
    assert lib.drrt_sensor_splat(buf, buf, null, 1.0, 8, 7, v3, v3, v3, v3, 8, 1.0, buf, 0, null) == -1
    assert "layout" in err()
    assert lib.drrt_sensor_splat(buf, buf, null, 1.0, 8, 3, null, v3, v3, v3, 8, 1.0, buf, 0, null) == -1
    assert lib.drrt_sensor_splat_backward(buf, buf, null, 1.0, 8, 3, v3, v3, v3, v3, 8, 1.0, null, buf, buf, null, 0, null) == -1
    assert "dimage" in err()
    assert lib.drrt_sensor_gather(buf, buf, 8, 3, v3, v3, v3, v3, null, 8, 1.0, buf, 0, null) == -1
    assert lib.drrt_sensor_gather(buf, buf, -1, 3, v3, v3, v3, v3, buf, 8, 1.0, buf, 0, null) == -1
    assert lib.drrt_source_rand(6, null, 0, buf, 1, 4, 4, 1, 1.0, 0.0, 90.0, 0, null, 0.0, buf, buf, buf, null, null, null, null) == -1
    assert "kind" in err()
    assert lib.drrt_source_rand(0, null, 0, buf, 1, 0, 4, 1, 1.0, 0.0, 90.0, 0, null, 0.0, buf, buf, buf, null, null, null,
                                null) == -1
    assert lib.drrt_plane_source_rand(null, 0, null, 1, 4, 4, 1, 1.0, 0.0, 0, null, 0.0, buf, buf, buf, null, null) == -1
    assert lib.drrt_plane_source_rand(null, 0, null, 0, 4, 4, 1, 1.0, 0.0, 0, null, 0.0, null, null, null, null, null) == 0
    assert lib.drrt_backtrace_rays(buf, null, 4, 4, 4, buf, buf, buf, buf, null, 8, 3, 0.25, 0.125, buf, null, null,
                                   null, 0, null, 0, null) == -1
    assert "dpos" in err()
    assert lib.drrt_backtrace_rays(buf, null, 4, 4, 1, buf, buf, buf, buf, null, 8, 3, 0.25, 0.0, buf, buf, buf,
                                   null, 0, null, 0, null) < 0
    assert lib.drrt_adam_step(buf, buf, buf, buf, 4, 4, 4, 0, 1e-3, 0.9, 0.999, 1e-8, 1.0, 0, null) == -1
    # sharded optimiser step: slabs must lie inside the grid, an empty slab is a no-op, 1..16 ranks
    assert lib.drrt_adam_step_slab(buf, buf, buf, buf, 4, 4, 4, 3, 2, 1, 1e-3, 0.9, 0.999, 1e-8, 1.0, 0, null) == -2
    assert "slab" in err()
    assert lib.drrt_adam_step_slab(buf, buf, buf, buf, 4, 4, 4, -1, 2, 1, 1e-3, 0.9, 0.999, 1e-8, 1.0, 0, null) == -2
    assert lib.drrt_adam_step_slab(null, null, null, null, 4, 4, 4, 2, 0, 1, 1e-3, 0.9, 0.999, 1e-8, 1.0, 0, null) == 0
    assert lib.drrt_adam_step_slab(buf, null, buf, buf, 4, 4, 4, 0, 2, 1, 1e-3, 0.9, 0.999, 1e-8, 1.0, 0, null) == -1
    tbl = (ctypes.c_void_p * 17)(*([ctypes.addressof(buf)] * 17))
    assert lib.drrt_adam_step_peers(tbl, tbl, 17, 0, buf, buf, 4, 4, 4, 0, 4, 1, 1e-3, 0.9, 0.999, 1e-8, 1.0, 0, null) == -1
    assert "ranks" in err()
    assert lib.drrt_adam_step_peers(tbl, tbl, 2, 2, buf, buf, 4, 4, 4, 0, 2, 1, 1e-3, 0.9, 0.999, 1e-8, 1.0, 0, null) == -1
    assert lib.drrt_adam_step_peers(null, tbl, 2, 0, buf, buf, 4, 4, 4, 0, 2, 1, 1e-3, 0.9, 0.999, 1e-8, 1.0, 0, null) == -1
    assert lib.drrt_adam_step_peers(tbl, tbl, 2, 0, buf, buf, 4, 4, 4, 2, 3, 1, 1e-3, 0.9, 0.999, 1e-8, 1.0, 0, null) == -2
    assert lib.drrt_ipc_export(null, buf, buf) == -1 and lib.drrt_ipc_open(null, 0, buf) == -1
    assert lib.drrt_ipc_close_all() == 0


def test_scratch_bytes_contract(cuda_lib):
    """drrt_backtrace_scratch_bytes: the int64 grid in deterministic mode, the private gradient copies of
    small grids otherwise (at most 16 copies and 2^23 floats in total), nothing for large grids."""
    DET, NOREP = 2, 1 << 19
    f = cuda_lib.drrt_backtrace_scratch_bytes
    assert f(0, 0) == 0 and f(-5, DET) == 0
    assert f(32 ** 3, 0) == 15 * 4 * 32 ** 3          # 16 copies, one of them `grad` itself
    assert f(65 ** 3, 0) == 15 * 4 * 65 ** 3
    assert f(128 ** 3, 0) == 3 * 4 * 128 ** 3         # 2^23 / 2^21 = 4 copies
    assert f(129 ** 3, 0) == 2 * 4 * 129 ** 3
    assert f(256 ** 3, 0) == 0 and f(512 ** 3, 0) == 0
    assert f(9 ** 3, 0) == 0 and f(129, 0) == 0       # tiny grids and the cable's radial profile
    assert f(32 ** 3, NOREP) == 0
    assert f(32 ** 3, DET) >= 8 * 32 ** 3 and f(32 ** 3, DET) == f(32 ** 3, DET | NOREP)
    assert f(256 ** 3, DET) >= 8 * 256 ** 3
    REP = lambda n: n << 20                            # DRRT_REPLICAS(n): the caller's count, any grid size
    assert f(256 ** 3, REP(3)) == 2 * 4 * 256 ** 3 and f(32 ** 3, REP(2)) == 4 * 32 ** 3
    assert f(256 ** 3, REP(3) | NOREP) == 0 and f(256 ** 3, REP(3) | DET) == f(256 ** 3, DET)
