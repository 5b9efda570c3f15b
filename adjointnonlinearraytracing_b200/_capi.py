"""ctypes binding of ``libdrrt_b200.so`` (C ABI: ``include/drrt_b200.h``).

There is no fallback: if the CUDA library is missing, importing the compute entry
points raises.  Build it with ``python -c "import __graft_entry__ as g; g.build()"``
or ``make -C adjointnonlinearraytracing_b200/csrc``.
"""
from __future__ import annotations

import ctypes
import os
from ctypes import POINTER, c_char_p, c_float, c_int, c_int64, c_size_t, c_uint, c_uint8, c_void_p

_HERE = os.path.dirname(os.path.abspath(__file__))
# DRRT_B200_LIB: development override (tuning builds of the same library); unset in production
LIB_PATH = os.environ.get("DRRT_B200_LIB") or os.path.join(_HERE, "lib", "libdrrt_b200.so")

DRRT_SOA, DRRT_AOS = 1, 3
DRRT_REFERENCE_ARITH = 1 << 0
DRRT_DETERMINISTIC = 1 << 1
DRRT_ACCUMULATE = 1 << 2
DRRT_FAST_ARITH = 1 << 3
DRRT_WARP_AGGREGATION = 1 << 4
DRRT_NO_INTERLEAVE = 1 << 5
DRRT_SENSOR_FAR = 1 << 6
DRRT_NO_FACE_CARRY = 1 << 7
DRRT_AUTO_INTERLEAVE = 1 << 18
DRRT_NO_REPLICAS = 1 << 19
DRRT_FACE_CARRY = 1 << 24
DRRT_PAIR_RAYS = 1 << 25
DRRT_NO_PAIR_RAYS = 1 << 26
DRRT_SMEM_GRID = 1 << 27
DRRT_AGG_RUN_WARP = 15 << 8


def DRRT_REPLICAS(n: int) -> int:
    return int(n) << 20


def DRRT_AGG_RUN(n: int) -> int:
    return int(n) << 8


def DRRT_INTERLEAVE(n: int) -> int:
    return int(n) << 12

DRRT_UNIQUE_ID_BYTES = 128
DRRT_IPC_HANDLE_BYTES = 64
DRRT_MAX_PEERS = 16
DRRT_NOALLOC_MAX_VOXELS = 2146689  # 129^3: largest grid of drrt_backtrace_noalloc

# every symbol include/drrt_b200.h declares (checked by tests/test_capi_symbols.py)
_F, _I, _L, _U, _P = c_float, c_int, c_int64, c_uint, c_void_p
PROTOTYPES = {
    "drrt_abi_version": (c_int, []),
    "drrt_last_error": (c_char_p, []),
    "drrt_launch_count": (c_int64, []),
    "drrt_trace": (c_int, [_P, _I, _I, _I, _P, _P, _L, _I, _F, _F, _P, _P, _P, _U, _P]),
    "drrt_trace_plane": (c_int, [_P, _I, _I, _I, _P, _P, _P, _P, _L, _I, _F, _F, _P, _P, _P, _P, _U, _P]),
    "drrt_trace_target": (c_int, [_P, _I, _I, _I, _P, _P, _P, _L, _I, _F, _F, _P, _P, _P, _P, _U, _P]),
    "drrt_trace_sdf": (c_int, [_P, _P, _I, _I, _I, _P, _P, _L, _I, _F, _F, _P, _P, _P, _U, _P]),
    "drrt_trace_cable": (c_int, [_P, _I, _F, _F, _P, _P, _P, _L, _I, _F, _P, _P, _P, _P, _U, _P]),
    "drrt_backtrace_scratch_bytes": (c_size_t, [_L, _U]),
    "drrt_backtrace": (c_int, [_P, _I, _I, _I, _P, _P, _P, _P, _L, _I, _F, _F, _P, _P, c_size_t, _P, _U, _P]),
    "drrt_backtrace_noalloc": (c_int, [_P, _I, _I, _I, _P, _P, _P, _P, _L, _I, _F, _F, _P, _P, _U, _P]),
    "drrt_backtrace_sdf": (c_int, [_P, _P, _I, _I, _I, _P, _P, _P, _P, _L, _I, _F, _F, _P, _P, c_size_t, _P, _U, _P]),
    "drrt_backtrace_cable": (c_int, [_P, _I, _F, _F, _P, _P, _P, _P, _L, _I, _F, _P, _P, c_size_t, _P, _U, _P]),
    "drrt_backtrace_rays": (c_int, [_P, _P, _I, _I, _I, _P, _P, _P, _P, _P, _L, _I, _F, _F, _P, _P, _P, _P, c_size_t, _P, _U, _P]),
    "drrt_volume_eval": (c_int, [_P, _I, _I, _I, _P, _L, _I, _F, _P, _P, _P, _U, _P]),
    "drrt_volume_splat": (c_int, [_P, _I, _I, _I, _P, _P, _P, _L, _I, _F, _P]),
    "drrt_cylinder_eval": (c_int, [_P, _I, _F, _F, _P, _L, _I, _P, _P, _P, _P]),
    "drrt_cylinder_splat": (c_int, [_P, _I, _F, _F, _P, _P, _P, _L, _I, _P]),
    "drrt_adam_step": (c_int, [_P, _P, _P, _P, _I, _I, _I, _I, _F, _F, _F, _F, _F, _U, _P]),
    "drrt_adam_step_slab": (c_int, [_P, _P, _P, _P, _I, _I, _I, _I, _I, _I, _F, _F, _F, _F, _F, _U, _P]),
    "drrt_upres_volume": (c_int, [_P, _I, _I, _I, _P, _I, _I, _I, _P]),
    "drrt_sensor_splat": (c_int, [_P, _P, _P, _F, _L, _I, _P, _P, _P, _P, _I, _F, _P, _U, _P]),
    "drrt_sensor_splat_backward": (c_int, [_P, _P, _P, _F, _L, _I, _P, _P, _P, _P, _I, _F, _P, _P, _P, _P, _U, _P]),
    "drrt_plane_source_rand": (c_int, [_P, ctypes.c_uint64, _P, _I, _I, _I, _I, _F, _F, _I, _P, _F, _P, _P, _P, _P, _P]),
    "drrt_sensor_gather": (c_int, [_P, _P, _L, _I, _P, _P, _P, _P, _P, _I, _F, _P, _U, _P]),
    "drrt_sensor_gather_backward": (c_int, [_P, _P, _L, _I, _P, _P, _P, _P, _P, _I, _F, _P, _P, _P, _U, _P]),
    "drrt_source_rand": (c_int, [_I, _P, ctypes.c_uint64, _P, _I, _I, _I, _I, _F, _F, _F, _I, _P, _F, _P, _P, _P, _P, _P, _P, _P]),
    "drrt_trace_host": (c_int, [_P, _I, _I, _I, _P, _P, _L, _I, _F, _F, _P, _P, _P, _U]),
    "drrt_backtrace_host": (c_int, [_P, _I, _I, _I, _P, _P, _P, _P, _L, _I, _F, _F, _P, _P, _U]),
    "drrt_host_cache_grid": (None, [_I]),
    "drrt_host_invalidate_grid": (None, []),
    "drrt_host_release": (None, []),
    "drrt_comm_unique_id": (c_int, [_P]),
    "drrt_comm_init_rank": (c_int, [POINTER(c_void_p), _I, _P, _I]),
    "drrt_comm_destroy": (c_int, [_P]),
    "drrt_allreduce_grad": (c_int, [_P, c_size_t, _P, _P]),
    "drrt_allreduce_multimem": (c_int, [_P, c_size_t, _I, _I, _P]),
    "drrt_reduce_scatter_grad": (c_int, [_P, c_size_t, _I, _P, _P]),
    "drrt_adam_step_peers": (c_int, [_P, _P, _I, _I, _P, _P, _I, _I, _I, _I, _I, _I, _F, _F, _F, _F, _F, _U, _P]),
    "drrt_adam_step_multimem": (c_int, [_P, _P, _P, _P, _P, _I, _I, _I, _I, _I, _I, _F, _F, _F, _F, _F, _U, _P]),
    "drrt_ipc_export": (c_int, [_P, _P, _P]),
    "drrt_ipc_open": (c_int, [_P, ctypes.c_uint64, _P]),
    "drrt_ipc_close_all": (c_int, []),
    "drrt_allgather_grid": (c_int, [_P, c_size_t, _I, _P, _P]),
}

_lib = None


class DrrtError(RuntimeError):
    """Raised for any non-zero return of the C ABI (the reference raises RuntimeError
    from std::runtime_error, reference src/volume.cpp:28,37,115,124)."""


def load() -> ctypes.CDLL:
    """Load libdrrt_b200.so; fail loudly when it has not been built."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise ImportError(
                f"{LIB_PATH} is missing: the CUDA library has not been built "
                "(run __graft_entry__.build()); there is no CPU fallback")
        lib = ctypes.CDLL(LIB_PATH)
        for name, (res, args) in PROTOTYPES.items():
            fn = getattr(lib, name)
            fn.restype = res
            fn.argtypes = args
        _lib = lib
    return _lib


def check(rc: int) -> None:
    if rc != 0:
        msg = load().drrt_last_error()
        raise DrrtError((msg or b"drrt error").decode() + f" (code {rc})")


def launch_count() -> int:
    return int(load().drrt_launch_count())
