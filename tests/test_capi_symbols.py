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
    assert cuda_lib.drrt_abi_version() == 1


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
