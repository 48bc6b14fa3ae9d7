"""CPU: the C-ABI library loads and exports every symbol include/modelopt_b200.h declares."""

import ctypes
import os
import re

from conftest import ROOT


def _declared():
    src = open(os.path.join(ROOT, "include", "modelopt_b200.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(mo_[a-z0-9_]+)\s*\(", src)))


def test_library_exports_every_declared_symbol():
    from modelopt_b200 import _lib

    lib = _lib.load()
    names = _declared()
    assert len(names) >= 15
    for n in names:
        assert hasattr(lib, n), f"{n} declared in the header but not exported"
    assert set(_lib.SIGNATURES) | {"mo_last_error"} == set(names), "ctypes table and header disagree"
    assert lib.mo_abi_version() == _lib.ABI_VERSION


def test_argument_errors_without_gpu():
    """Argument validation happens before any CUDA call, so it is checkable on a CPU box."""
    from modelopt_b200 import _lib

    lib = _lib.load()
    assert lib.mo_qdq_int(None, None, 16, 2, None, 1, 16, 8, 0, 0, None, None) == -4  # MO_ERR_NULL
    buf = ctypes.create_string_buffer(64)
    p = ctypes.addressof(buf)
    assert lib.mo_qdq_int(p, p, 16, 2, p, 1, 16, 40, 0, 0, None, None) == -3  # MO_ERR_ARG (num_bits)
    assert b"num_bits" in lib.mo_last_error()
    assert lib.mo_qdq_nvfp4_dynamic(p, p, 1, 16, 2, p, 8, 0, None) == -3  # block not in {16,32,64}
    assert lib.mo_amax(p, 1, 0, 16, 2, p, None) == -2  # MO_ERR_SHAPE
    assert lib.mo_qdq_fp8(p, p, 0, 2, p, 1, 0, 0, None) == 0  # empty tensor is a no-op


def test_cpu_tensor_is_rejected_loudly():
    import pytest
    import torch

    from modelopt_b200 import ops

    with pytest.raises(RuntimeError, match="no CPU fallback"):
        ops.qdq_fp8(torch.randn(4, 16), torch.tensor(1.0))
