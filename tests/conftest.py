"""pytest configuration: registers the ``gpu`` marker and puts the product package + oracle on sys.path."""

import os
import sys

import numpy as np
import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = os.path.join(ROOT, "tensorrt-model-optimizer_b200")
for p in (ROOT, PKG):
    if p not in sys.path:
        sys.path.insert(0, p)

GOLDEN = os.path.join(ROOT, "tests", "golden")

DTYPES = {"f32": torch.float32, "f16": torch.float16, "bf16": torch.bfloat16}
_BITVIEW = {torch.bfloat16: torch.int16, torch.float16: torch.int16, torch.float8_e4m3fn: torch.uint8}


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


def pytest_collection_modifyitems(config, items):
    if torch.cuda.is_available():
        return
    skip = pytest.mark.skip(reason="no CUDA device")
    for item in items:
        if "gpu" in item.keywords:
            item.add_marker(skip)


def load_golden(name):
    return np.load(os.path.join(GOLDEN, name))


def from_np(a, dtype):
    """Inverse of oracle.gen_golden.to_np: integer bit patterns -> low-precision float tensor."""
    t = torch.from_numpy(np.ascontiguousarray(a))
    if dtype in _BITVIEW:
        return t.view(dtype)
    return t.to(dtype)


def bits_equal(a: torch.Tensor, b: torch.Tensor) -> bool:
    """Value equality with NaN == NaN and -0 == +0 (dequantized outputs); shapes must match."""
    if a.shape != b.shape or a.dtype != b.dtype:
        return False
    af, bf = a.double(), b.double()
    return bool(torch.all((af == bf) | (torch.isnan(af) & torch.isnan(bf))))


def n_mismatch(a: torch.Tensor, b: torch.Tensor) -> int:
    af, bf = a.double(), b.double()
    return int((~((af == bf) | (torch.isnan(af) & torch.isnan(bf)))).sum())
