"""INT4QTensor (mirror of modelopt/torch/quantization/qtensor/int4_tensor.py); implementation in packed.py."""

from .packed import INT4QTensor  # noqa: F401

__all__ = ["INT4QTensor"]
