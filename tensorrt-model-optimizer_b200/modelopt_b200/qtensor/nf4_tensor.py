"""NF4QTensor (mirror of modelopt/torch/quantization/qtensor/nf4_tensor.py); implementation in packed.py."""

from .packed import NF4QTensor  # noqa: F401

__all__ = ["NF4QTensor"]
