"""FP8QTensor (mirror of modelopt/torch/quantization/qtensor/fp8_tensor.py); implementation in packed.py."""

from .packed import FP8QTensor  # noqa: F401

__all__ = ["FP8QTensor"]
