"""INT8QTensor (mirror of modelopt/torch/quantization/qtensor/int8_tensor.py); implementation in packed.py."""

from .packed import INT8QTensor  # noqa: F401

__all__ = ["INT8QTensor"]
