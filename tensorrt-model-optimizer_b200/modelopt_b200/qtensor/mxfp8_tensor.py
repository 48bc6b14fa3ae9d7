"""MXFP8QTensor (mirror of modelopt/torch/quantization/qtensor/mxfp8_tensor.py); implementation in packed.py."""

from .packed import MXFP8QTensor  # noqa: F401

__all__ = ["MXFP8QTensor"]
