"""MXFP4QTensor (mirror of modelopt/torch/quantization/qtensor/mxfp4_tensor.py); implementation in packed.py."""

from .packed import MXFP4QTensor  # noqa: F401

__all__ = ["MXFP4QTensor"]
