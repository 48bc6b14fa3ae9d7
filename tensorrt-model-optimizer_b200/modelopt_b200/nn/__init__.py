"""Quantized modules (mirror of modelopt.torch.quantization.nn)."""

from .tensor_quantizer import *  # noqa: F401,F403
from .quant_module import *  # noqa: F401,F403
