"""Real-quantized tensors (mirror of modelopt.torch.quantization.qtensor for NVFP4)."""

from .nvfp4_tensor import *  # noqa: F401,F403
