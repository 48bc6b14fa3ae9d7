"""Real-quantized tensors (mirror of modelopt.torch.quantization.qtensor)."""

from .nvfp4_tensor import *  # noqa: F401,F403
from .packed import *  # noqa: F401,F403
