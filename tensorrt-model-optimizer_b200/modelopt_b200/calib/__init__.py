"""Calibrators (mirror of modelopt.torch.quantization.calib)."""

from .calibrator import *  # noqa: F401,F403
from .histogram import *  # noqa: F401,F403
from .max import *  # noqa: F401,F403
from .mse import *  # noqa: F401,F403
