"""modelopt_b200 — B200 (sm_100a) native implementation of ModelOpt's quantization hot path.

Mirror of the reference interface for this path (same names, argument meaning and error behaviour as
``modelopt.torch.quantization``)::

    import modelopt_b200 as mtq
    model = mtq.quantize(model, mtq.NVFP4_DEFAULT_CFG, forward_loop)

functionals in :mod:`modelopt_b200.tensor_quant`, calibrators in :mod:`modelopt_b200.calib`, ``TensorQuantizer`` /
quant modules in :mod:`modelopt_b200.nn`, the NVFP4 packer in :mod:`modelopt_b200.qtensor`, calibration algorithms in
:mod:`modelopt_b200.model_calib`.  Everything computes in ``lib/libmodelopt_b200.so`` (hand-written CUDA behind a
C ABI, include/modelopt_b200.h); there is no Triton, no eager and no CPU fallback.
"""

from . import _lib, ops  # noqa: F401
from . import calib, config, nn, tensor_quant, utils  # noqa: F401
from .config import *  # noqa: F401,F403
from .conversion import *  # noqa: F401,F403
from .model_calib import *  # noqa: F401,F403
from .model_quant import *  # noqa: F401,F403
from .backends import *  # noqa: F401,F403
from .compress import *  # noqa: F401,F403
from . import gptq  # noqa: F401  (module: gptq.gptq(model, forward_loop, ...) / algorithm "gptq")
from .nn import (QuantModuleRegistry, TensorQuantizer, is_registered_quant_backend, register_quant_backend,  # noqa: F401
                 unregister_quant_backend)

__version__ = "0.2.0"
