"""modelopt_b200 — B200 (sm_100a) native implementation of ModelOpt's quantization hot path.

Mirror of the reference interface for this path (same names, argument meaning and error behaviour as
``modelopt.torch.quantization``): functionals in :mod:`modelopt_b200.tensor_quant`, calibrators in
:mod:`modelopt_b200.calib`, ``TensorQuantizer`` / quant modules in :mod:`modelopt_b200.nn`, the NVFP4
packer in :mod:`modelopt_b200.qtensor`, ``quantize(model, cfg, forward_loop)`` in
:mod:`modelopt_b200.model_quant`.  Everything computes in ``lib/libmodelopt_b200.so`` (hand-written
CUDA behind a C ABI, include/modelopt_b200.h); there is no Triton, no eager and no CPU fallback.
"""

from . import _lib, ops  # noqa: F401

__version__ = "0.1.0"
