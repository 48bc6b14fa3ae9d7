"""User-facing API (interface of modelopt/torch/quantization/model_quant.py: ``quantize`` :147, ``calibrate`` :64,
``print_quant_summary`` :709, ``fold_weight`` :728, ``disable_quantizer`` / ``enable_quantizer`` :650-680)."""

from __future__ import annotations

import fnmatch
import warnings
from collections.abc import Callable
from typing import Any

from torch import nn

from . import model_calib
from .conversion import replace_quant_module, set_quantizer_attributes_partial, set_quantizer_by_cfg
from .nn.quant_module import QuantModule
from .nn.tensor_quantizer import TensorQuantizer
from .utils import is_quantized

__all__ = ["quantize", "calibrate", "print_quant_summary", "fold_weight", "disable_quantizer", "enable_quantizer",
           "postprocess_amax"]

ForwardLoop = Callable[[nn.Module], None]


def _algorithms() -> dict[str | None, Callable]:
    table: dict[str | None, Callable] = {"max": model_calib.max_calibrate}
    table["smoothquant"] = model_calib.smoothquant
    table["mse"] = model_calib.mse_calibrate
    table["local_hessian"] = model_calib.local_hessian_calibrate
    from . import gptq as _gptq

    table["gptq"] = lambda model, forward_loop=None, **kw: _gptq.gptq(model, forward_loop, **kw)
    for name in ("awq_lite", "awq_clip", "awq_full"):  # mode.py:236-238: every awq method goes through awq()
        table[name] = (lambda method: lambda model, forward_loop=None, **kw:
                       model_calib.awq(model, forward_loop, algorithm=method, **kw))(name)
    return table


def calibrate(model: nn.Module, algorithm: str | dict | None = "max", forward_loop: ForwardLoop | None = None):
    """model_quant.py:64-144.  ``algorithm``: name, ``{"method": name, **kwargs}``, or None (skip calibration)."""
    if algorithm is None:
        return model
    if isinstance(algorithm, dict):
        kwargs = dict(algorithm)
        method = kwargs.pop("method")
    else:
        method, kwargs = algorithm, {}
    table = _algorithms()
    if method not in table:
        raise ValueError(f"Unsupported calibration algorithm {method!r}; available: {sorted(k for k in table if k)}")
    was_training = model.training
    model.eval()
    table[method](model, forward_loop=forward_loop, **kwargs)
    model.train(was_training)
    _validate_after_calibration(model)
    return model


def _validate_after_calibration(model: nn.Module):
    """model_quant.py:119-122: warn on negative / inf / NaN amax or pre_quant_scale after calibration.  The reference asks
    each quantizer (three device->host reads per attribute: ~2700 syncs on a Llama-3-8B); here every attribute of a device
    is checked by ONE reduction and one read, and only a failing model pays for the per-quantizer warnings."""
    import torch

    found = [(name, module, attr) for name, module in model.named_modules() if isinstance(module, TensorQuantizer)
             for attr in ("_amax", "_pre_quant_scale")
             if isinstance(getattr(module, attr, None), torch.Tensor) and not getattr(module, attr).is_meta]
    by_dev: dict = {}
    for item in found:
        by_dev.setdefault(getattr(item[1], item[2]).device, []).append(item)
    for items in by_dev.values():
        flat = torch.cat([getattr(m, attr).detach().reshape(-1).float() for _, m, attr in items])
        # >= 0 is False for NaN too; inf is the only bad value that passes it
        if bool(torch.all((flat >= 0) & ~torch.isinf(flat))):
            continue
        for name, module, attr in items:
            module.validate_attr(attr_name=attr, warn_error=True, name=name)


def quantize(model: nn.Module, config: dict[str, Any], forward_loop: ForwardLoop | None = None) -> nn.Module:
    """Insert quantizers per ``config["quant_cfg"]`` and calibrate with ``config["algorithm"]``; in place.

    ``forward_loop(model)`` feeds the calibration data; every TensorQuantizer it reaches collects statistics with a
    CUDA kernel (no fake-quant runs during calibration, as in the reference)."""
    if not is_quantized(model):
        replace_quant_module(model)
    set_quantizer_by_cfg(model, config.get("quant_cfg", []))
    return calibrate(model, config.get("algorithm", "max"), forward_loop)


def _toggle(model, pattern, enable: bool):
    set_quantizer_attributes_partial(model, pattern, {"enable": enable})


def disable_quantizer(model: nn.Module, wildcard_or_filter_func):
    _toggle(model, wildcard_or_filter_func, False)


def enable_quantizer(model: nn.Module, wildcard_or_filter_func):
    _toggle(model, wildcard_or_filter_func, True)


def print_quant_summary(model: nn.Module, output_dir: str | None = None):
    lines = [f"{name:80} {mod}" for name, mod in model.named_modules() if isinstance(mod, TensorQuantizer)]
    lines.append(f"{len(lines)} TensorQuantizers found in model")
    text = "\n".join(lines)
    if output_dir:
        import os

        with open(os.path.join(output_dir, ".quant_summary.txt"), "w") as f:
            f.write(text)
    else:
        print(text)


def fold_weight(model: nn.Module, keep_attrs: bool = False):
    """Fold every weight quantizer into its weight (the weight becomes its fake-quantized value)."""
    for module in model.modules():
        if isinstance(module, QuantModule):
            module.fold_weight(keep_attrs)


def postprocess_amax(model: nn.Module, key: str, post_process_fn):
    """Apply ``post_process_fn`` to the amax of every quantizer whose name matches ``key``."""
    for name, module in model.named_modules():
        if isinstance(module, TensorQuantizer) and fnmatch.fnmatch(name, key) and module.amax is not None:
            module.amax = post_process_fn(module.amax)
            if not module.validate_attr(attr_name="_amax"):
                warnings.warn(f"{name}: amax is invalid after post-processing")
