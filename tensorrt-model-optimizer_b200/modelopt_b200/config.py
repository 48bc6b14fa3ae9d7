"""Quantization configs: the attribute set of a TensorQuantizer and the presets BASELINE.json names.

Mirrors modelopt/torch/quantization/config.py (``QuantizerAttributeConfig`` :322-700, presets :1681-1780,
``normalize_quant_cfg_list`` :1447-1560) without pydantic: a config is a validated plain mapping.  The preset
dictionaries have the reference's shape — ``{"quant_cfg": [ordered entries], "algorithm": ...}`` with entries
``{"quantizer_name": wildcard, "enable": bool, "cfg": {...}, "parent_class": "nn.X"}`` — so a user config
written for the reference is accepted unchanged (including the legacy flat-dict form).
"""

from __future__ import annotations

import copy
import warnings
from collections.abc import Mapping, Sequence
from typing import Any

__all__ = ["QuantizerAttributeConfig", "normalize_quant_cfg_list", "INT8_DEFAULT_CFG", "INT8_SMOOTHQUANT_CFG",
           "NVFP4_W4A4_WEIGHT_LOCAL_HESSIAN_CFG", "NVFP4_W4A4_WEIGHT_MSE_FP8_SWEEP_CFG",
           "MXFP8_DEFAULT_CFG", "MXFP6_DEFAULT_CFG", "MXFP4_DEFAULT_CFG", "MXINT8_DEFAULT_CFG", "W4A8_MXFP4_FP8_CFG",
           "FP8_DEFAULT_CFG", "NVFP4_DEFAULT_CFG", "INT4_AWQ_CFG", "INT4_BLOCKWISE_WEIGHT_ONLY_CFG",
           "NVFP4_AWQ_LITE_CFG", "W4A8_AWQ_BETA_CFG", "NVFP4_AWQ_CLIP_CFG", "NVFP4_AWQ_FULL_CFG", "choices"]

_DEFAULTS: dict[str, Any] = {
    "enable": True,
    "num_bits": 8,
    "effective_bits": None,
    "axis": None,
    "fake_quant": True,
    "unsigned": False,
    "narrow_range": False,
    "learn_amax": False,
    "type": "static",
    "block_sizes": None,
    "bias": None,
    "trt_high_precision_dtype": "Float",
    "calibrator": "max",
    "rotate": False,
    "pass_through_bwd": True,
    "backend": None,
    "backend_extra_args": None,
    "use_constant_amax": False,
    "constant_amax": None,
}


class QuantizerAttributeConfig(Mapping):
    """Validated attribute set of one TensorQuantizer (config.py:322-700).  Behaves as a read-only mapping over
    ALL fields (unspecified ones hold their defaults, as ``model_dump()`` of the reference's pydantic model)."""

    model_fields = _DEFAULTS  # name kept: TensorQuantizer.set_from_attribute_config checks against it

    def __init__(self, **kwargs):
        unknown = set(kwargs) - set(_DEFAULTS)
        if unknown:
            raise ValueError(f"{sorted(unknown)} are not valid `TensorQuantizer` attributes")
        d = copy.deepcopy(_DEFAULTS)
        d.update(kwargs)
        self._d = d
        self._validate()

    def _validate(self):
        d = self._d
        nb = d["num_bits"]
        if isinstance(nb, list):
            d["num_bits"] = nb = tuple(nb)
        if isinstance(nb, int):
            if nb < 1:
                raise ValueError("num_bits must be a positive integer or a tuple of positive integers.")
        elif isinstance(nb, tuple):
            if len(nb) != 2 or not all(isinstance(x, int) and x >= 0 for x in nb):
                raise ValueError("num_bits must be a positive integer or a tuple of positive integers.")
            if d["block_sizes"] is None and nb not in ((4, 3),):
                raise ValueError(f"Supported FPx quantization formats without block scaling: (4, 3); got {nb}")
        else:
            raise ValueError(f"Invalid num_bits {nb!r}")
        if d["type"] not in ("static", "dynamic"):
            raise ValueError(f"Invalid quantizer type {d['type']!r}")
        if isinstance(d["axis"], list):
            d["axis"] = tuple(d["axis"])
        bs = d["block_sizes"]
        if bs is not None:
            if d["axis"] is not None:
                raise ValueError("axis must be None when block_sizes is not None.")
            bs = dict(bs)
            for k in ("scale_bits",):
                if isinstance(bs.get(k), list):
                    bs[k] = tuple(bs[k])
            for k, v in bs.items():
                if isinstance(k, str):
                    if k not in ("type", "scale_bits", "scale_block_sizes", "four_over_six"):
                        raise ValueError(f"Invalid block_sizes key {k!r}")
                    if k == "type" and v not in ("static", "dynamic"):
                        raise ValueError("block_sizes['type'] must be 'static' or 'dynamic'")
                elif not (isinstance(k, int) and (v is None or (isinstance(v, int) and v > 0))):
                    raise ValueError(f"Invalid block_sizes entry {k!r}: {v!r}")
            d["block_sizes"] = bs
        if d["constant_amax"] is not None and not d["constant_amax"] > 0:
            raise ValueError("constant_amax must be positive")

    @staticmethod
    def _get_block_quant_axes_and_sizes(block_sizes):
        if block_sizes is None:
            return None
        return {k: v for k, v in block_sizes.items()
                if k not in ("type", "scale_bits", "scale_block_sizes", "four_over_six")}

    # Mapping protocol
    def __getitem__(self, k):
        return self._d[k]

    def __iter__(self):
        return iter(self._d)

    def __len__(self):
        return len(self._d)

    def __getattr__(self, k):
        try:
            return self.__dict__["_d"][k]
        except KeyError:
            raise AttributeError(k) from None

    def model_copy(self, update=None):
        return QuantizerAttributeConfig(**{**self._d, **(update or {})})

    def model_dump(self):
        return copy.deepcopy(self._d)

    def __repr__(self):
        changed = {k: v for k, v in self._d.items() if v != _DEFAULTS[k]}
        return f"QuantizerAttributeConfig({changed})"


# ------------------------------------------------------------------------------------------------
def normalize_quant_cfg_list(v) -> list[dict]:
    """config.py:1447-1560: any accepted quant_cfg form -> ordered list of explicit entries
    ``{"quantizer_name", "cfg" (None | dict | list), "enable" (bool), "parent_class" (optional)}``."""
    if isinstance(v, Mapping):
        warnings.warn("Passing quant_cfg in the legacy dict format is deprecated; use the list-of-dicts format with "
                      "explicit 'quantizer_name' keys.", DeprecationWarning, stacklevel=3)
        v = [{k: val} for k, val in v.items()]
    elif not isinstance(v, Sequence) or isinstance(v, (str, bytes)):
        raise ValueError(f"quant_cfg must be a sequence of entries (or a legacy flat mapping), got {type(v).__name__}")

    def from_legacy(key, value, parent=None):
        if key == "default":
            key = "*"
        if isinstance(key, str) and key.startswith("nn.") and parent is None:
            if not isinstance(value, Mapping):
                raise ValueError(f"Entry {key!r} must map quantizer names to configs")
            out = []
            for k2, v2 in value.items():
                out += from_legacy(k2, v2, parent=key)
            return out
        e: dict[str, Any] = {"quantizer_name": key}
        if parent:
            e["parent_class"] = parent
        if isinstance(value, (list, tuple)):
            e["cfg"] = [dict(x) for x in value]
        elif isinstance(value, (Mapping, QuantizerAttributeConfig)):
            value = dict(value)
            if set(value) == {"enable"}:
                e["enable"] = value["enable"]
            else:
                if "enable" in value:
                    e["enable"] = value.pop("enable")
                e["cfg"] = value
        else:
            raise ValueError(f"Unrecognised quant_cfg entry {key!r}: {value!r}")
        return [e]

    out = []
    for entry in v:
        if not isinstance(entry, Mapping):
            raise ValueError(f"quant_cfg entries must be dicts, got {entry!r}")
        if "quantizer_name" in entry:
            entries = [dict(entry)]
        elif len(entry) == 1:
            (k, val), = entry.items()
            entries = from_legacy(k, val)
        else:
            raise ValueError(f"Unrecognised quant_cfg entry {entry!r}")
        for e in entries:
            cfg, enable = e.get("cfg"), e.get("enable")
            if cfg is None and enable is None:
                raise ValueError(f"Entry {e!r} must specify 'cfg', 'enable', or both")
            if enable is None:
                enable = True
            if cfg is not None and enable:
                items = cfg if isinstance(cfg, (list, tuple)) else [cfg]
                if not items or any(not (isinstance(c, QuantizerAttributeConfig) or (isinstance(c, Mapping) and c))
                                    for c in items):
                    raise ValueError(f"'cfg' of enabled entry {e['quantizer_name']!r} must be a non-empty dict or a "
                                     "non-empty list of non-empty dicts")
            n = {"quantizer_name": e["quantizer_name"], "cfg": cfg, "enable": bool(enable)}
            if e.get("parent_class"):
                n["parent_class"] = e["parent_class"]
            out.append(n)
    return out


# ------------------------------------------------------------------------------------------------ presets
# modelopt_recipes/configs/ptq/units/default_disabled_quantizers.yaml
_DISABLED_NAMES = ["*block_sparse_moe.gate*", "*linear_attn.conv1d*", "*linear_attn.in_proj_a*",
                   "*linear_attn.in_proj_b*", "*lm_head*", "*mixer.conv1d*", "*mlp.gate.*",
                   "*mlp.shared_expert_gate.*", "*output_layer*", "*proj_out.*", "*router*", "mtp.*", "output.*",
                   "*embed_vision*", "*vision_tower*", "*visual*", "*vision_model*", "*multi_modal_projector*"]
_DISABLED_PARENTS = ["nn.BatchNorm1d", "nn.BatchNorm2d", "nn.BatchNorm3d", "nn.LeakyReLU", "nn.Embedding"]


def _preset(weight, inputs, algorithm):
    q: list[dict[str, Any]] = [{"quantizer_name": "*", "enable": False}]
    for name, c in (("*weight_quantizer", weight), ("*input_quantizer", inputs)):
        q.append({"quantizer_name": name, "enable": False} if c is None else {"quantizer_name": name, "cfg": c})
    q += [{"quantizer_name": n, "enable": False} for n in _DISABLED_NAMES]
    q += [{"quantizer_name": "*", "parent_class": p, "enable": False} for p in _DISABLED_PARENTS]
    return {"quant_cfg": q, "algorithm": algorithm}


_NVFP4 = {"num_bits": (2, 1), "block_sizes": {-1: 16, "type": "dynamic", "scale_bits": (4, 3)}, "effective_bits": 4.5}
_INT4_BLK = {"num_bits": 4, "block_sizes": {-1: 128, "type": "static"}}

INT8_DEFAULT_CFG = _preset({"num_bits": 8, "axis": 0}, {"num_bits": 8, "axis": None}, "max")
INT8_SMOOTHQUANT_CFG = _preset({"num_bits": 8, "axis": 0}, {"num_bits": 8, "axis": None}, "smoothquant")
FP8_DEFAULT_CFG = _preset({"num_bits": (4, 3), "axis": None}, {"num_bits": (4, 3), "axis": None}, "max")
NVFP4_DEFAULT_CFG = _preset(copy.deepcopy(_NVFP4), copy.deepcopy(_NVFP4), "max")
INT4_BLOCKWISE_WEIGHT_ONLY_CFG = _preset(copy.deepcopy(_INT4_BLK), None, "max")
INT4_AWQ_CFG = _preset(copy.deepcopy(_INT4_BLK), None, {"method": "awq_lite", "alpha_step": 0.1})
NVFP4_AWQ_LITE_CFG = _preset(copy.deepcopy(_NVFP4), copy.deepcopy(_NVFP4), "awq_lite")
W4A8_AWQ_BETA_CFG = _preset([copy.deepcopy(_INT4_BLK), {"num_bits": (4, 3)}], {"num_bits": (4, 3)}, "awq_lite")
NVFP4_AWQ_CLIP_CFG = _preset(copy.deepcopy(_NVFP4), copy.deepcopy(_NVFP4), {"method": "awq_clip"})
NVFP4_AWQ_FULL_CFG = _preset(copy.deepcopy(_NVFP4), copy.deepcopy(_NVFP4), {"method": "awq_full", "alpha_step": 0.1})



def _mx(num_bits):
    # modelopt_recipes/configs/numerics/mx{fp8,fp6,fp4,int8}.yaml
    return {"num_bits": num_bits, "block_sizes": {-1: 32, "type": "dynamic", "scale_bits": (8, 0)}}


# modelopt_recipes/configs/ptq/presets/model/mx*.yaml, w4a8_mxfp4_fp8.yaml (config.py:1708-1714): algorithm is None
MXFP8_DEFAULT_CFG = _preset(_mx((4, 3)), _mx((4, 3)), None)
MXFP6_DEFAULT_CFG = _preset(_mx((3, 2)), _mx((3, 2)), None)
MXFP4_DEFAULT_CFG = _preset(_mx((2, 1)), _mx((2, 1)), None)
MXINT8_DEFAULT_CFG = _preset(_mx(8), _mx(8), None)
W4A8_MXFP4_FP8_CFG = _preset(_mx((2, 1)), {"num_bits": (4, 3), "axis": None}, None)

# modelopt_recipes/configs/ptq/presets/model/nvfp4_w4a4_weight_local_hessian.yaml, nvfp4_w4a4_weight_mse_fp8_sweep.yaml
_NVFP4_STATIC_W = {"num_bits": (2, 1), "block_sizes": {-1: 16, "type": "static", "scale_bits": (4, 3)}}
NVFP4_W4A4_WEIGHT_LOCAL_HESSIAN_CFG = _preset(copy.deepcopy(_NVFP4_STATIC_W), copy.deepcopy(_NVFP4),
                                              {"method": "local_hessian", "fp8_scale_sweep": True})
NVFP4_W4A4_WEIGHT_MSE_FP8_SWEEP_CFG = _preset(copy.deepcopy(_NVFP4_STATIC_W), copy.deepcopy(_NVFP4),
                                              {"method": "mse", "fp8_scale_sweep": True})

choices = {"NVFP4_W4A4_WEIGHT_LOCAL_HESSIAN_CFG", "NVFP4_W4A4_WEIGHT_MSE_FP8_SWEEP_CFG", "MXFP8_DEFAULT_CFG", "MXFP6_DEFAULT_CFG", "MXFP4_DEFAULT_CFG", "MXINT8_DEFAULT_CFG", "W4A8_MXFP4_FP8_CFG",
           "INT8_DEFAULT_CFG", "INT8_SMOOTHQUANT_CFG", "FP8_DEFAULT_CFG", "NVFP4_DEFAULT_CFG", "INT4_AWQ_CFG",
           "INT4_BLOCKWISE_WEIGHT_ONLY_CFG", "NVFP4_AWQ_LITE_CFG", "W4A8_AWQ_BETA_CFG", "NVFP4_AWQ_CLIP_CFG",
           "NVFP4_AWQ_FULL_CFG"}
