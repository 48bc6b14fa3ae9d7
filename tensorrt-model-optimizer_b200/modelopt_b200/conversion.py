"""Model conversion: insert quantizers, apply a quant_cfg (interface of modelopt/torch/quantization/conversion.py:
``replace_quant_module`` :214, ``set_quantizer_by_cfg`` :245, ``set_quantizer_attributes_full/partial`` :367-500,
``quantizer_state`` :205)."""

from __future__ import annotations

import contextlib
import fnmatch
import re
from collections.abc import Callable
from typing import Any

from torch import nn

from .config import QuantizerAttributeConfig, normalize_quant_cfg_list
from .nn.quant_module import QuantModuleRegistry
from .nn.tensor_quantizer import SequentialQuantizer, TensorQuantizer
from .utils import is_quantized

__all__ = ["replace_quant_module", "set_quantizer_by_cfg", "set_quantizer_by_cfg_context", "set_quantizer_attribute", "set_quantizer_attributes_full",
           "set_quantizer_attributes_partial", "quantizer_state", "restore_quantizer_state"]

_EXPERT_LIST_RE = re.compile(r"(weight_quantizers?|input_quantizers?)\.\d+(?=$|\.)")


def replace_quant_module(model: nn.Module, registry=QuantModuleRegistry):
    """Recursively swap every registered module type for its quantized version (in place)."""
    assert not is_quantized(model), "Model must not be quantized!"
    if type(model) in registry:
        model = registry.convert(model)

    def walk(parent):
        for name, child in parent.named_children():
            if type(child) in registry:
                setattr(parent, name, registry.convert(child))
            walk(getattr(parent, name))

    walk(model)
    n = sum(isinstance(m, TensorQuantizer) for m in model.modules())
    print(f"Inserted {n} quantizers")
    return model


_PATTERN_CACHE: dict = {}


def _compiled(pattern: str):
    rx = _PATTERN_CACHE.get(pattern)
    if rx is None:
        rx = _PATTERN_CACHE[pattern] = re.compile(fnmatch.translate(pattern))
    return rx


def _match(pattern: str | Callable, name: str, module, parent_class, model, parents=None, norm=None) -> bool:
    if not isinstance(module, (TensorQuantizer, SequentialQuantizer)):
        return False
    if isinstance(pattern, str):
        rx = _compiled(pattern)  # fnmatch semantics (case-sensitive on POSIX), compiled once per pattern
        if rx.match(name) is None:
            if norm is None:
                norm = _EXPERT_LIST_RE.sub(lambda m: m.group(1).removesuffix("s"), name)
            if norm == name or rx.match(norm) is None:
                return False
    elif callable(pattern):
        if not pattern(name):
            return False
    else:
        raise NotImplementedError(f"Unsupported type {type(pattern)}")
    if parent_class is None:
        return True
    pname = name.rpartition(".")[0]
    parent = parents[pname] if parents is not None else model.get_submodule(pname)
    return isinstance(parent, parent_class)


class _QuantizerIndex:
    """(name, quantizer) pairs of a model, listed ONCE per `set_quantizer_by_cfg` instead of one `named_modules()` walk of
    the whole model per config entry (a Llama-3-8B has 1100 modules and a preset about 25 entries).  Entries are
    refreshed in place when a quantizer object is swapped (TensorQuantizer <-> SequentialQuantizer)."""

    def __init__(self, model):
        self.model = model
        self.parents = dict(model.named_modules())
        self.items = [(n, m) for n, m in self.parents.items() if isinstance(m, (TensorQuantizer, SequentialQuantizer))]
        # expert-list names ("weight_quantizers.3" -> "weight_quantizer") normalised once, not once per config entry
        self.norm = {n: _EXPERT_LIST_RE.sub(lambda m: m.group(1).removesuffix("s"), n) for n, _ in self.items}

    def replace(self, name, module):
        parent = self.parents[name.rpartition(".")[0]]
        setattr(parent, name.rpartition(".")[2], module)
        fresh = _QuantizerIndex(self.model)  # the children of the old / new object changed too; rare, so just re-list
        self.parents, self.items, self.norm = fresh.parents, fresh.items, fresh.norm


def set_quantizer_attributes_full(model, pattern, attributes, parent_class=None, _index=None):
    """Replace ALL attributes of the matched quantizers (unspecified fields revert to their defaults); a list of
    configs turns the quantizer into a SequentialQuantizer."""
    if not isinstance(attributes, (QuantizerAttributeConfig, list)):
        raise ValueError(f"Invalid type for attributes: {type(attributes)}")
    idx = _index if _index is not None else _QuantizerIndex(model)
    for name, module in list(idx.items):  # snapshot: quantizers created by a replacement below are not visited
        if idx.parents.get(name) is not module or not _match(pattern, name, module, parent_class, model, idx.parents, idx.norm.get(name)):
            continue
        if isinstance(attributes, list):
            if not isinstance(module, SequentialQuantizer) or len(module) != len(attributes):
                module = SequentialQuantizer(*(TensorQuantizer() for _ in attributes))
                idx.replace(name, module)
            module.set_from_attribute_config(attributes)
        else:
            if isinstance(module, SequentialQuantizer):
                module = TensorQuantizer()
                idx.replace(name, module)
            module.set_from_attribute_config(attributes)


def set_quantizer_attributes_partial(model, pattern, attributes: dict[str, Any], parent_class=None, _index=None):
    """Update only the given attributes of the matched quantizers."""
    idx = _index if _index is not None else _QuantizerIndex(model)
    for name, module in idx.items:
        if _match(pattern, name, module, parent_class, model, idx.parents, idx.norm.get(name)):
            targets = module if isinstance(module, SequentialQuantizer) else [module]
            for q in targets:
                q.set_from_attribute_config(attributes)


def set_quantizer_attribute(model, pattern, attribute, parent_class=None):
    """Legacy name: a dict updates partially, a QuantizerAttributeConfig / list replaces fully."""
    if isinstance(attribute, dict):
        set_quantizer_attributes_partial(model, pattern, attribute, parent_class)
    else:
        set_quantizer_attributes_full(model, pattern, attribute, parent_class)


def set_quantizer_by_cfg(model: nn.Module, quant_cfg):
    """Apply an ordered quant_cfg list: later entries override earlier ones; an entry with ``cfg`` is a complete
    replacement, an entry with only ``enable`` toggles."""
    index = _QuantizerIndex(model)
    for entry in normalize_quant_cfg_list(quant_cfg):
        name, cfg, enable = entry["quantizer_name"], entry["cfg"], entry["enable"]
        pc = entry.get("parent_class")
        if pc:
            try:
                parent_class = QuantModuleRegistry[pc]
            except KeyError:
                raise ValueError(f"parent_class {pc!r} not found in QuantModuleRegistry.") from None
        else:
            parent_class = None
        if cfg is None:
            set_quantizer_attributes_partial(model, name, {"enable": enable}, parent_class, index)
        elif isinstance(cfg, QuantizerAttributeConfig):
            set_quantizer_attributes_full(model, name, cfg.model_copy(update={"enable": enable}), parent_class, index)
        elif isinstance(cfg, dict):
            set_quantizer_attributes_full(model, name, QuantizerAttributeConfig(**cfg, enable=enable), parent_class, index)
        else:
            attrs = [c.model_copy(update={"enable": enable}) if isinstance(c, QuantizerAttributeConfig)
                     else QuantizerAttributeConfig(**c, enable=enable) for c in cfg]
            set_quantizer_attributes_full(model, name, attrs, parent_class, index)


@contextlib.contextmanager
def set_quantizer_by_cfg_context(model: nn.Module, quant_cfg):
    """conversion.py:507-599: apply ``quant_cfg`` for the duration of the context, then put every TensorQuantizer's
    properties (not its buffers or calibrator objects) back."""
    quant_cfg = normalize_quant_cfg_list(quant_cfg)
    for entry in quant_cfg:
        if isinstance(entry.get("cfg"), list):
            raise ValueError("Sequential cfg lists are not allowed in set_quantizer_by_cfg_context. "
                             "Use only single-dict cfg entries.")
    saved = {n: m.get_modelopt_state(properties_only=True) for n, m in model.named_modules()
             if isinstance(m, TensorQuantizer)}
    try:
        set_quantizer_by_cfg(model, quant_cfg)
        yield
    finally:
        for n, m in model.named_modules():
            if isinstance(m, TensorQuantizer) and n in saved:
                m.set_from_modelopt_state(saved[n], properties_only=True)


def quantizer_state(model: nn.Module) -> dict[str, Any]:
    """name -> TensorQuantizer.get_modelopt_state() (what a checkpoint needs beside the state_dict)."""
    return {n: m.get_modelopt_state() for n, m in model.named_modules()
            if isinstance(m, (TensorQuantizer, SequentialQuantizer))}


def restore_quantizer_state(model: nn.Module, state: dict[str, Any]):
    """Inverse of ``quantizer_state`` on a model that went through ``replace_quant_module``."""
    for n, m in model.named_modules():
        if isinstance(m, TensorQuantizer) and n in state:
            m.set_from_modelopt_state(state[n])
    return model
