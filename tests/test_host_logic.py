"""CPU: host-side logic of the API mirror — config normalisation, wildcard matching, module conversion, quantizer
attribute handling — exercised without any compute call (no GPU)."""

import copy
import warnings

import pytest
import torch
from torch import nn


@pytest.fixture(scope="module")
def mtq():
    import modelopt_b200

    return modelopt_b200


def tiny():
    return nn.Sequential(nn.Linear(32, 64), nn.LayerNorm(64), nn.ReLU(), nn.Linear(64, 32), nn.Embedding(4, 4))


def test_presets_have_the_reference_shape(mtq):
    for name in mtq.config.choices:
        cfg = getattr(mtq, name)
        assert set(cfg) == {"quant_cfg", "algorithm"}
        entries = mtq.normalize_quant_cfg_list(cfg["quant_cfg"])
        assert entries[0] == {"quantizer_name": "*", "cfg": None, "enable": False}
        assert {"quantizer_name": "*lm_head*", "cfg": None, "enable": False} in entries
    w = mtq.NVFP4_DEFAULT_CFG["quant_cfg"][1]["cfg"]
    assert w["num_bits"] == (2, 1) and w["block_sizes"] == {-1: 16, "type": "dynamic", "scale_bits": (4, 3)}
    assert mtq.INT4_AWQ_CFG["algorithm"] == {"method": "awq_lite", "alpha_step": 0.1}


def test_legacy_flat_dict_is_accepted(mtq):
    with warnings.catch_warnings():
        warnings.simplefilter("ignore", DeprecationWarning)
        e = mtq.normalize_quant_cfg_list({"*weight_quantizer": {"num_bits": 8, "axis": 0},
                                          "*input_quantizer": {"enable": False},
                                          "nn.BatchNorm1d": {"*": {"enable": False}}, "default": {"enable": False}})
    assert e[0] == {"quantizer_name": "*weight_quantizer", "cfg": {"num_bits": 8, "axis": 0}, "enable": True}
    assert e[1] == {"quantizer_name": "*input_quantizer", "cfg": None, "enable": False}
    assert e[2]["parent_class"] == "nn.BatchNorm1d" and e[3]["quantizer_name"] == "*"
    with pytest.raises(ValueError):
        mtq.normalize_quant_cfg_list([{"quantizer_name": "*"}])
    with pytest.raises(ValueError):
        mtq.QuantizerAttributeConfig(num_bits=8, bogus=1)
    with pytest.raises(ValueError):
        mtq.QuantizerAttributeConfig(num_bits=4, axis=0, block_sizes={-1: 128})


def test_conversion_inserts_quantizers_and_applies_cfg_in_order(mtq, capsys):
    m = tiny()
    sd = copy.deepcopy(m.state_dict())
    mtq.replace_quant_module(m)
    assert "Inserted 9 quantizers" in capsys.readouterr().out
    assert set(m.state_dict()) == set(sd)  # the original state_dict layout is untouched
    mtq.set_quantizer_by_cfg(m, mtq.INT8_DEFAULT_CFG["quant_cfg"])
    lin, ln, emb = m[0], m[1], m[4]
    assert lin.weight_quantizer.is_enabled and lin.weight_quantizer.axis == 0 and lin.weight_quantizer.num_bits == 8
    assert lin.input_quantizer.is_enabled and lin.input_quantizer.axis is None
    assert not lin.output_quantizer.is_enabled
    assert ln.input_quantizer.is_enabled            # "*input_quantizer" matches LayerNorm's too (GPT-2 config 1)
    assert not emb.weight_quantizer.is_enabled      # parent_class nn.Embedding rule
    # an enable-only entry keeps the attributes; a cfg entry is a full replacement
    mtq.set_quantizer_by_cfg(m, [{"quantizer_name": "0.weight_quantizer", "enable": False}])
    assert not lin.weight_quantizer.is_enabled and lin.weight_quantizer.axis == 0
    mtq.set_quantizer_by_cfg(m, [{"quantizer_name": "0.weight_quantizer", "cfg": {"num_bits": 4}}])
    assert lin.weight_quantizer.is_enabled and lin.weight_quantizer.axis is None and lin.weight_quantizer.num_bits == 4
    with pytest.raises(ValueError):
        mtq.set_quantizer_by_cfg(m, [{"quantizer_name": "*", "parent_class": "nn.Nope", "enable": False}])


def test_sequential_quantizer_from_list_cfg(mtq):
    m = nn.Sequential(nn.Linear(8, 8))
    mtq.replace_quant_module(m)
    mtq.set_quantizer_by_cfg(m, [{"quantizer_name": "*weight_quantizer",
                                  "cfg": [{"num_bits": 4, "block_sizes": {-1: 128}}, {"num_bits": (4, 3)}]}])
    q = m[0].weight_quantizer
    assert isinstance(q, mtq.nn.SequentialQuantizer) and len(q) == 2 and q[1].num_bits == (4, 3)
    with mtq.nn.SequentialQuantizer.convert_to_single_quantizer(m):
        assert isinstance(m[0].weight_quantizer, mtq.TensorQuantizer)
    assert isinstance(m[0].weight_quantizer, mtq.nn.SequentialQuantizer)


def test_tensor_quantizer_attributes_and_state(mtq):
    q = mtq.TensorQuantizer(mtq.QuantizerAttributeConfig(num_bits=(2, 1),
                                                         block_sizes={-1: 16, "type": "dynamic", "scale_bits": (4, 3)}))
    assert q.is_nvfp4_dynamic and not q.is_static_block_quant and q.maxbound == 6.0 and not q._dynamic
    q2 = mtq.TensorQuantizer(mtq.QuantizerAttributeConfig(num_bits=4, block_sizes={-1: 128, "type": "static"}))
    assert q2.is_static_block_quant and q2.maxbound == 7
    q2.amax = torch.ones(4, 1)
    with pytest.raises(RuntimeError, match="Changing shape"):
        q2.amax = torch.ones(5, 1)
    st = q2.get_modelopt_state()
    assert st["_num_bits"] == 4 and st["_pytorch_state_metadata"]["buffers"]["_amax"]["shape"] == (4, 1)
    q3 = mtq.TensorQuantizer()
    q3.set_from_modelopt_state(st)
    assert q3.num_bits == 4 and q3._amax.shape == (4, 1) and q3.block_sizes == {-1: 128, "type": "static"}
    q3.disable()
    x = torch.randn(3, 5)
    assert q3(x) is x  # disabled quantizer: identity, no compute (CPU tensor would raise otherwise)
    q4 = mtq.TensorQuantizer(mtq.QuantizerAttributeConfig(num_bits=(4, 3), block_sizes={-1: 32, "type": "dynamic",
                                                                                             "scale_bits": (8, 0)}))
    assert q4.is_mx_format and q4._pass_through_bwd and q4.amax is None  # tensor_quantizer.py:289-290, :360
    q4._real_quantize  # noqa: B018  (exists; MXFP8 / MXFP4 / INT4 / INT8 / FP8 / NF4 packers are GPU-tested)
    assert q4._is_real_quantize_support()


def test_block_quant_reshape_bookkeeping(mtq):
    """_setup_for_blockquant / _process_for_blockquant / _reset_to_original_shape with a stub backend."""
    mtq.register_quant_backend("ident", lambda x, tq: x)
    q = mtq.TensorQuantizer(mtq.QuantizerAttributeConfig(num_bits=4, block_sizes={-1: 128, "type": "static"},
                                                         backend="ident"))
    x = torch.randn(6, 200)  # 200 is not a multiple of 128: padded to 256, sliced back
    y = q(x)
    assert torch.equal(y, x) and q._block_reshape_size == torch.Size([-1, 128]) and q._padding == (0, 56)
    assert q._amax_shape_for_export == (6, -1) and q.axis == (0,)
    with pytest.raises(ValueError, match="Input shape has changed"):
        q(torch.randn(6, 300))
    mtq.unregister_quant_backend("ident")


def test_cpu_tensors_fail_loudly_through_the_api(mtq):
    m = nn.Sequential(nn.Linear(16, 16))
    mtq.replace_quant_module(m)
    mtq.set_quantizer_by_cfg(m, mtq.FP8_DEFAULT_CFG["quant_cfg"])
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        m(torch.randn(2, 16))


def test_library_ops_registered_with_fake_impls():
    """tensor_quant.py:216-276: the functionals are torch.library operators that trace without running a kernel."""
    from torch._subclasses import FakeTensorMode

    from modelopt_b200 import tensor_quant as tq

    assert tq.OP_NAMESPACE in ("tensorrt", "modelopt_b200")
    ns = getattr(torch.ops, tq.OP_NAMESPACE)
    with FakeTensorMode():
        x = torch.empty(4, 32, dtype=torch.bfloat16, device="cuda")
        amax = torch.empty((), device="cuda")
        assert ns.quantize_op(x, amax, 8, 0, False, True).shape == x.shape
        assert ns.quantize_op(x, amax, 8, 4, False, False).dtype == torch.bfloat16
        assert ns.dynamic_block_quantize_op(x, 16, amax, 4, 2, 8, 4).shape == x.shape
    schema = str(ns.quantize_op.default._schema)
    assert "Tensor input, Tensor amax, int num_bits, int exponent_bits, bool unsigned, bool narrow_range" in schema


def test_int_cast_ste():
    from modelopt_b200.tensor_quant import int_cast_ste

    x = torch.tensor([-9.6, -8.0, 0.4, 0.5, 1.5, 7.7, 100.0], requires_grad=True)
    y = int_cast_ste(x, 4, False, False)
    assert y.tolist() == [-8.0, -8.0, 0.0, 0.0, 2.0, 7.0, 7.0]
    y.sum().backward()
    assert x.grad.tolist() == [0.0, 1.0, 1.0, 1.0, 1.0, 0.0, 0.0]
    assert int_cast_ste(torch.tensor([-3.0, 20.0]), 4, True, True).tolist() == [0.0, 15.0]
    assert int_cast_ste(torch.tensor([-9.0]), 4, False, True).tolist() == [-7.0]


def test_scale_map_layouts(mtq):
    """ops._scale_map: where each element's scale lives for per-tensor / per-axis / N-D block layouts (host logic of
    mo_q8_pack / mo_q8_unpack): dims = {O, A, I, bo, ba, bi, so, sa, si}."""
    from modelopt_b200 import ops

    assert ops._scale_map([256, 512], [1, 1]) == [1, 1, 256 * 512, 1, 1, 256 * 512, 1, 1, 1]          # per tensor
    assert ops._scale_map([256, 512], [256, 1]) == [1, 256, 512, 1, 1, 512, 256, 1, 1]                # per row
    assert ops._scale_map([256, 512], [1, 512]) == [1, 256, 512, 1, 256, 1, 512, 512, 1]              # per column
    assert ops._scale_map([256, 512], [2, 4]) == [1, 256, 512, 1, 128, 128, 8, 4, 1]                  # 128 x 128 blocks
    assert ops._scale_map([4, 256, 512], [4, 2, 4]) == [4, 256, 512, 1, 128, 128, 8, 4, 1]            # per-expert blocks
    with pytest.raises(AssertionError):
        ops._scale_map([256, 512], [3, 1])                                                            # does not tile
    with pytest.raises(NotImplementedError):
        ops._scale_map([4, 6, 8, 10], [2, 6, 2, 10])                                                  # 4 index groups


def test_gemm_registry_and_presets(mtq):
    from modelopt_b200 import config as qcfg
    from modelopt_b200.backends import GEMMRegistry

    r = GEMMRegistry()
    f1, f2 = (lambda *a: 1), (lambda *a: 2)
    r.register(f1, lambda m, x, a, k: x == "one")
    r.register(f2, None)
    assert r.find_match(None, "one") is f1 and r.find_match(None, "other") is f2 and f1 in r
    r.unregister(f2)
    assert r.find_match(None, "other") is None and f2 not in r
    for name in ("MXFP8_DEFAULT_CFG", "MXFP6_DEFAULT_CFG", "MXFP4_DEFAULT_CFG", "MXINT8_DEFAULT_CFG", "W4A8_MXFP4_FP8_CFG",
                 "NVFP4_W4A4_WEIGHT_LOCAL_HESSIAN_CFG", "NVFP4_W4A4_WEIGHT_MSE_FP8_SWEEP_CFG"):
        cfg = getattr(qcfg, name)
        assert name in qcfg.choices and isinstance(cfg["quant_cfg"], list)
    mx = next(e for e in qcfg.MXFP8_DEFAULT_CFG["quant_cfg"] if e.get("quantizer_name") == "*weight_quantizer")["cfg"]
    assert mx["num_bits"] == (4, 3) and mx["block_sizes"] == {-1: 32, "type": "dynamic", "scale_bits": (8, 0)}
    assert qcfg.MXFP8_DEFAULT_CFG["algorithm"] is None
    assert qcfg.NVFP4_W4A4_WEIGHT_LOCAL_HESSIAN_CFG["algorithm"] == {"method": "local_hessian", "fp8_scale_sweep": True}
    from modelopt_b200.model_quant import _algorithms

    assert {"max", "mse", "local_hessian", "gptq", "smoothquant", "awq_lite", "awq_clip", "awq_full"} <= set(_algorithms())
