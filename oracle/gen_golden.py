"""Generate golden vectors by running the REAL reference on CPU.  TEST INFRASTRUCTURE ONLY.

Run in the build container only (needs the read-only checkout at /root/reference):

    PYTHONDONTWRITEBYTECODE=1 python oracle/gen_golden.py

Imports the reference with three import shims (oracle/ref_shims: package metadata, a stub
``omegaconf`` and a stub ``pulp`` — none is called on this path) and writes
``tests/golden/*.npz``.  Nothing in tests/, smoke() or bench.py reads /root/reference at run time:
they read the committed ``.npz`` files.  bf16 / fp16 / fp8 arrays are stored as integer bit patterns.
"""

from __future__ import annotations

import os
import sys
import warnings

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
REF = os.environ.get("MODELOPT_REFERENCE", "/root/reference")
sys.path[:0] = [REF, os.path.join(HERE, "ref_shims"), os.path.dirname(HERE)]
sys.dont_write_bytecode = True
warnings.filterwarnings("ignore")

OUT = os.path.join(HERE, "..", "tests", "golden")

_INT_VIEW = {torch.bfloat16: torch.int16, torch.float16: torch.int16, torch.float8_e4m3fn: torch.uint8}


def to_np(t):
    """Tensor -> numpy, low-precision float dtypes as integer bit patterns."""
    if isinstance(t, (int, float)):
        return np.asarray(t)
    t = t.detach().cpu().contiguous()
    if t.dtype in _INT_VIEW:
        return t.view(_INT_VIEW[t.dtype]).numpy()
    return t.numpy()


DTYPES = {"f32": torch.float32, "f16": torch.float16, "bf16": torch.bfloat16}


def gen_qdq():
    from modelopt.torch.quantization.tensor_quant import fake_tensor_quant, scaled_e4m3

    g = {}
    torch.manual_seed(1234)
    base = torch.randn(48, 256) * 3.0
    base[3, 7] = 40.0  # outlier so clipping and saturation are exercised
    base[5, :16] = 0.0
    for dn, dt in DTYPES.items():
        x = base.to(dt)
        g[f"x_{dn}"] = to_np(x)
        # ---------------- INT per-tensor -------------------------------------------------
        for bits in (3, 4, 5, 7, 8, 11):
            for unsigned in (False, True):
                for narrow in (False, True):
                    xi = x.abs() if unsigned else x
                    amax = (xi.float().abs().max() * 0.7).to(dt)
                    y = fake_tensor_quant(xi, amax, None, bits, unsigned, narrow)
                    k = f"int_{dn}_b{bits}_u{int(unsigned)}_n{int(narrow)}"
                    g[k + "_amax"], g[k + "_y"] = to_np(amax.float()), to_np(y)
        # ---------------- INT per-channel / per-block ------------------------------------
        for axis in (0, 1):
            red = 1 - axis
            amax = (x.float().abs().amax(dim=red, keepdim=True) * 0.9).to(dt)
            y = fake_tensor_quant(x, amax, None, 8, False, False)
            g[f"int_{dn}_axis{axis}_amax"], g[f"int_{dn}_axis{axis}_y"] = to_np(amax.float()), to_np(y)
        xb = x.reshape(-1, 128)
        amax = xb.float().abs().amax(dim=1, keepdim=True).to(dt)
        y = fake_tensor_quant(xb, amax, None, 4, False, False)  # INT4 block-128 (INT4_AWQ_CFG)
        g[f"int_{dn}_blk128_amax"], g[f"int_{dn}_blk128_y"] = to_np(amax.float()), to_np(y)
        # tiny / zero amax conventions
        for nm, am in (("tiny", 1e-26), ("zero", 0.0), ("eps", 1.0 / (1 << 24)), ("eps2", 1.5 / (1 << 24))):
            y = fake_tensor_quant(x, torch.tensor(am), None, 8, False, False)
            g[f"int_{dn}_{nm}_y"] = to_np(y)
        # ---------------- FP8 --------------------------------------------------------------
        for nm, am in (("max", x.float().abs().max()), ("clip", x.float().abs().max() * 0.5),
                       ("zero", torch.tensor(0.0)), ("small", torch.tensor(1e-3))):
            y = scaled_e4m3(x, am.to(dt), None, 4, 3)
            g[f"fp8_{dn}_{nm}_amax"], g[f"fp8_{dn}_{nm}_y"] = to_np(am.to(dt).float()), to_np(y)
        y = scaled_e4m3(x, None, None, 4, 3)
        g[f"fp8_{dn}_none_y"] = to_np(y)
        x3 = x.reshape(6, 8, 256)
        for axis in (0, 1, 2):
            red = [i for i in range(3) if i != axis]
            amax = x3.float().abs().amax(dim=red, keepdim=True).to(dt)
            y = scaled_e4m3(x3, amax, None, 4, 3)
            g[f"fp8_{dn}_axis{axis}_amax"], g[f"fp8_{dn}_axis{axis}_y"] = to_np(amax.float()), to_np(y)
    np.savez_compressed(os.path.join(OUT, "qdq_cpu.npz"), **g)
    print("qdq_cpu.npz", len(g), "arrays")


def gen_fp4_scales():
    from modelopt.torch.kernels.quantization.gemm.fp4_kernel import compute_fp4_scales

    g = {}
    torch.manual_seed(7)
    amax = (torch.randn(4096).abs() * 4.0)
    amax[:8] = torch.tensor([0.0, 1e-8, 6.0 * 2**-11, 6.0 * 2**-10, 6.0, 448.0 * 6, 1e6, 3.3e-4])
    g["amax"] = to_np(amax)
    for nm, ga in (("auto", None), ("given", torch.tensor(17.0)), ("small", torch.tensor(1e-3))):
        for q in (True, False):
            for m in (448.0, 256.0):
                s = compute_fp4_scales(amax, ga, q, m)
                g[f"scale_{nm}_q{int(q)}_m{int(m)}"] = to_np(s)
        if ga is not None:
            g[f"ga_{nm}"] = to_np(ga)
    np.savez_compressed(os.path.join(OUT, "fp4_scales_cpu.npz"), **g)
    print("fp4_scales_cpu.npz", len(g), "arrays")


def gen_amax():
    from modelopt.torch.quantization.calib import MaxCalibrator
    from modelopt.torch.quantization.utils import reduce_amax, reduce_block_amax

    g = {}
    torch.manual_seed(99)
    for dn, dt in DTYPES.items():
        xs = [(torch.randn(6, 40, 96) * (i + 1)).to(dt) for i in range(3)]
        for i, x in enumerate(xs):
            g[f"x{i}_{dn}"] = to_np(x)
        g[f"amax_all_{dn}"] = to_np(reduce_amax(xs[0]).float())
        for axis in (0, 1, 2, -1, (0, 2)):
            an = str(axis).replace(" ", "").replace("(", "").replace(")", "").replace(",", "_")
            c = MaxCalibrator(8, axis, False)
            for x in xs:
                c.collect(x)
            g[f"calib_axis{an}_{dn}"] = to_np(c.compute_amax().float())
        c = MaxCalibrator(8, None, False)
        for x in xs:
            c.collect(x)
        g[f"calib_none_{dn}"] = to_np(c.compute_amax().float())
        xb = xs[1].reshape(-1, 96)
        g[f"blk16_{dn}"] = to_np(reduce_block_amax(xb, {-1: 16}).float())
        g[f"blk32_{dn}"] = to_np(reduce_block_amax(xb, {-1: 32}).float())
    np.savez_compressed(os.path.join(OUT, "amax_cpu.npz"), **g)
    print("amax_cpu.npz", len(g), "arrays")


def gen_nvfp4_pack():
    from modelopt.torch.quantization.qtensor.nvfp4_tensor import NVFP4QTensor

    g = {}
    torch.manual_seed(2024)
    base = torch.randn(32, 192) * 2.0
    base[0, :16] = 0.0                       # zero block: scale 0 -> 1.0
    base[1, :16] = base[1, :16] * 1e-6       # tiny block: scale clamps to 2^-9
    base[2, :8] = torch.tensor([0.25, 0.75, 1.25, 1.75, 2.5, 3.5, 5.0, 6.0])
    base[3, 5] = 300.0
    for dn, dt in DTYPES.items():
        x = base.to(dt)
        g[f"x_{dn}"] = to_np(x)
        qt, sf, sf2 = NVFP4QTensor.quantize(x.clone(), 16)
        g[f"codes_{dn}"], g[f"sf_{dn}"], g[f"sf2_{dn}"] = to_np(qt._quantized_data), to_np(sf), to_np(sf2)
        for fast in (False,):
            d = qt.dequantize(dtype=dt, scale=sf, double_scale=sf2, block_sizes={-1: 16}, fast=fast)
            g[f"deq_{dn}"] = to_np(d)
        # externally supplied global scale (the quantizer's calibrated amax)
        sf2b = torch.tensor(5.0) / (6.0 * 448.0)
        qt, sf, _ = NVFP4QTensor.quantize(x.clone(), 16, weights_scaling_factor_2=sf2b)
        g[f"codes_ext_{dn}"], g[f"sf_ext_{dn}"], g[f"sf2_ext_{dn}"] = to_np(qt._quantized_data), to_np(sf), to_np(sf2b)
        # ragged last dim (padded to the block)
        xr = x[:, :184].contiguous()
        qt, sf, sf2 = NVFP4QTensor.quantize(xr.clone(), 16)
        g[f"codes_rag_{dn}"], g[f"sf_rag_{dn}"], g[f"sf2_rag_{dn}"] = to_np(qt._quantized_data), to_np(sf), to_np(sf2)
    np.savez_compressed(os.path.join(OUT, "nvfp4_pack_cpu.npz"), **g)
    print("nvfp4_pack_cpu.npz", len(g), "arrays")


def _tiny_model(dt=torch.float32):
    from torch import nn

    torch.manual_seed(77)
    m = nn.Sequential(nn.Linear(256, 512), nn.ReLU(), nn.Linear(512, 256, bias=False)).to(dt)
    return m


def gen_quantize():
    """mtq.quantize end to end on CPU with the reference: calibrated amax of every quantizer and model outputs."""
    import copy

    import modelopt.torch.quantization as mtq

    g = {}
    gen = torch.Generator().manual_seed(5)
    batches = [torch.randn(8, 16, 256, generator=gen) * (1 + i) for i in range(3)]
    test = torch.randn(4, 16, 256, generator=gen) * 2
    base = _tiny_model()
    for k, v in base.state_dict().items():
        g[f"w_{k}"] = to_np(v)
    for i, b in enumerate(batches):
        g[f"batch{i}"] = to_np(b)
    g["test"] = to_np(test)
    for name in ("INT8_DEFAULT_CFG", "FP8_DEFAULT_CFG", "INT4_BLOCKWISE_WEIGHT_ONLY_CFG", "NVFP4_DEFAULT_CFG",
                 "INT8_SMOOTHQUANT_CFG", "INT4_AWQ_CFG"):
        m = copy.deepcopy(base)

        def loop(model):
            for b in batches:
                model(b)

        m = mtq.quantize(m, copy.deepcopy(getattr(mtq, name)), loop)
        for n, mod in m.named_modules():
            if isinstance(mod, mtq.nn.TensorQuantizer):
                g[f"{name}|{n}|enabled"] = np.asarray(mod.is_enabled)
                for buf in ("_amax", "_pre_quant_scale"):
                    if hasattr(mod, buf):
                        g[f"{name}|{n}|{buf}"] = to_np(getattr(mod, buf).float())
        for k, v in m.state_dict().items():
            if "quantizer" not in k and not torch.equal(v, base.state_dict()[k]):
                g[f"{name}|weight_after|{k}"] = to_np(v)
        if name != "NVFP4_DEFAULT_CFG":  # NVFP4 fake-quant has no CPU path in the reference
            with torch.no_grad():
                g[f"{name}|out"] = to_np(m(test))
    np.savez_compressed(os.path.join(OUT, "quantize_cpu.npz"), **g)
    print("quantize_cpu.npz", len(g), "arrays")


def gen_mse():
    """MseCalibrator of the reference on CPU (calib/mse.py:31-172) + mtq.quantize(algorithm='mse') end to end."""
    import copy

    import modelopt.torch.quantization as mtq
    from modelopt.torch.quantization.calib import MseCalibrator
    from modelopt.torch.quantization.tensor_quant import fake_tensor_quant, scaled_e4m3

    g = {}
    gen = torch.Generator().manual_seed(99)
    w = torch.randn(96, 320, generator=gen) * torch.rand(96, 1, generator=gen) * 3
    w[5] = 0
    w[7, 3] = 40.0  # outlier row: the search should clip it
    g["w"] = to_np(w)
    fmts = {"int8": lambda x, a: fake_tensor_quant(x, a, None, 8, False, True),
            "int4": lambda x, a: fake_tensor_quant(x, a, None, 4, False, True),
            "fp8": lambda x, a: scaled_e4m3(x, a, None, 4, 3)}
    for fn, f in fmts.items():
        for an, (axis, amax) in {"tensor": (None, w.abs().max()), "row": (0, w.abs().amax(dim=1, keepdim=True))}.items():
            if fn == "fp8" and an == "row":
                amax = amax.clamp(min=1e-3)
            c = MseCalibrator(amax=amax.clone(), axis=axis, quant_func=f)
            c.collect(w)
            c.collect(w * 0.5)  # two batches accumulate
            g[f"{fn}_{an}_amax0"] = to_np(amax)
            g[f"{fn}_{an}_losses"] = to_np(torch.stack(c._losses_sum))
            g[f"{fn}_{an}_best"] = to_np(c.compute_amax())
    base = _tiny_model()
    for k, v in base.state_dict().items():
        g[f"w_{k}"] = to_np(v)
    batches = [torch.randn(8, 16, 256, generator=gen) * (1 + i) for i in range(2)]
    for i, b in enumerate(batches):
        g[f"batch{i}"] = to_np(b)
    for name in ("INT8_DEFAULT_CFG", "FP8_DEFAULT_CFG", "INT4_BLOCKWISE_WEIGHT_ONLY_CFG"):
        cfg = copy.deepcopy(getattr(mtq, name))
        cfg["algorithm"] = {"method": "mse", "step_size": 0.25}
        m = mtq.quantize(copy.deepcopy(base), cfg, lambda model: [model(b) for b in batches])
        for n, mod in m.named_modules():
            if isinstance(mod, mtq.nn.TensorQuantizer) and hasattr(mod, "_amax"):
                g[f"{name}|{n}|_amax"] = to_np(mod._amax.float())
    np.savez_compressed(os.path.join(OUT, "mse_cpu.npz"), **g)
    print("mse_cpu.npz", len(g), "arrays")


def gen_awq_clip():
    """The reference's awq_clip / awq_full on CPU (model_calib.py:1724-1940) on the tiny two-linear model: per-ratio
    block losses (debug=True keeps them), the selected clip amax and the model output."""
    import copy

    import modelopt.torch.quantization as mtq

    g = {}
    gen = torch.Generator().manual_seed(11)
    batches = [torch.randn(8, 16, 256, generator=gen) * (1 + i) for i in range(2)]
    test = torch.randn(4, 16, 256, generator=gen) * 2
    base = _tiny_model()
    with torch.no_grad():  # a few outlier weights so that clipping below 1.0 wins in some blocks
        base[0].weight[::7, ::13] *= 6.0
        base[2].weight[::5, ::11] *= 6.0
    for k, v in base.state_dict().items():
        g[f"w_{k}"] = to_np(v)
    for i, b in enumerate(batches):
        g[f"batch{i}"] = to_np(b)
    g["test"] = to_np(test)
    cases = {
        "int4_clip_t64": ("INT4_AWQ_CFG", {"method": "awq_clip", "debug": True}),
        "int4_clip_t100": ("INT4_AWQ_CFG", {"method": "awq_clip", "max_tokens_per_batch": 100, "debug": True}),
        "int4_clip_step": ("INT4_AWQ_CFG", {"method": "awq_clip", "min_clip_ratio": 0.3, "shrink_step": 0.05,
                                            "debug": True}),
        "int4_full": ("INT4_AWQ_CFG", {"method": "awq_full", "alpha_step": 0.1, "debug": True}),
        "w4a8_clip": ("W4A8_AWQ_BETA_CFG", {"method": "awq_clip", "debug": True}),
    }
    for case, (name, algo) in cases.items():
        cfg = copy.deepcopy(getattr(mtq, name))
        cfg["algorithm"] = algo
        m = mtq.quantize(copy.deepcopy(base), cfg, lambda model: [model(b) for b in batches])
        for n, mod in m.named_modules():
            if isinstance(mod, mtq.nn.TensorQuantizer):
                g[f"{case}|{n}|enabled"] = np.asarray(mod.is_enabled)
                for buf in ("_amax", "_pre_quant_scale"):
                    if hasattr(mod, buf):
                        g[f"{case}|{n}|{buf}"] = to_np(getattr(mod, buf).float())
            if hasattr(mod, "awq_clip"):
                h = mod.awq_clip
                g[f"{case}|{n}|ratios"] = np.asarray(list(h.loss.keys()), dtype=np.float64)
                g[f"{case}|{n}|loss"] = to_np(torch.stack([v.float() for v in h.loss.values()]))
                g[f"{case}|{n}|w_amax"] = to_np(h.w_amax.float())
                g[f"{case}|{n}|num_tokens"] = np.asarray(h.num_tokens)
        with torch.no_grad():
            g[f"{case}|out"] = to_np(m(test))
    np.savez_compressed(os.path.join(OUT, "awq_clip_cpu.npz"), **g)
    print("awq_clip_cpu.npz", len(g), "arrays")


def gen_gpt2():
    """The reference's mtq.quantize(INT8_DEFAULT_CFG) on GPT-2 small on CPU: every quantizer's amax + a logits slice."""
    import copy

    import modelopt.torch.quantization as mtq

    from oracle.models import fingerprint, gpt2_model_and_batches

    model, batches = gpt2_model_and_batches()
    g = {"w_fingerprint": to_np(fingerprint(model))}

    def loop(m):
        with torch.no_grad():
            for b in batches:
                m(b)

    model = mtq.quantize(model, copy.deepcopy(mtq.INT8_DEFAULT_CFG), loop)
    n = 0
    for name, mod in model.named_modules():
        if isinstance(mod, mtq.nn.TensorQuantizer):
            g[f"{name}|enabled"] = np.asarray(mod.is_enabled)
            if hasattr(mod, "_amax"):
                g[f"{name}|_amax"] = to_np(mod._amax.float())
                n += 1
    with torch.no_grad():
        g["logits"] = to_np(model(batches[0][:1, :64]).logits[0, :, :256])
    np.savez_compressed(os.path.join(OUT, "gpt2_int8_cpu.npz"), **g)
    print("gpt2_int8_cpu.npz", len(g), "arrays,", n, "calibrated quantizers")


def gen_hist():
    from modelopt.torch.quantization.calib import HistogramCalibrator

    g = {}
    torch.manual_seed(31)
    cases = {
        "grow": [torch.randn(32, 512) * (1.0 + 0.25 * i) for i in range(3)],       # range grows every batch
        "same": [torch.randn(16, 1024).clamp(-3, 3) for _ in range(2)] + [torch.randn(16, 1024) * 0.5],
        "pos": [torch.rand(24, 256) * 4.0, torch.rand(24, 256) * 5.0],              # non-negative (no abs path)
    }
    cases["same"][0][0, 0] = 3.0
    for cn, xs in cases.items():
        for dn, dt in (("f32", torch.float32), ("bf16", torch.bfloat16)):
            xs_d = [x.to(dt) for x in xs]
            for bits, uns in ((8, False), (4, False), (8, True)):
                if uns and cn != "pos":
                    continue
                c = HistogramCalibrator(bits, None, uns)
                for i, x in enumerate(xs_d):
                    c.collect(x)
                    g[f"{cn}_{dn}_x{i}"] = to_np(x)
                    g[f"{cn}_{dn}_hist{i}"] = to_np(c._calib_hist.clone())
                    g[f"{cn}_{dn}_edges{i}"] = to_np(c._calib_bin_edges.clone())
                k = f"{cn}_{dn}_b{bits}_u{int(uns)}"
                for pct in (99.0, 99.9, 99.99, 100.0):
                    g[f"{k}_pct{pct}"] = to_np(c.compute_amax("percentile", percentile=pct))
                g[f"{k}_mse"] = to_np(c.compute_amax("mse"))
                g[f"{k}_mse_s4"] = to_np(c.compute_amax("mse", stride=4, start_bin=256))
                # entropy LAST: it writes hist[0] = hist[1] into a copy (.int() copies), the calibrator is unchanged
                g[f"{k}_entropy"] = to_np(c.compute_amax("entropy"))
                g[f"{k}_entropy_s3"] = to_np(c.compute_amax("entropy", stride=3, start_bin=200))
    np.savez_compressed(os.path.join(OUT, "hist_cpu.npz"), **g)
    print("hist_cpu.npz", len(g), "arrays")


if __name__ == "__main__":
    os.makedirs(OUT, exist_ok=True)
    which = sys.argv[1:] or ["qdq", "fp4_scales", "amax", "nvfp4_pack", "hist", "quantize", "mse", "gpt2", "awq_clip"]
    for w in which:
        globals()["gen_" + w]()
