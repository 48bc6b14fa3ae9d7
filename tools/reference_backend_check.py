"""INTEGRATION.md section 3, executed literally against the UNMODIFIED reference on a GPU.

A fresh process imports the reference package (from /root/reference in the build container, or from a copy under
oracle/_ref on the GPU box: `mkdir -p oracle/_ref && cp -r /root/reference/modelopt /root/reference/modelopt_recipes
oracle/_ref/`, git-ignored, deleted after the call), binds libmodelopt_b200.so with ctypes using nothing but the declarations of
include/modelopt_b200.h (modelopt_b200's Python is NOT imported), registers it with the reference's own
`register_quant_backend("b200", fn)` and runs the reference's `mtq.quantize` twice on identical models and batches: once with
the reference's kernels (JIT CUDA extensions / Triton), once with `{"backend": "b200"}` on every quantizer.  The calibrated
`_amax` of every quantizer and the quantized models' outputs must agree bit for bit.

TEST INFRASTRUCTURE (run by tests/test_gpu_reference_backend.py in a subprocess; skipped when the reference is absent)."""
import copy
import ctypes as C
import math
import os
import sys
import warnings

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF = next((p for p in ("/root/reference", os.path.join(ROOT, "oracle", "_ref")) if os.path.isdir(os.path.join(p, "modelopt"))), None)
if REF is None:
    print("reference not present: nothing to check")
    sys.exit(77)
sys.path[:0] = [REF, os.path.join(ROOT, "oracle", "ref_shims")]
sys.dont_write_bytecode = True
warnings.filterwarnings("ignore")

import torch  # noqa: E402
from torch import nn  # noqa: E402

import modelopt.torch.quantization as mtq  # noqa: E402  (the reference)
from modelopt.torch.quantization.nn import TensorQuantizer, register_quant_backend  # noqa: E402

assert "modelopt_b200" not in sys.modules

# ---- the binding: include/modelopt_b200.h through ctypes, as INTEGRATION.md section 1 shows -------------------------------
_DT = {torch.float32: 0, torch.float16: 1, torch.bfloat16: 2}
_lib = C.CDLL(os.environ.get("MODELOPT_B200_LIB", os.path.join(ROOT, "tensorrt-model-optimizer_b200", "modelopt_b200", "lib",
                                                                "libmodelopt_b200.so")))
_p, _i, _i64 = C.c_void_p, C.c_int, C.c_int64
_lib.mo_qdq_int.argtypes = [_p, _p, _i64, _i, _p, _i64, _i64, _i, _i, _i, _p, _p]
_lib.mo_qdq_fp8.argtypes = [_p, _p, _i64, _i, _p, _i64, _i64, _i, _p]
_lib.mo_qdq_nvfp4_dynamic.argtypes = [_p, _p, _i64, _i64, _i, _p, _i, _i, _p]
_lib.mo_last_error.restype = C.c_char_p
MO_NUMERICS_GPU = 0


def _ck(rc, what):
    if rc:
        raise RuntimeError(f"{what}: {_lib.mo_last_error().decode()} (status {rc})")


def _stream(t):
    return torch.cuda.current_stream(t.device).cuda_stream


def b200_backend(inputs: torch.Tensor, tq) -> torch.Tensor:
    """fn(inputs, tq) of register_quant_backend (nn/modules/tensor_quantizer.py:92-129): dispatch on the quantizer's format."""
    x = inputs.contiguous()
    y = torch.empty_like(x)
    amax = tq._get_amax(x)
    am = amax.detach().to(torch.float32).contiguous().view(-1)
    bs = tq.block_sizes
    if bs is not None and bs.get("type") == "dynamic" and tq.num_bits == (2, 1) and bs.get("scale_bits") == (4, 3):
        if am.numel() != 1:
            am = am.amax().reshape(1)
        cols = x.shape[-1]
        _ck(_lib.mo_qdq_nvfp4_dynamic(x.data_ptr(), y.data_ptr(), x.numel() // cols, cols, _DT[x.dtype], am.data_ptr(),
                                      bs[-1], MO_NUMERICS_GPU, _stream(x)), "mo_qdq_nvfp4_dynamic")
        return y
    if am.numel() == 1:
        axis_size, inner = 1, x.numel()
    else:
        a = list(amax.shape).index(amax.numel())
        axis_size, inner = x.shape[a], math.prod(x.shape[a + 1:])
    if tq.num_bits == (4, 3):
        _ck(_lib.mo_qdq_fp8(x.data_ptr(), y.data_ptr(), x.numel(), _DT[x.dtype], am.data_ptr(), axis_size, inner,
                            MO_NUMERICS_GPU, _stream(x)), "mo_qdq_fp8")
        return y
    assert isinstance(tq.num_bits, int)
    _ck(_lib.mo_qdq_int(x.data_ptr(), y.data_ptr(), x.numel(), _DT[x.dtype], am.data_ptr(), axis_size, inner, tq.num_bits,
                        int(tq.unsigned), int(tq.narrow_range), None, _stream(x)), "mo_qdq_int")
    return y


register_quant_backend("b200", b200_backend)


def with_backend(cfg):
    c = copy.deepcopy(cfg)
    for e in c["quant_cfg"]:
        if isinstance(e.get("cfg"), dict):
            e["cfg"]["backend"] = "b200"
    return c


def gpt2_small():
    from transformers import GPT2Config, GPT2LMHeadModel

    torch.manual_seed(0)
    return GPT2LMHeadModel(GPT2Config()).cuda().eval()


def toy(dtype):
    torch.manual_seed(0)
    return nn.Sequential(nn.Linear(256, 512), nn.ReLU(), nn.Linear(512, 256)).cuda().to(dtype).eval()


def run(name, make_model, batches, cfg_name, call):
    results = []
    for cfg in (getattr(mtq, cfg_name), with_backend(getattr(mtq, cfg_name))):
        model = make_model()
        with torch.no_grad():
            mtq.quantize(model, copy.deepcopy(cfg), lambda m: [call(m, b) for b in batches])
            out = call(model, batches[0])
        amax = {n: q._amax.detach().float().clone() for n, q in model.named_modules()
                if isinstance(q, TensorQuantizer) and getattr(q, "_amax", None) is not None}
        used = sum(1 for _, q in model.named_modules() if isinstance(q, TensorQuantizer) and q.backend == "b200" and q.is_enabled)
        results.append((out.float(), amax, used))
    (o_ref, a_ref, u_ref), (o_b, a_b, u_b) = results
    assert u_ref == 0 and u_b > 0, (u_ref, u_b)
    same_amax = a_ref.keys() == a_b.keys() and all(torch.equal(a_ref[k], a_b[k]) for k in a_ref)
    same_out = torch.equal(o_ref, o_b)
    nz = int((o_ref != o_b).sum())
    print(f"{name:10s} {cfg_name:22s} quantizers on the b200 backend: {u_b:3d}  calibrated amax: {len(a_ref):3d} identical: {same_amax}  "
          f"outputs bit-identical: {same_out} ({nz} of {o_ref.numel()} differ)", flush=True)
    return same_amax and same_out


def main():
    ok = True
    for dtype in (torch.bfloat16, torch.float32):
        g = torch.Generator(device="cuda").manual_seed(1)
        batches = [torch.randn(64, 256, device="cuda", generator=g).to(dtype) * (1 + i) for i in range(2)]
        for cfg_name in ("INT8_DEFAULT_CFG", "FP8_DEFAULT_CFG", "NVFP4_DEFAULT_CFG", "INT4_BLOCKWISE_WEIGHT_ONLY_CFG"):
            ok &= run(f"toy/{str(dtype)[6:]}", lambda: toy(dtype), batches, cfg_name, lambda m, b: m(b))
    ids = [torch.randint(0, 50257, (4, 512), device="cuda", generator=torch.Generator(device="cuda").manual_seed(3 + i))
           for i in range(2)]
    for cfg_name in ("INT8_DEFAULT_CFG", "FP8_DEFAULT_CFG"):
        ok &= run("gpt2-small", gpt2_small, ids, cfg_name, lambda m, b: m(input_ids=b).logits)
    print("REFERENCE BACKEND CHECK", "OK" if ok else "FAILED", f"({torch.cuda.get_device_name()}, reference at {REF})")
    sys.exit(0 if ok else 1)


if __name__ == "__main__":
    main()
