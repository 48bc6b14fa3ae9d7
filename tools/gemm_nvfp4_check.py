"""Correctness + speed of the block-scaled NVFP4 linear: against an fp32 matmul of the DEQUANTIZED operands (the exact
value the instruction should produce up to fp32 accumulation order) and, for context, against the fake-quant bf16 path."""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "tensorrt-model-optimizer_b200")]
from modelopt_b200 import ops  # noqa: E402
from tools.kbench import timeit  # noqa: E402


def operands(M, N, K, dt=torch.bfloat16):
    torch.manual_seed(M + N + K)
    x = torch.randn(M, K, device="cuda", dtype=dt) * 2
    w = torch.randn(N, K, device="cuda", dtype=dt) * 0.05
    return x, w


def check(M, N, K, bias=True):
    x, w = operands(M, N, K)
    b = torch.randn(N, device="cuda", dtype=torch.bfloat16) if bias else None
    ax, aw = x.float().abs().max(), w.float().abs().max()
    xc, xs, xg = ops.nvfp4_quantize_operand(x, ax, 128)
    wc, ws, wg = ops.nvfp4_quantize_operand(w, aw, 256)
    y = ops.linear_nvfp4(xc, xs, xg, wc, ws, wg, b)
    torch.cuda.synchronize()
    # dequantized operands in fp32 from the same codes / scales
    _, xsf = ops.nvfp4_pack(x, 16, xg)
    _, wsf = ops.nvfp4_pack(w, 16, wg)
    xd = ops.nvfp4_unpack(xc, xsf, xg, 16, torch.float32)
    wd = ops.nvfp4_unpack(wc, wsf, wg, 16, torch.float32)
    want = xd.double() @ wd.double().t() + (b.double() if bias else 0)
    err = (y.double() - want).abs().max().item()
    scale = want.abs().max().item()
    rel = ((y.double() - want).abs() / (want.abs() + 1e-3 * scale)).max().item()
    # the fake-quant linear the reference simulates (operands rounded to bf16 after dequantization)
    fq = torch.nn.functional.linear(ops.qdq_nvfp4_dynamic(x, ax, 16), ops.qdq_nvfp4_dynamic(w, aw, 16), b)
    d_fq = (y.float() - fq.float()).abs().max().item()
    ok = rel < 2.0 ** -7
    print(f"M={M:6d} N={N:6d} K={K:6d}: max|err| vs fp64 matmul of dequantized operands {err:.4g} (max|out| {scale:.3g}, "
          f"max rel {rel:.3g}) | vs fake-quant bf16 linear max|diff| {d_fq:.4g}  {'OK' if ok else 'MISMATCH'}", flush=True)
    return ok


def speed(M, N, K):
    x, w = operands(M, N, K)
    ax, aw = x.float().abs().max(), w.float().abs().max()
    xc, xs, xg = ops.nvfp4_quantize_operand(x, ax, 128)
    wc, ws, wg = ops.nvfp4_quantize_operand(w, aw, 256)
    y = torch.empty(M, N, device="cuda", dtype=torch.bfloat16)
    fl = 2.0 * M * N * K
    xq = torch.empty_like(x)
    for name, fn in (("nvfp4 block-scaled linear (operands packed)", lambda: ops.linear_nvfp4(xc, xs, xg, wc, ws, wg, None, out=y)),
                     ("  + activation pack + scale swizzle", lambda: ops.linear_nvfp4(*ops.nvfp4_quantize_operand(x, ax, 128), wc, ws, wg, None, out=y)),
                     ("fake-quant: nvfp4 QDQ kernel + bf16 tcgen05 linear", lambda: ops.linear(ops.qdq_nvfp4_dynamic(x, ax, 16, out=xq), w, None, out=y)),
                     ("cuBLAS bf16 F.linear", lambda: torch.nn.functional.linear(x, w))):
        med, best = timeit(fn, 20, 5)
        print(f"  {name:52s} M={M} N={N} K={K}: {med:8.3f} ms  {fl / med / 1e9:8.1f} TFLOP/s (best {fl / best / 1e9:8.1f})", flush=True)


if __name__ == "__main__":
    ok = True
    for shp in [(128, 256, 256), (128, 256, 1024), (256, 512, 512), (1000, 264, 768), (77, 8, 256), (4096, 4096, 4096)]:
        ok &= check(*shp)
    ok &= check(512, 1024, 512, bias=False)
    print("ALL OK" if ok else "MISMATCH")
    if ok and "--speed" in sys.argv:
        for shp in [(16384, 4096, 4096), (16384, 14336, 4096), (16384, 4096, 14336)]:
            speed(*shp)
