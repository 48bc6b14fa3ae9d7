"""NVFP4 W4A4 GEMM: the CTA-pair kernel (mo_linear_nvfp4_pair) against the single-CTA kernel (mo_linear_nvfp4) - results
(same operands: outputs must be identical, the products are exact and both accumulate the K steps in order) and
TFLOP/s under SUSTAINED load (each candidate back to back for ~0.5 s, SM clock sampled).  Developer tool:
    python tools/gemm_nvfp4_lab.py [--check-only] [--secs 0.5]
Every launch runs under the kernels' bounded mbarrier waits (a protocol bug traps instead of hanging)."""
import argparse
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "tensorrt-model-optimizer_b200")]
from modelopt_b200 import ops  # noqa: E402
from tools.gemm_sustained import sustained  # noqa: E402


def operands(M, N, K, dt=torch.bfloat16):
    torch.manual_seed(M + N + K)
    x = torch.randn(M, K, device="cuda", dtype=dt) * 2
    w = torch.randn(N, K, device="cuda", dtype=dt) * 0.05
    ax, aw = x.float().abs().max(), w.float().abs().max()
    return ops.nvfp4_quantize_operand(x, ax, 128), ops.nvfp4_quantize_operand(w, aw, 256)


def check(M, N, K, bias=True, dt=torch.bfloat16):
    a, b = operands(M, N, K, dt)
    bv = torch.randn(N, device="cuda", dtype=dt) if bias else None
    y1 = ops.linear_nvfp4(*a, *b, bv, out_dtype=dt, pair=False)
    y2 = ops.linear_nvfp4(*a, *b, bv, out_dtype=dt, pair=True)
    y3 = ops.linear_nvfp4(*a, *b, bv, out_dtype=dt, pair=True)
    torch.cuda.synchronize()
    same, rep = torch.equal(y1, y2), torch.equal(y2, y3)
    d = (y1.float() - y2.float()).abs().max().item()
    print(f"M={M:6d} N={N:6d} K={K:6d} {str(dt)[6:]:8s} bias={int(bias)}: pair == single {same} (max|diff| {d:.4g}, "
          f"max|out| {y1.float().abs().max().item():.4g}), repeatable {rep}", flush=True)
    return same and rep


VARIANTS = [  # label, MO_LAB_F4P, MO_LAB_F4P_GROUPM
    ("overlapped acc", 0, 0), ("single acc", 1, 0), ("overlapped, group_m 2", 0, 2), ("overlapped, group_m 8", 0, 8),
    ("overlapped, group_m 32", 0, 32),
]


def variants(secs):
    M = 16384
    for N, K in [(4096, 4096), (4096, 14336)]:
        xa, wb = operands(M, N, K)
        y = torch.empty(M, N, device="cuda", dtype=torch.bfloat16)
        ref = ops.linear_nvfp4(*xa, *wb, None, pair=False)
        fl = 2.0 * M * N * K
        ms, mhz = sustained(lambda: ops.linear_nvfp4(*xa, *wb, None, out=y, pair=False), secs)
        print(f"N={N} K={K}: single-CTA kernel {ms:7.3f} ms {fl / ms / 1e9:7.1f} TF ({mhz} MHz)", flush=True)
        for label, lab, dbg in VARIANTS:
            os.environ["MO_LAB_F4P"], os.environ["MO_LAB_F4P_GROUPM"] = str(lab), str(dbg)
            try:
                ms, mhz = sustained(lambda: ops.linear_nvfp4(*xa, *wb, None, out=y, pair=True), secs)
                same = torch.equal(y, ref)
                print(f"   {label:36s} {ms:7.3f} ms {fl / ms / 1e9:7.1f} TF ({mhz} MHz)"
                      + ("" if same is None else f"  identical={same}"), flush=True)
            finally:
                os.environ["MO_LAB_F4P"], os.environ["MO_LAB_F4P_GROUPM"] = "0", "0"
    return 0


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--check-only", action="store_true")
    ap.add_argument("--secs", type=float, default=0.5)
    ap.add_argument("--variants", action="store_true", help="time the lab variants of the pair kernel (MO_LAB_F4P / _DBG)")
    a = ap.parse_args()
    if a.variants:
        return variants(a.secs)
    ok = True
    for shp in [(256, 256, 256), (256, 256, 1024), (129, 256, 512), (512, 512, 2560), (1000, 264, 768), (300, 8, 256),
                (4096, 4096, 4096), (2048, 14336, 4096), (1111, 4096, 14336)]:
        ok &= check(*shp)
    ok &= check(512, 1024, 512, bias=False, dt=torch.float16)
    print("ALL IDENTICAL" if ok else "MISMATCH", flush=True)
    if a.check_only or not ok:
        return 0 if ok else 1
    print("sustained TFLOP/s, M = 16384 (SM MHz during the run)")
    for N, K in [(4096, 4096), (14336, 4096), (4096, 14336), (6144, 4096), (28672, 8192), (8192, 28672)]:
        M = 16384
        xa, wb = operands(M, N, K)
        y = torch.empty(M, N, device="cuda", dtype=torch.bfloat16)
        fl = 2.0 * M * N * K
        line = f"N={N:6d} K={K:6d}:"
        for name, fn in (("single", lambda: ops.linear_nvfp4(*xa, *wb, None, out=y, pair=False)),
                         ("pair", lambda: ops.linear_nvfp4(*xa, *wb, None, out=y, pair=True))):
            ms, mhz = sustained(fn, a.secs)
            line += f"  {name} {ms:7.3f} ms {fl / ms / 1e9:7.1f} TF ({mhz} MHz)"
        print(line, flush=True)
    return 0


if __name__ == "__main__":
    sys.exit(main())
