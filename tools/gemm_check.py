"""Correctness + speed of the tcgen05 linear kernel against torch (cuBLAS) on the same operands."""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "tensorrt-model-optimizer_b200")]
from modelopt_b200 import ops  # noqa: E402
from tools.kbench import timeit  # noqa: E402


def check(M, N, K, dt=torch.bfloat16, bias=True):
    torch.manual_seed(M + N + K)
    x = torch.randn(M, K, device="cuda", dtype=dt)
    w = (torch.randn(N, K, device="cuda", dtype=dt) * 0.05)
    b = torch.randn(N, device="cuda", dtype=dt) if bias else None
    y = ops.linear(x, w, b)
    torch.cuda.synchronize()
    want32 = x.float() @ w.float().t() + (b.float() if bias else 0)
    ref = torch.nn.functional.linear(x, w, b)
    err = (y.float() - want32).abs().max().item()
    err_ref = (ref.float() - want32).abs().max().item()
    diff = (y != ref).float().mean().item()
    acc = torch.zeros(1, dtype=torch.float64, device="cuda")
    target = (ref.float() * 1.01).to(dt)
    ops.linear_loss_into(x, w, b, target, acc)
    want_loss = (y - target).float().pow(2).double().mean().item()
    print(f"M={M:6d} N={N:6d} K={K:6d} {str(dt)[6:]:9s} max|err| mine {err:.4f} cublas {err_ref:.4f} | "
          f"elements != cublas {diff * 100:.3f}% | loss {acc.item():.6g} want {want_loss:.6g}", flush=True)
    return err <= max(2 * err_ref, 1e-2)


def check_fused(M, N, K, dt=torch.bfloat16):
    torch.manual_seed(M + N + K + 1)
    x = torch.randn(M, K, device="cuda", dtype=dt) * 2
    w = torch.randn(N, K, device="cuda", dtype=dt) * 0.05
    b = torch.randn(N, device="cuda", dtype=dt)
    pqs = (torch.rand(K, device="cuda") + 0.5).to(dt)
    am = (x.float().abs().max() * 0.7)
    ok = True
    for fmt in ((4, 3), 8, 4):
        for ps in (None, pqs):
            if fmt == (4, 3):
                xq = ops.qdq_fp8(x, am, prescale=ps)
            else:
                xq = ops.qdq_int(x, am, fmt, False, False, prescale=ps)
            want = ops.linear(xq, w, b)
            got = ops.linear_qdq(x, w, b, am, fmt, False, prescale=ps)
            same = torch.equal(want, got)
            ok &= same
            print(f"fused prologue M={M} N={N} K={K} fmt={fmt} pqs={ps is not None}: identical to two-kernel path = {same}"
                  + ("" if same else f"  max|diff| {(want.float() - got.float()).abs().max().item():.4g}"), flush=True)
    return ok


def speed(M, N, K):
    x = torch.randn(M, K, device="cuda", dtype=torch.bfloat16)
    w = torch.randn(N, K, device="cuda", dtype=torch.bfloat16) * 0.05
    y = torch.empty(M, N, device="cuda", dtype=torch.bfloat16)
    acc = torch.zeros(1, dtype=torch.float64, device="cuda")
    fl = 2.0 * M * N * K
    am = x.float().abs().max()
    xq = torch.empty_like(x)
    for name, fn in (("tcgen05 linear", lambda: ops.linear(x, w, None, out=y)),
                     ("fp8 QDQ kernel + tcgen05 linear", lambda: ops.linear(ops.qdq_fp8(x, am, out=xq), w, None, out=y)),
                     ("tcgen05 linear, fp8 QDQ in prologue", lambda: ops.linear_qdq(x, w, None, am, (4, 3), out=y)),
                     ("int8 QDQ kernel + tcgen05 linear", lambda: ops.linear(ops.qdq_int(x, am, 8, False, False, out=xq), w, None, out=y)),
                     ("tcgen05 linear, int8 QDQ in prologue", lambda: ops.linear_qdq(x, w, None, am, 8, out=y)),
                     ("tcgen05 linear + fused loss", lambda: ops.linear_loss_into(x, w, None, y, acc)),
                     ("cuBLAS F.linear", lambda: torch.nn.functional.linear(x, w))):
        med, best = timeit(fn, 20, 5)
        print(f"  {name:38s} M={M} N={N} K={K}: {med:8.3f} ms  {fl / med / 1e9:8.1f} TFLOP/s (best {fl / best / 1e9:8.1f})",
              flush=True)


if __name__ == "__main__":
    ok = True
    for shp in [(128, 256, 64), (128, 256, 256), (256, 512, 128), (1000, 264, 72), (77, 8, 8), (4096, 4096, 4096)]:
        ok &= check(*shp)
    ok &= check(300, 512, 320, torch.float16)
    ok &= check(512, 1024, 512, torch.bfloat16, bias=False)
    for shp in [(128, 256, 64), (1000, 264, 72), (2048, 1024, 1024)]:
        ok &= check_fused(*shp)
    ok &= check_fused(300, 512, 320, torch.float16)
    print("ALL OK" if ok else "MISMATCH")
    if ok and "--speed" in sys.argv:
        for shp in [(16384, 4096, 4096), (16384, 14336, 4096), (16384, 4096, 14336), (16384, 1024, 4096)]:
            speed(*shp)
