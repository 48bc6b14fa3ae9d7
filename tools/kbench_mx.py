"""Kernel timings for the MX fake-quant and the real-quant packers (CUDA events, median of 20 after 5 warm-ups, inputs
larger than L2).  Developer tool: `gpurun -- python tools/kbench_mx.py > gpurun_out/kbench_mx.txt`."""

import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "tensorrt-model-optimizer_b200")]
from modelopt_b200 import ops  # noqa: E402
from modelopt_b200.qtensor import FP8QTensor, INT8QTensor  # noqa: E402

PEAK = 6448.7
try:
    PEAK = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"]
except Exception:  # noqa: BLE001
    pass


def timeit(fn, n=20, warm=5):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    ts = []
    for _ in range(n):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        fn()
        b.record()
        torch.cuda.synchronize()
        ts.append(a.elapsed_time(b))
    ts.sort()
    return ts[len(ts) // 2]


def row(name, fn, nbytes):
    ms = timeit(fn)
    gbs = nbytes / ms / 1e6
    print(f"{name:58s} {ms:8.3f} ms  {gbs:8.1f} GB/s  {100 * gbs / PEAK:5.1f} % of {PEAK:.0f}")


def main():
    torch.manual_seed(0)
    R, C = 16384, 14336
    x = (torch.randn(R, C, device="cuda") * 2).to(torch.bfloat16)
    n = x.numel()
    y = torch.empty_like(x)
    print(f"# {torch.cuda.get_device_name(0)}  [{R}, {C}] bf16 = {n * 2 / 1e6:.0f} MB; algorithmic bytes in the 3rd column's rate")
    for fmt, bits in (("MXFP8 E4M3", (4, 3)), ("MXFP4 E2M1", (2, 1)), ("MXFP6 E3M2", (3, 2)), ("MXINT8", 8), ("E2M3", (2, 3)),
                      ("E5M2", (5, 2))):
        row(f"qdq_mx {fmt} block 32 (gpu numerics)", lambda b=bits: ops.qdq_mx(x, 32, b, out=y), 4 * n)
    row("qdq_mx MXFP8 block 32 (cpu numerics)", lambda: ops.qdq_mx(x, 32, (4, 3), numerics="cpu", out=y), 4 * n)
    row("qdq_mx E4M3 block 16", lambda: ops.qdq_mx(x, 16, (4, 3), out=y), 4 * n)
    row("qdq_mx E2M1/E4M3 scales + global amax (generic kernel)",
        lambda: ops.qdq_mx(x, 16, (2, 1), (4, 3), x[0, :1].float().abs() + 8, out=y), 4 * n)
    row("torch copy (reference point)", lambda: y.copy_(x), 4 * n)
    w = x  # the whole 470 MB tensor: kernel times well above the Python launch overhead (~0.1 ms per call)
    nw = w.numel()
    row("mxfp8_pack (amax + E8M0 + e4m3 codes)", lambda: ops.mxfp8_pack(w), nw * (2 + 1 + 1 / 32))
    q8, s8 = ops.mxfp8_pack(w)
    row("mx_unpack MXFP8 -> bf16", lambda: ops.mx_unpack(q8, s8, 32, torch.bfloat16), nw * (1 + 2 + 1 / 32))
    row("mxfp4_pack block 32", lambda: ops.mxfp4_pack(w, 32), nw * (2 + 0.5 + 1 / 32))
    q4, s4 = ops.mxfp4_pack(w, 32)
    row("mx_unpack MXFP4 -> bf16", lambda: ops.mx_unpack(q4.reshape(-1, 16), s4, 32, torch.bfloat16, fp4=True), nw * (0.5 + 2 + 1 / 32))
    flat = w.reshape(-1)
    row("int4_pack block 128 (amax + scale + codes, one pass)", lambda: ops.int4_pack(flat, 128), nw * (2 + 0.5 + 2 / 128))
    qi, si = ops.int4_pack(flat, 128)
    row("int4_unpack block 128 -> bf16", lambda: ops.int4_unpack(qi, si, 128), nw * (0.5 + 2 + 2 / 128))
    row("nf4_pack block 64", lambda: ops.nf4_pack(flat, 64), nw * (2 + 0.5 + 2 / 64))
    qn, sn = ops.nf4_pack(flat, 64)
    row("nf4_unpack block 64 -> bf16", lambda: ops.nf4_unpack(qn, sn.float().reshape(-1), 64, True), nw * (0.5 + 2 + 4 / 64))
    for name, cls, kind in (("INT8", INT8QTensor, "int8"), ("FP8", FP8QTensor, "fp8")):
        s_pc = ops.reduce_amax(w, axis=1) / (127.0 if kind == "int8" else 448.0)
        row(f"q8_pack {name} per-channel (scales given)", lambda k=kind, s=s_pc: ops.q8_pack(w, s, k), nw * 3)
        row(f"{name}QTensor.quantize per-channel (amax pass + pack)", lambda c=cls: c.quantize(w, axis=0), nw * 5)
        row(f"{name}QTensor.quantize 128x128 blocks (amax passes + pack)",
            lambda c=cls: c.quantize(w, block_sizes={-1: 128, -2: 128}), nw * 5)
        qq = ops.q8_pack(w, s_pc, kind)
        row(f"q8_unpack {name} per-channel -> bf16", lambda k=kind, q=qq, s=s_pc: ops.q8_unpack(q, s, torch.bfloat16, k), nw * 3)


def with_clocks(fn):
    """Run fn() with the SM clock and throttle reasons sampled (bench.py's sampler) and print them: a kernel figure without
    the clocks it was taken at is a claim, not a measurement."""
    sys.path[:0] = [ROOT]
    from bench import Clocks

    with Clocks(torch.cuda.current_device()) as clk:
        fn()
    print("# clocks during this run:", clk.summary(), flush=True)


if __name__ == "__main__":
    with_clocks(main)
