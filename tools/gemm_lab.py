"""Kernel lab for the CTA-pair bf16 linear: builds variants of csrc/gemm_2cta.cu with -DMO_GEMM_EXP=<bits> / other -D
overrides into tools/_lab/ (not part of the product library) and times them under sustained load against cuBLAS.
  build (CPU box):  python tools/gemm_lab.py build
  run (GPU box):    python tools/gemm_lab.py run
-DMO_GEMM_EXP bits: 1 = no operand traffic (MMAs on stale shared memory), 2 = no epilogue work, 4 = TMEM read only;
-DMO_LAB_MH=1|2 forces the 256 x 256 / 512 x 256 pair tile, -DMO_LAB_GROUPM / -DMO_LAB_STAGES / -DMO_LAB_L2PROMO override
the tile walk, ring depth and L2 promotion; LAB_VARIANTS="name=flags;..." adds variants, LAB_LOSS=1 times the loss mode."""
import ctypes as C
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LAB = os.path.join(ROOT, "tools", "_lab")
CSRC = os.path.join(ROOT, "tensorrt-model-optimizer_b200", "csrc")
VARIANTS = {"base": [], "mh1": ["-DMO_LAB_MH=1"], "mh2": ["-DMO_LAB_MH=2"]}
for spec in os.environ.get("LAB_VARIANTS", "").split(";"):
    if spec:
        name, flags = spec.split("=", 1)
        VARIANTS[name] = flags.split()


def build():
    os.makedirs(LAB, exist_ok=True)
    for name, flags in VARIANTS.items():
        out = os.path.join(LAB, f"libgemm_{name}.so")
        cmd = ["/usr/local/cuda/bin/nvcc", "-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-std=c++17", "-lineinfo",
               "-Xcompiler", "-fPIC", "--expt-relaxed-constexpr", "-shared", *flags,
               os.path.join(CSRC, "gemm_2cta.cu"), os.path.join(CSRC, "runtime.cu"), "-o", out, "-lcudart"]
        subprocess.run(cmd, check=True)
        print("built", out)


def run():
    import torch

    sys.path[:0] = [ROOT, os.path.join(ROOT, "tools"), os.path.join(ROOT, "tensorrt-model-optimizer_b200")]
    from gemm_sustained import sustained

    shapes = [(4096, 4096), (14336, 4096), (4096, 14336)]
    M = 16384
    import glob

    libs = {}
    only = [v for v in os.environ.get("LAB_ONLY", "").split(",") if v]
    for path in sorted(glob.glob(os.path.join(LAB, "libgemm_*.so"))):
        name = os.path.basename(path)[len("libgemm_"):-3]
        if only and name not in only:
            continue
        lib = C.CDLL(path)
        lib.mo_linear_pair.argtypes = [C.c_void_p] * 4 + [C.c_int64] * 3 + [C.c_int] + [C.c_void_p] * 3
        lib.mo_linear_fp8_pair.argtypes = [C.c_void_p] * 6 + [C.c_int64] * 3 + [C.c_int, C.c_void_p]
        libs[name] = lib
    for N, K in shapes:
        x = torch.randn(M, K, device="cuda", dtype=torch.bfloat16)
        w = (torch.randn(N, K, device="cuda") * 0.02).to(torch.bfloat16)
        out = torch.empty(M, N, device="cuda", dtype=torch.bfloat16)
        flops = 2.0 * M * N * K
        stream = torch.cuda.current_stream().cuda_stream
        cands = {"cublas": lambda: torch.nn.functional.linear(x, w)}
        loss_mode = os.environ.get("LAB_LOSS", "0") == "1"
        fp8 = os.environ.get("LAB_FP8", "0") == "1"
        if fp8:
            a8, b8 = x.to(torch.float8_e4m3fn), w.to(torch.float8_e4m3fn)
            one = torch.ones(1, device="cuda")
            flops = 2.0 * M * N * K
            cands = {"scaled_mm": lambda: torch._scaled_mm(a8, b8.t(), one, one, None, out_dtype=torch.bfloat16, use_fast_accum=True)}
        ref = torch.randn(M, N, device="cuda", dtype=torch.bfloat16)
        acc = torch.zeros(1, dtype=torch.float64, device="cuda")
        for name, lib in libs.items():
            def fn(lib=lib):
                if fp8:
                    rc = lib.mo_linear_fp8_pair(a8.data_ptr(), b8.data_ptr(), one.data_ptr(), one.data_ptr(), None, out.data_ptr(),
                                                M, N, K, 2, stream)
                elif loss_mode:
                    rc = lib.mo_linear_pair(x.data_ptr(), w.data_ptr(), None, None, M, N, K, 2, ref.data_ptr(), acc.data_ptr(), stream)
                else:
                    rc = lib.mo_linear_pair(x.data_ptr(), w.data_ptr(), None, out.data_ptr(), M, N, K, 2, None, None, stream)
                assert rc == 0, rc
            cands[name] = fn
        res = {k: [] for k in cands}
        for _ in range(2):  # two interleaved rounds: the box's thermal state drifts
            for name, fn in cands.items():
                ms, mhz = sustained(fn, 0.35)
                res[name].append((flops / ms / 1e9, mhz))
        line = f"N={N:6d} K={K:6d}:"
        for name, r in res.items():
            line += f"  {name} {r[0][0]:6.0f}/{r[1][0]:6.0f} @{r[1][1]}"
        print(line, flush=True)


if __name__ == "__main__":
    {"build": build, "run": run}[sys.argv[1]]()
