"""Compile the small part of the reference that builds from its own sources with g++ (TEST INFRASTRUCTURE ONLY):
oracle/_ref/libref_mx.so = the host functions of gemm/tensor_quant_mx.h behind a C ABI (oracle/ref_mx_wrap.cpp).
Needs /root/reference (build container); the .so is git-ignored and travels to the GPU box with the snapshot."""

from __future__ import annotations

import os
import subprocess

HERE = os.path.dirname(os.path.abspath(__file__))
REF = os.environ.get("MODELOPT_REFERENCE", "/root/reference")
HDR_DIR = os.path.join(REF, "modelopt", "torch", "kernels", "quantization", "gemm")
OUT = os.path.join(HERE, "_ref", "libref_mx.so")


def build(force: bool = False) -> str | None:
    if not os.path.isdir(HDR_DIR):
        return OUT if os.path.exists(OUT) else None
    src = os.path.join(HERE, "ref_mx_wrap.cpp")
    if not force and os.path.exists(OUT) and os.path.getmtime(OUT) >= os.path.getmtime(src):
        return OUT
    os.makedirs(os.path.dirname(OUT), exist_ok=True)
    cuda_inc = os.path.join(os.environ.get("CUDA_HOME", "/usr/local/cuda"), "include")
    subprocess.check_call(["g++", "-O2", "-std=c++17", "-shared", "-fPIC", "-I", HDR_DIR, "-I", cuda_inc, src, "-o", OUT])
    return OUT


if __name__ == "__main__":
    print(build(force=True))
