"""bench.py — ModelOpt's quantization hot path on B200: Llama-3 calibration, kernel level and model level.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference] [--workload nvfp4|fp8] [--model 8b|70b]

One JSON line.  Its parts:

`value` / `roofline` — the quantizer STEP (BASELINE.json configs[2], NVFP4_DEFAULT_CFG; `--workload fp8` = configs[1]):
  the work every TensorQuantizer of a random-init Llama-3-8B does for one micro-batch of [8, 2048] = 16384 synthetic
  tokens, model GEMMs excluded:
    calibration pass : MaxCalibrator.collect on the input of each of the 7 x 32 quantized linears (activation
                       [16384, Cin] bf16 -> one amax word in the arena) and on each weight;
    fake-quant pass  : TensorQuantizer._fake_quantize of the same activations and weights with the calibrated amax.
  Data-parallel: every rank processes its own micro-batch (weak scaling), one all-reduce(MAX) over the amax
  arena per step.  Inputs exceed L2 (each tensor >= 134 MB, reuse distance > 1.2 GB): no L2 flush needed.
`e2e` — the same step through the HOST-buffer C ABI (`mo_calib_qdq_host_batch`): activations come from pinned host memory
  every step (weights live on the GPU, as in every real use), the result read back is what max_calibrate needs (the 448
  running amax words); `copy_ceiling_gbs` is a plain pinned H2D copy of the same bytes on all ranks at once.
`workloads` — the other configs as rooflines of their own: fp8 (configs[1]) quantizer step, awq (configs[3]: INT4_AWQ
  awq_lite on one full decoder layer through `quantize`), hist (HistogramCalibrator.collect).
`model_level` — configs[1] / [2] / [4] as written: `modelopt_b200.quantize(LlamaForCausalLM(<dims>), cfg, forward_loop)` over
  512 x 2048 synthetic tokens split over the ranks (STRONG scaling), beside the plain bf16 forward.
`cpu_baseline` / `--impl reference` — the reference's CPU path (oracle port) on the host cores over a bounded sample.
"""

from __future__ import annotations

import argparse
import gc
import json
import os
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path[:0] = [ROOT, os.path.join(ROOT, "tensorrt-model-optimizer_b200"), os.path.join(ROOT, "tools")]

import torch  # noqa: E402

# (name, Cout, Cin) of the quantized linears of one decoder layer; SURVEY.md section 8 config table
SHAPES = {
    "8b": {"name": "Llama-3-8B", "layers": 32,
           "linears": [("q_proj", 4096, 4096), ("k_proj", 1024, 4096), ("v_proj", 1024, 4096), ("o_proj", 4096, 4096),
                       ("gate_proj", 14336, 4096), ("up_proj", 14336, 4096), ("down_proj", 4096, 14336)]},
    "70b": {"name": "Llama-3-70B", "layers": 80,
            "linears": [("q_proj", 8192, 8192), ("k_proj", 1024, 8192), ("v_proj", 1024, 8192), ("o_proj", 8192, 8192),
                        ("gate_proj", 28672, 8192), ("up_proj", 28672, 8192), ("down_proj", 8192, 28672)]},
}
TOKENS = 8 * 2048
METRIC = "calib tokens/s"
UNIT = "tokens/s"
KERNEL = {"nvfp4": "nvfp4_dynamic_kernel<bf16,16>", "fp8": "qdq_stream_kernel<bf16,Fp8Op>"}


def parse():
    p = argparse.ArgumentParser()
    p.add_argument("--gpus", type=int, default=1)
    p.add_argument("--steps", type=int, default=10)
    p.add_argument("--warmup", type=int, default=3)
    p.add_argument("--impl", default="b200", choices=["b200", "reference"])
    p.add_argument("--workload", default="nvfp4", choices=["nvfp4", "fp8"])
    p.add_argument("--model", default="8b", choices=list(SHAPES))
    p.add_argument("--layers", type=int, default=None)
    p.add_argument("--tokens", type=int, default=TOKENS)
    p.add_argument("--e2e-steps", type=int, default=2)
    p.add_argument("--seqs", type=int, default=512, help="model-level calibration set (sequences of 2048 tokens)")
    p.add_argument("--micro-batch", type=int, default=None, help="model-level micro-batch (default 8; 70b: 4)")
    p.add_argument("--model-presets", default=None, help="comma list for the model-level leg (default nvfp4,fp8,hist; 70b: nvfp4,hist)")
    p.add_argument("--no-e2e", action="store_true")
    p.add_argument("--no-cpu-baseline", action="store_true")
    p.add_argument("--no-workloads", action="store_true")
    p.add_argument("--no-model-level", action="store_true")
    a = p.parse_args()
    a.linears = SHAPES[a.model]["linears"]
    a.layers = a.layers or SHAPES[a.model]["layers"]
    return a


PRESET_TEXT = {"nvfp4": "NVFP4_DEFAULT_CFG (block-16 E4M3 scales, fp32 global amax)", "fp8": "FP8_DEFAULT_CFG (per-tensor E4M3)"}


def workload_name(a, preset=None):
    return (f"{SHAPES[a.model]['name']} random-init {PRESET_TEXT[preset or a.workload]} max calibration + fake-quant, "
            f"quantizer hot path only, micro-batch {a.tokens} tokens/GPU, {a.layers} layers x 7 linears")


# =================================================================================================
# clocks sampler (nvml)
# =================================================================================================
class Clocks:
    def __init__(self, index):
        self.index, self.samples, self.reasons, self.max_mhz = index, [], set(), None
        self._stop = threading.Event()
        self._ready = threading.Event()
        self._t = None

    def _run(self):
        try:
            import pynvml as nv

            nv.nvmlInit()
            h = nv.nvmlDeviceGetHandleByIndex(self.index)
            self.max_mhz = nv.nvmlDeviceGetMaxClockInfo(h, nv.NVML_CLOCK_SM)
            names = {"hw_slowdown": 0x8, "sw_power_cap": 0x4, "sw_thermal_slowdown": 0x20,
                     "hw_thermal_slowdown": 0x40, "hw_power_brake": 0x80, "sync_boost": 0x10}
            while not self._stop.is_set():
                self.samples.append(nv.nvmlDeviceGetClockInfo(h, nv.NVML_CLOCK_SM))
                r = nv.nvmlDeviceGetCurrentClocksEventReasons(h) if hasattr(nv, "nvmlDeviceGetCurrentClocksEventReasons") \
                    else nv.nvmlDeviceGetCurrentClocksThrottleReasons(h)
                for n, bit in names.items():
                    if r & bit:
                        self.reasons.add(n)
                self._ready.set()
                time.sleep(0.01)
        except Exception as e:  # noqa: BLE001
            self.reasons.add(f"sampler_error:{type(e).__name__}")
        self._ready.set()

    def __enter__(self):
        self._t = threading.Thread(target=self._run, daemon=True)
        self._t.start()
        self._ready.wait(timeout=5)
        return self

    def __exit__(self, *a):
        self._stop.set()
        self._t.join(timeout=2)

    def summary(self):
        s = sorted(self.samples)
        return {"sm_mhz": s[len(s) // 2] if s else None, "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons),
                "samples": len(s)}


def hbm_peak():
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        return float(peaks["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)", peaks
    except Exception:  # noqa: BLE001
        return 6650.0, "fallback (B200_PROFILING.md)", {}


# =================================================================================================
# the quantizer step, device-resident
# =================================================================================================
class DeviceWorkload:
    """Device-resident tensors of one rank + the per-step launch sequence."""

    def __init__(self, a, device, rank, preset, weights=None):
        from modelopt_b200 import ops
        from modelopt_b200.arena import AmaxArena

        self.ops, self.a, self.device, self.preset = ops, a, device, preset
        g = torch.Generator(device=device).manual_seed(1234 + rank)
        bf = torch.bfloat16
        T = a.tokens
        L = a.linears
        # weights: every layer has its own (random-init, std 0.02)
        self.w = weights if weights is not None else \
            [[torch.randn((co, ci), generator=g, device=device, dtype=bf) * 0.02 for _, co, ci in L] for _ in range(a.layers)]
        # activations: 7 distinct inputs per layer slot, reused across layers (reuse distance 1.27 GB >> L2)
        self.x = [torch.randn((T, ci), generator=g, device=device, dtype=bf) for _, _, ci in L]
        self.y = [torch.empty_like(t) for t in self.x]
        self.wy = [torch.empty_like(t) for t in self.w[0]]
        self.arena = AmaxArena(device, 8 * 2 * 7 * a.layers + 64)
        self.slot_x = [[self.arena.alloc(1) for _ in L] for _ in range(a.layers)]
        self.slot_w = [[self.arena.alloc(1) for _ in L] for _ in range(a.layers)]
        self.elems_act = sum(t.numel() for t in self.x) * a.layers
        self.elems_w = sum(t.numel() for layer in self.w for t in layer)
        self.big = max(range(len(L)), key=lambda i: self.x[i].numel())  # down_proj input [T, ffn]
        self.kernel_events = None  # when a list: (start, stop) pairs around the dominant kernel's launches

    def algorithmic_bytes(self):
        # collect: 2 B/elem (one bf16 read); fake-quant: 4 B/elem (read + write)   SURVEY.md 8(d)
        n = self.elems_act + self.elems_w
        return 2 * n + 4 * n

    def _qdq(self, x, amax_f32, out):
        if self.preset == "nvfp4":
            return self.ops.qdq_nvfp4_dynamic(x, amax_f32, 16, None, out=out)
        return self.ops.qdq_fp8(x, amax_f32, out=out)

    def step(self):
        ops, arena = self.ops, self.arena
        n = len(self.a.linears)
        for layer in range(self.a.layers):
            # calibration pass (TensorQuantizer.collect -> MaxCalibrator.collect): 14 tensors
            for i in range(n):
                ops.amax_into(self.x[i], 1, 1, self.x[i].numel(), arena.bits(self.slot_x[layer][i]))
                w = self.w[layer][i]
                ops.amax_into(w, 1, 1, w.numel(), arena.bits(self.slot_w[layer][i]))
            # fake-quant pass (TensorQuantizer._fake_quantize) with the running amax; the tensor read here
            # was last touched 1.7 GB of traffic ago, so nothing is served from L2
            for i in range(n):
                ev = self.kernel_events is not None and i == self.big
                if ev:
                    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                    e0.record()
                self._qdq(self.x[i], arena.amax(self.slot_x[layer][i]), self.y[i])
                if ev:
                    e1.record()
                    self.kernel_events.append((e0, e1))
                self._qdq(self.w[layer][i], arena.amax(self.slot_w[layer][i]), self.wy[i])
        arena.sync()  # DP exchange: one all-reduce(MAX) over every quantizer's amax


def read_traffic(preset, shape):
    """Per-launch DRAM bytes of the dominant kernel from the committed ncu capture (profiles/traffic.json), or None when
    the capture is of another kernel or shape than the one timed here (a stale figure is worse than none)."""
    try:
        with open(os.path.join(ROOT, "profiles", "traffic.json")) as f:
            t = json.load(f).get(preset, {})
        if t.get("kernel") not in (None, KERNEL[preset]) or (t.get("shape") is not None and list(t["shape"]) != list(shape)):
            return None
        return t.get("dram_bytes_per_launch")
    except Exception:  # noqa: BLE001
        return None


def run_step_leg(a, preset, steps, warmup, device, rank, world, barrier, keep=False):
    """K timed steps of the quantizer step; returns (result dict, workload or None)."""
    import torch.distributed as dist

    from modelopt_b200 import ops

    wl = DeviceWorkload(a, device, rank, preset)
    for _ in range(warmup):
        wl.step()
    barrier()
    wl.kernel_events = []
    launches0 = ops.LAUNCHES[0]
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    with Clocks(device.index) as clk:
        barrier()
        e0.record()
        for _ in range(steps):
            wl.step()
        e1.record()
        barrier()
    launches = ops.LAUNCHES[0] - launches0
    t = torch.tensor([e0.elapsed_time(e1) / steps], dtype=torch.float64, device=device)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms = float(t.item())
    kms = [s.elapsed_time(e) for s, e in wl.kernel_events]  # dominant kernel: the fake-quant of the down_proj input
    k_avg_ms = sum(kms) / len(kms)
    k_bytes = 4 * wl.x[wl.big].numel()
    wl.kernel_events = None
    peak, peak_src, _ = hbm_peak()
    achieved = k_bytes / (k_avg_ms * 1e-3) / 1e9
    shape = list(wl.x[wl.big].shape)
    roofline = {"bound": "hbm", "kernel": KERNEL[preset], "shape": shape, "achieved": achieved, "peak": peak,
                "unit": "GB/s", "frac": achieved / peak, "frac_of_nominal_8tbs": achieved / 8000.0, "peak_source": peak_src,
                "algorithmic_bytes_per_launch": k_bytes, "avg_launch_ms": k_avg_ms, "launches_timed": len(kms),
                "traffic": read_traffic(preset, shape),
                "whole_step": {"algorithmic_bytes": wl.algorithmic_bytes(),
                               "achieved": wl.algorithmic_bytes() / (ms * 1e-3) / 1e9,
                               "frac": wl.algorithmic_bytes() / (ms * 1e-3) / 1e9 / peak}}
    res = {"value": a.tokens * world / (ms * 1e-3), "ms_per_step": ms, "steps": steps, "warmup": warmup,
           "roofline": roofline, "gpu_launches": launches, "clocks": clk.summary(), "workload": workload_name(a, preset)}
    if keep:
        return res, wl
    del wl
    gc.collect()
    torch.cuda.empty_cache()
    return res, None


# =================================================================================================
# e2e: the same step through the HOST-buffer C ABI
# =================================================================================================
class HostWorkload:
    """Activations in pinned host memory (uploaded every step), weights resident on the GPU (uploaded once, before the
    timed region: in every real use they already live there), results read back = the running amax of every quantizer."""

    def __init__(self, a, rank, weights, preset):
        from modelopt_b200 import ops

        self.ops, self.a, self.preset = ops, a, preset
        g = torch.Generator().manual_seed(4321 + rank)
        bf = torch.bfloat16

        def pinned(shape):
            t = torch.empty(shape, dtype=bf).pin_memory()
            blk = torch.randn(1 << 20, generator=g).to(bf)  # cheap fill: a random block tiled over the tensor
            flat = t.view(-1)
            for off in range(0, flat.numel(), blk.numel()):
                n = min(blk.numel(), flat.numel() - off)
                flat[off:off + n] = blk[:n]
            return t

        self.x = [pinned((a.tokens, ci)) for _, _, ci in a.linears]
        self.w = weights
        self.srcs = [t for layer in range(a.layers) for i in range(len(a.linears)) for t in (self.x[i], self.w[layer][i])]
        self.h2d = 2 * sum(t.numel() for t in self.x) * a.layers  # activations only; weights are resident
        self.running = [0.0] * len(self.srcs)  # MaxCalibrator state per quantizer
        self.d2h = 4 * len(self.srcs)
        self.last = None

    def step(self):
        # collect (running max) + fake-quant of the step's 2 x 7 x layers tensors in ONE pipelined call: the upload of
        # activation i+1 overlaps the amax + fake-quant of tensor i; the fake-quantized tensors stay on the device
        self.ops.calib_qdq_host_batch(self.srcs, self.running, self.preset, 16, None, download=False)
        self.last = float(sum(self.running))  # the D2H result (the amax words) is read on the host

    def copy_only(self, device):
        """The copy ceiling: the same activation bytes, plain pinned H2D copies on one stream, nothing else."""
        bufs = [torch.empty_like(t, device=device) for t in self.x]
        for _ in range(self.a.layers):
            for src, dst in zip(self.x, bufs):
                dst.copy_(src, non_blocking=True)


def run_e2e(a, device, rank, world, barrier, weights, preset):
    import torch.distributed as dist

    hw = HostWorkload(a, rank, weights, preset)
    hw.step()  # warm-up (allocates the library's pipeline buffers)

    def timed(fn, n):
        barrier()
        t0 = time.perf_counter()
        for _ in range(n):
            fn()
        torch.cuda.synchronize()
        dt = torch.tensor([(time.perf_counter() - t0) / n], dtype=torch.float64, device=device)
        if world > 1:
            dist.all_reduce(dt, op=dist.ReduceOp.MAX)
        return float(dt.item())

    s = timed(hw.step, a.e2e_steps)
    hw.copy_only(device)
    c = timed(lambda: hw.copy_only(device), 1)
    ceiling = hw.h2d / c / 1e9
    return {"value": a.tokens * world / s, "unit": UNIT, "h2d_bytes_per_step": hw.h2d, "d2h_bytes_per_step": hw.d2h,
            "ms_per_step": s * 1e3, "steps": a.e2e_steps, "h2d_gbs_per_gpu": hw.h2d / s / 1e9,
            "copy_ceiling_gbs": ceiling, "copy_ceiling_note": "plain pinned H2D copy of the same activation bytes on every rank "
            "at the same time, GB/s per GPU", "frac_of_copy_ceiling": (hw.h2d / s / 1e9) / ceiling,
            "api": "mo_calib_qdq_host_batch (include/modelopt_b200.h): collect + fake-quant of the step's tensors in one pipelined "
                   "call; activations from pinned host memory every step, weights resident on the GPU, the fake-quantized "
                   "tensors stay on the device, the running amax words return to the host"}


# =================================================================================================
# histogram workload: HistogramCalibrator.collect through the calibrator class
# =================================================================================================
def run_hist(a, device, rank, world, barrier, steps=20):
    import torch.distributed as dist

    from modelopt_b200.calib import HistogramCalibrator

    g = torch.Generator(device=device).manual_seed(77 + rank)
    xs = [torch.randn((a.tokens, ci), generator=g, device=device, dtype=torch.bfloat16) for _, _, ci in a.linears]
    cals = [HistogramCalibrator(8, None, False) for _ in xs]
    for c, x in zip(cals, xs):  # first batch fixes the range; scale it so later batches stay inside it (steady state)
        c.collect(x * 1.5)
    for c, x in zip(cals, xs):
        c.collect(x)
    barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps):
        for c, x in zip(cals, xs):
            c.collect(x)
    e1.record()
    barrier()
    t = torch.tensor([e0.elapsed_time(e1) / steps], dtype=torch.float64, device=device)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms = float(t.item())
    nbytes = 2 * sum(x.numel() for x in xs)
    peak, peak_src, _ = hbm_peak()
    gbs = nbytes / (ms * 1e-3) / 1e9
    return {"api": "HistogramCalibrator.collect (2048 bins, steady state: range fixed by the first batch)",
            "tensors": [list(x.shape) for x in xs], "ms_per_pass": ms, "passes": steps,
            "roofline": {"bound": "hbm", "kernel": "hist_kernel<bf16> (whole collect: kernel + host logic)", "achieved": gbs,
                         "peak": peak, "unit": "GB/s", "frac": gbs / peak, "algorithmic_bytes_per_pass": nbytes,
                         "peak_source": peak_src}}


# =================================================================================================
# CPU arm: the oracle port (torch restatement of the reference's CPU path) on the host cores
# =================================================================================================
CPU_ROW_CUT = 8  # the CPU sample keeps 1/8 of every tensor's rows of ONE layer


def host_threads():
    """All the host threads this process may use (torchrun exports OMP_NUM_THREADS=1, which would silently turn the CPU arm
    into a one-core run)."""
    try:
        n = len(os.sched_getaffinity(0))
    except Exception:  # noqa: BLE001
        n = os.cpu_count() or 1
    torch.set_num_threads(max(1, n))
    return torch.get_num_threads()


def reference_fp8():
    """The reference's own CPU functions for the FP8 path when the reference is importable (build container only), else None."""
    ref = "/root/reference"
    if not os.path.isdir(os.path.join(ref, "modelopt")):
        return None
    try:
        sys.path[:0] = [ref, os.path.join(ROOT, "oracle", "ref_shims")]
        from modelopt.torch.quantization import tensor_quant as rtq
        from modelopt.torch.quantization.utils import reduce_amax as rra

        return rtq, rra
    except Exception:  # noqa: BLE001
        return None


def cpu_sample_step(a, tensors, ref=None):
    """One pass of the same per-tensor work (collect + fake-quant) over the sample tensors, on CPU."""
    from oracle import amax as oamax
    from oracle import fake_quant as ofq

    for t in tensors:
        if ref is not None and a.workload == "fp8":
            am = ref[1](t).float()
            ref[0].scaled_e4m3(t, am, None, 4, 3)
            continue
        am = oamax.reduce_amax(t).float()
        if a.workload == "nvfp4":
            ofq.nvfp4_dynamic_qdq(t, am, 16)
        else:
            ofq.fp8_qdq(t, am)


def cpu_sample_tensors(a):
    g = torch.Generator().manual_seed(99)
    T = max(16, a.tokens // CPU_ROW_CUT)
    ts = [torch.randn((T, ci), generator=g).to(torch.bfloat16) for _, _, ci in a.linears]
    ts += [(torch.randn((max(16, co // CPU_ROW_CUT), ci), generator=g) * 0.02).to(torch.bfloat16) for _, co, ci in a.linears]
    return ts


def cpu_tokens_per_s(a, seconds_per_sample_step):
    # the sample is 1/CPU_ROW_CUT of one layer; a full step is CPU_ROW_CUT x layers of it
    return a.tokens / (seconds_per_sample_step * CPU_ROW_CUT * a.layers)


def cpu_kind(a, ref):
    if a.workload == "fp8" and ref is not None:
        return "reference", "the reference's own CPU path (scaled_e4m3 -> _fp8_eager, reduce_amax) imported from /root/reference"
    if a.workload == "nvfp4":
        return "port", "oracle/ torch restatement (the reference has no CPU NVFP4 path: IEEE restatement of its Triton kernel)"
    return "port", "oracle/ torch restatement of the reference's CPU path (_fp8_eager)"


def cpu_baseline(a, reps=1):
    cores = host_threads()
    ref = reference_fp8() if a.workload == "fp8" else None
    ts = cpu_sample_tensors(a)
    cpu_sample_step(a, ts[:2], ref)  # warm-up (allocator, thread pool)
    total, n = 0.0, 0
    while n < max(reps, 1) or (total < 10.0 and n < 64):  # about 10 s of CPU work; report the mean pass
        t0 = time.perf_counter()
        cpu_sample_step(a, ts, ref)
        total += time.perf_counter() - t0
        n += 1
    mean = total / n
    kind, path = cpu_kind(a, ref)
    return {"value": cpu_tokens_per_s(a, mean), "unit": UNIT, "cores": cores, "kind": kind,
            "sample": f"layer 0 of {a.layers}, every activation and weight cut to 1/{CPU_ROW_CUT} of its rows; {n} passes, "
                      f"{total:.1f} s of CPU work in total ({mean:.2f} s per pass, scaled x{CPU_ROW_CUT}x{a.layers}); {path}"}


def run_reference(a, rank):
    """--impl reference: the reference's CPU implementation of the path (oracle port; the real reference for FP8 when it
    is importable), all host threads, rank 0 only."""
    if rank != 0:
        return
    cores = host_threads()
    ref = reference_fp8() if a.workload == "fp8" else None
    ts = cpu_sample_tensors(a)
    warm = max(a.warmup, 3)
    for _ in range(warm):
        cpu_sample_step(a, ts[:4], ref)
    steps = max(1, a.steps)  # exactly K steps; each step is a bounded sample (under 1 s) so the run stays short
    t0 = time.perf_counter()
    for _ in range(steps):
        cpu_sample_step(a, ts, ref)
    per = (time.perf_counter() - t0) / steps
    v = cpu_tokens_per_s(a, per)
    kind, path = cpu_kind(a, ref)
    sample = (f"layer 0 of {a.layers}, every tensor cut to 1/{CPU_ROW_CUT} of its rows; {steps} sample steps of "
              f"{per:.2f} s on {cores} threads, scaled x{CPU_ROW_CUT}x{a.layers}; {path}")
    print(json.dumps({
        "impl": "reference", "metric": METRIC, "value": v, "unit": UNIT, "n_gpus": a.gpus, "steps": steps,
        "warmup": warm, "ms_per_step": per * CPU_ROW_CUT * a.layers * 1e3, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
        "config": {"workload": workload_name(a), "sample": sample},
        "cpu_baseline": {"value": v, "unit": UNIT, "cores": cores, "kind": kind, "sample": sample},
        "e2e": {"value": v, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }), flush=True)


# =================================================================================================
def main():
    a = parse()
    rank = int(os.environ.get("RANK", 0))
    world = int(os.environ.get("WORLD_SIZE", 1))
    local = int(os.environ.get("LOCAL_RANK", 0))
    if a.impl == "reference":
        run_reference(a, rank)
        return
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the product path has no CPU fallback "
                         "(use --impl reference for the CPU arm)")
    import torch.distributed as dist

    torch.cuda.set_device(local)
    device = torch.device("cuda", local)
    if world > 1:
        # one process per GPU: keep the rank's threads and its pinned host buffers (first touch) on the CPU cores next
        # to its GPU, so the host-buffer path does not cross sockets
        try:
            import pynvml

            pynvml.nvmlInit()
            pynvml.nvmlDeviceSetCpuAffinity(pynvml.nvmlDeviceGetHandleByIndex(local))
        except Exception:  # noqa: BLE001  (affinity is an optimisation only)
            pass
        dist.init_process_group("nccl", device_id=device)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    t_start = time.perf_counter()
    warm = max(a.warmup, 3)
    main_leg, wl = run_step_leg(a, a.workload, a.steps, warm, device, rank, world, barrier, keep=not a.no_e2e)

    # ---- e2e: the same step through the HOST-buffer C-ABI call (activations from pinned host memory) ----
    e2e = None
    if not a.no_e2e:
        weights = wl.w
        del wl
        gc.collect()
        torch.cuda.empty_cache()
        e2e = run_e2e(a, device, rank, world, barrier, weights, a.workload)
        del weights
        gc.collect()
        torch.cuda.empty_cache()

    # ---- the other configs as workloads of their own ----
    workloads = None
    if not a.no_workloads:
        import model_level as ml

        workloads = {}
        other = "fp8" if a.workload == "nvfp4" else "nvfp4"
        leg, _ = run_step_leg(a, other, max(3, min(a.steps, 10)), 3, device, rank, world, barrier)
        workloads[other] = {"metric": METRIC, "unit": UNIT, **leg}
        workloads["hist"] = run_hist(a, device, rank, world, barrier)
        gc.collect()
        torch.cuda.empty_cache()
        _, _, peaks = hbm_peak()
        awq = ml.awq_layer(device, rank, world)
        tpeak = float(peaks.get("bf16_tflops_sustained", 1394.4))
        awq["roofline"] = {"bound": "tensor", "kernel": "linear_pair_kernel<bf16, loss> (awq_lite search GEMMs)",
                           "achieved": awq["search_tflops"], "peak": tpeak, "unit": "TFLOP/s",
                           "frac": awq["search_tflops"] / tpeak,
                           "peak_source": "measured (MEASURED_PEAKS.json bf16_tflops_sustained)" if peaks else "fallback",
                           "note": "flops of the 11 candidate GEMMs / the whole search pass (which also runs the weight and "
                                   "activation scaling + fake-quant kernels and the reference GEMM)"}
        workloads["awq"] = awq
        gc.collect()
        torch.cuda.empty_cache()

    # ---- the configs as written: quantize(LlamaForCausalLM, cfg, forward_loop), strong scaling ----
    model_level = None
    if not a.no_model_level:
        import model_level as ml

        presets = tuple((a.model_presets or ("nvfp4,hist" if a.model == "70b" else "nvfp4,fp8,hist")).split(","))
        mb = a.micro_batch or (4 if a.model == "70b" else 8)
        with Clocks(local) as mclk:
            res = ml.run(a.model, presets, a.seqs, mb, 2048, device, rank, world)
        model_level = res[presets[0]]
        model_level["clocks"] = mclk.summary()
        for p in presets[1:]:
            model_level[p] = res[p]

    cpu = None
    if rank == 0 and world == 1 and not a.no_cpu_baseline:
        cpu = cpu_baseline(a)

    if rank == 0:
        print(json.dumps({
            "metric": METRIC, "value": main_leg["value"], "unit": UNIT, "n_gpus": world, "steps": a.steps,
            "warmup": warm, "ms_per_step": main_leg["ms_per_step"], "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
            "config": {"workload": workload_name(a), "preset": a.workload, "tokens_per_gpu_per_step": a.tokens,
                       "parallelism": f"dp{world}", "l2": "inputs larger than L2 (each tensor >= 134 MB, reuse "
                       "distance >= 1.7 GB); no flush needed"},
            "roofline": main_leg["roofline"], "cpu_baseline": cpu, "e2e": e2e, "gpu_launches": main_leg["gpu_launches"],
            "clocks": main_leg["clocks"], "workloads": workloads, "model_level": model_level,
            "bench_wall_s": time.perf_counter() - t_start,
        }), flush=True)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
