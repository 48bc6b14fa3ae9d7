"""bench.py — the quantizer hot path of a Llama-3-8B calibration micro-batch on B200.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference] [--workload nvfp4|fp8]

Workload (BASELINE.json configs[2], NVFP4_DEFAULT_CFG; `--workload fp8` = configs[1], FP8_DEFAULT_CFG):
one STEP = the work every TensorQuantizer of a random-init Llama-3-8B does for one micro-batch of
[8, 2048] = 16384 synthetic tokens, model GEMMs excluded (they are not on the path):

  calibration pass  : MaxCalibrator.collect on the input of each of the 7 x 32 quantized linears
                      (activation [16384, Cin] bf16 -> one amax word in the arena) and on each weight
                      (the reference re-collects weights on every forward);
  fake-quant pass   : TensorQuantizer._fake_quantize of the same activations and weights with the
                      calibrated amax (NVFP4: dynamic block-16 E4M3 scales under the fp32 global amax;
                      FP8: per-tensor E4M3) — what the post-calibration forward runs.

Data-parallel: every rank processes its own micro-batch (weak scaling), the only
exchange is ONE all-reduce(MAX) over the amax arena per step.  value = tokens of all ranks / max-over-
ranks device time.  Inputs exceed L2 (each tensor >= 134 MB, reuse distance > 1.2 GB), so no L2 flush.
"""

from __future__ import annotations

import argparse
import json
import os
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path[:0] = [ROOT, os.path.join(ROOT, "tensorrt-model-optimizer_b200")]

import torch  # noqa: E402

# Llama-3-8B: (name, Cout, Cin, which of the 4 per-layer activations feeds it)
LINEARS = [("q_proj", 4096, 4096), ("k_proj", 1024, 4096), ("v_proj", 1024, 4096), ("o_proj", 4096, 4096),
           ("gate_proj", 14336, 4096), ("up_proj", 14336, 4096), ("down_proj", 4096, 14336)]
N_LAYERS = 32
TOKENS = 8 * 2048
METRIC = "calib tokens/s"
UNIT = "tokens/s"


def parse():
    p = argparse.ArgumentParser()
    p.add_argument("--gpus", type=int, default=1)
    p.add_argument("--steps", type=int, default=10)
    p.add_argument("--warmup", type=int, default=3)
    p.add_argument("--impl", default="b200", choices=["b200", "reference"])
    p.add_argument("--workload", default="nvfp4", choices=["nvfp4", "fp8"])
    p.add_argument("--layers", type=int, default=N_LAYERS)
    p.add_argument("--tokens", type=int, default=TOKENS)
    p.add_argument("--e2e-steps", type=int, default=2)
    p.add_argument("--no-e2e", action="store_true")
    p.add_argument("--no-cpu-baseline", action="store_true")
    return p.parse_args()


def workload_name(a):
    cfg = "NVFP4_DEFAULT_CFG (block-16 E4M3 scales, fp32 global amax)" if a.workload == "nvfp4" else \
        "FP8_DEFAULT_CFG (per-tensor E4M3)"
    return (f"Llama-3-8B random-init {cfg} max calibration + fake-quant, quantizer hot path only, "
            f"micro-batch {a.tokens} tokens/GPU, {a.layers} layers x 7 linears")


# =================================================================================================
# clocks sampler (nvml, falls back to nvidia-smi)
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


# =================================================================================================
# the B200 arm
# =================================================================================================
class DeviceWorkload:
    """Device-resident tensors of one rank + the per-step launch sequence."""

    def __init__(self, a, device, rank):
        from modelopt_b200 import ops
        from modelopt_b200.arena import AmaxArena

        self.ops, self.a, self.device = ops, a, device
        g = torch.Generator(device=device).manual_seed(1234 + rank)
        bf = torch.bfloat16
        T = a.tokens
        # weights: every layer has its own (random-init, std 0.02)
        self.w = [[torch.randn((co, ci), generator=g, device=device, dtype=bf) * 0.02 for _, co, ci in LINEARS]
                  for _ in range(a.layers)]
        # activations: 7 distinct inputs per layer slot, reused across layers (reuse distance 1.27 GB >> L2)
        self.x = [torch.randn((T, ci), generator=g, device=device, dtype=bf) for _, _, ci in LINEARS]
        self.y = [torch.empty_like(t) for t in self.x]
        self.wy = [torch.empty_like(t) for t in self.w[0]]
        self.arena = AmaxArena(device, 8 * 2 * 7 * a.layers + 64)
        self.slot_x = [[self.arena.alloc(1) for _ in LINEARS] for _ in range(a.layers)]
        self.slot_w = [[self.arena.alloc(1) for _ in LINEARS] for _ in range(a.layers)]
        self.elems_act = sum(t.numel() for t in self.x) * a.layers
        self.elems_w = sum(t.numel() for layer in self.w for t in layer)
        self.big = max(range(len(LINEARS)), key=lambda i: self.x[i].numel())  # down_proj input [T, 14336]
        self.kernel_events = None  # when a list: (start, stop) pairs around the dominant kernel's launches

    def algorithmic_bytes(self):
        # collect: 2 B/elem (one bf16 read); fake-quant: 4 B/elem (read + write)   SURVEY.md 8(d)
        n = self.elems_act + self.elems_w
        return 2 * n + 4 * n

    def _qdq(self, x, amax_f32, out):
        if self.a.workload == "nvfp4":
            return self.ops.qdq_nvfp4_dynamic(x, amax_f32, 16, None, out=out)
        return self.ops.qdq_fp8(x, amax_f32, out=out)

    def step(self):
        ops, arena = self.ops, self.arena
        n = len(LINEARS)
        for layer in range(self.a.layers):
            # calibration pass (TensorQuantizer.collect -> MaxCalibrator.collect): 14 tensors, 1.7 GB
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


class HostWorkload:
    """The same step through the HOST-buffer C-ABI entry points (pinned host tensors in, host tensors out)."""

    def __init__(self, a, rank):
        from modelopt_b200 import ops

        self.ops, self.a = ops, a
        g = torch.Generator().manual_seed(4321 + rank)
        bf = torch.bfloat16
        T = a.tokens

        def pinned(shape, fill=True):
            t = torch.empty(shape, dtype=bf).pin_memory()
            if fill:  # cheap fill: a random block tiled over the tensor
                blk = torch.randn(1 << 20, generator=g).to(bf)
                flat = t.view(-1)
                for off in range(0, flat.numel(), blk.numel()):
                    n = min(blk.numel(), flat.numel() - off)
                    flat[off:off + n] = blk[:n]
            return t

        self.x = [pinned((T, ci)) for _, _, ci in LINEARS]
        self.y = [pinned((T, ci), False) for _, _, ci in LINEARS]
        self.w = [pinned((co, ci)) for _, co, ci in LINEARS]  # one layer's weights, re-sent for every layer
        self.wy = [pinned((co, ci), False) for _, co, ci in LINEARS]
        per_layer = sum(t.numel() for t in self.x) + sum(t.numel() for t in self.w)
        # one pipelined call per layer (mo_calib_qdq_host_batch): every tensor is uploaded ONCE (it stays resident on the
        # device between its amax and its fake-quant) and returned once
        self.h2d = 2 * per_layer * a.layers
        self.d2h = 2 * per_layer * a.layers
        self.running = [0.0] * (2 * len(LINEARS) * a.layers)  # MaxCalibrator state per quantizer
        self.last = None

    def step(self):
        ops = self.ops
        kind = "nvfp4" if self.a.workload == "nvfp4" else "fp8"
        srcs = [t for i in range(len(LINEARS)) for t in (self.x[i], self.w[i])]
        dsts = [t for i in range(len(LINEARS)) for t in (self.y[i], self.wy[i])]
        # collect (running max) + fake-quant of every layer's 14 tensors in ONE pipelined call: the upload of tensor i+1
        # overlaps the fake-quant and download of tensor i, with no bubble between layers
        ops.calib_qdq_host_batch(srcs * self.a.layers, self.running, kind, 16, None, outs=dsts * self.a.layers)
        self.last = float(self.wy[-1].view(-1)[0])  # the D2H result is read on the host


# =================================================================================================
# CPU arm: the oracle port (torch restatement of the reference's CPU path) on the host cores
# =================================================================================================
CPU_ROW_CUT = 8  # the CPU sample keeps 1/8 of every tensor's rows of ONE layer


def cpu_sample_step(a, tensors):
    """One pass of the same per-tensor work (collect + fake-quant) over the sample tensors, on CPU."""
    from oracle import amax as oamax
    from oracle import fake_quant as ofq

    for t in tensors:
        am = oamax.reduce_amax(t).float()
        if a.workload == "nvfp4":
            ofq.nvfp4_dynamic_qdq(t, am, 16)
        else:
            ofq.fp8_qdq(t, am)


def cpu_sample_tensors(a):
    g = torch.Generator().manual_seed(99)
    T = max(16, a.tokens // CPU_ROW_CUT)
    ts = [torch.randn((T, ci), generator=g).to(torch.bfloat16) for _, _, ci in LINEARS]
    ts += [(torch.randn((max(16, co // CPU_ROW_CUT), ci), generator=g) * 0.02).to(torch.bfloat16) for _, co, ci in LINEARS]
    return ts


def cpu_tokens_per_s(a, seconds_per_sample_step):
    # the sample is 1/CPU_ROW_CUT of one layer; a full step is CPU_ROW_CUT x layers of it
    return a.tokens / (seconds_per_sample_step * CPU_ROW_CUT * a.layers)


def cpu_baseline(a, reps=1):
    ts = cpu_sample_tensors(a)
    cpu_sample_step(a, ts[:2])  # warm-up (allocator, thread pool)
    # repeat the sample until about 10 s of CPU work has been timed (at least `reps` passes); report the mean pass
    total, n = 0.0, 0
    while n < max(reps, 1) or (total < 10.0 and n < 64):
        t0 = time.perf_counter()
        cpu_sample_step(a, ts)
        total += time.perf_counter() - t0
        n += 1
    mean = total / n
    path = ("the reference has no CPU NVFP4 path: IEEE restatement of its Triton kernel" if a.workload == "nvfp4"
            else "the reference's own CPU path (_fp8_eager)")
    return {"value": cpu_tokens_per_s(a, mean), "unit": UNIT, "cores": torch.get_num_threads(), "kind": "port",
            "sample": f"layer 0 of {a.layers}, every activation and weight cut to 1/{CPU_ROW_CUT} of its rows; {n} passes, "
                      f"{total:.1f} s of CPU work in total ({mean:.2f} s per pass, scaled x{CPU_ROW_CUT}x{a.layers}); oracle/ "
                      f"torch restatement ({path})"}


def run_reference(a, rank):
    """--impl reference: the reference's CPU implementation of the path (oracle port), all host threads."""
    if rank != 0:
        return
    ts = cpu_sample_tensors(a)
    warm = max(1, min(a.warmup, 3))
    for _ in range(warm):
        cpu_sample_step(a, ts[:2])
    steps = max(1, min(a.steps, 50))  # exactly K steps; each step is a bounded sample (about 1 s) so the run stays short
    t0 = time.perf_counter()
    for _ in range(steps):
        cpu_sample_step(a, ts)
    per = (time.perf_counter() - t0) / steps
    v = cpu_tokens_per_s(a, per)
    sample = (f"layer 0 of {a.layers}, every tensor cut to 1/{CPU_ROW_CUT} of its rows; {steps} sample steps of "
              f"{per:.2f} s, scaled x{CPU_ROW_CUT}x{a.layers}")
    print(json.dumps({
        "impl": "reference", "metric": METRIC, "value": v, "unit": UNIT, "n_gpus": a.gpus, "steps": steps,
        "warmup": warm, "ms_per_step": per * CPU_ROW_CUT * a.layers * 1e3, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
        "config": {"workload": workload_name(a), "sample": sample},
        "cpu_baseline": {"value": v, "unit": UNIT, "cores": torch.get_num_threads(), "kind": "port", "sample": sample},
        "e2e": {"value": v, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }), flush=True)


# =================================================================================================
def read_traffic(workload):
    """Per-launch DRAM bytes of the dominant kernel from the committed ncu capture (profiles/), or None."""
    try:
        with open(os.path.join(ROOT, "profiles", "traffic.json")) as f:
            return json.load(f).get(workload, {}).get("dram_bytes_per_launch")
    except Exception:  # noqa: BLE001
        return None


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
    from modelopt_b200 import ops

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    wl = DeviceWorkload(a, device, rank)
    for _ in range(max(a.warmup, 3)):
        wl.step()
    barrier()
    wl.kernel_events = []
    launches0 = ops.LAUNCHES[0]
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    with Clocks(local) as clk:
        barrier()
        e0.record()
        for _ in range(a.steps):
            wl.step()
        e1.record()
        barrier()
    launches = ops.LAUNCHES[0] - launches0
    ms = e0.elapsed_time(e1) / a.steps
    t = torch.tensor([ms], dtype=torch.float64, device=device)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms = float(t.item())
    # dominant kernel: the fake-quant of the down_proj input [tokens, 14336]
    kms = [s.elapsed_time(e) for s, e in wl.kernel_events]
    k_avg_ms = sum(kms) / len(kms)
    k_bytes = 4 * wl.x[wl.big].numel()
    wl.kernel_events = None
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        peak, peak_src = float(peaks["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    except Exception:  # noqa: BLE001
        peak, peak_src = 6650.0, "fallback (B200_PROFILING.md)"
    achieved = k_bytes / (k_avg_ms * 1e-3) / 1e9
    kname = "nvfp4_dynamic_kernel<bf16,16>" if a.workload == "nvfp4" else "qdq_stream_kernel<bf16,Fp8Op>"
    roofline = {"bound": "hbm", "kernel": kname, "shape": list(wl.x[wl.big].shape), "achieved": achieved, "peak": peak,
                "unit": "GB/s", "frac": achieved / peak, "peak_source": peak_src,
                "algorithmic_bytes_per_launch": k_bytes, "avg_launch_ms": k_avg_ms, "launches_timed": len(kms),
                "traffic": read_traffic(a.workload),
                "whole_step": {"algorithmic_bytes": wl.algorithmic_bytes(),
                               "achieved": wl.algorithmic_bytes() / (ms * 1e-3) / 1e9,
                               "frac": wl.algorithmic_bytes() / (ms * 1e-3) / 1e9 / peak}}
    value = a.tokens * world / (ms * 1e-3)

    # ---- e2e: the same step through the HOST-buffer C-ABI calls (pinned host in, host out) ----------
    e2e = None
    if not a.no_e2e:
        del wl
        torch.cuda.empty_cache()
        hw = HostWorkload(a, rank)
        hw.step()  # warm-up (allocates the library's staging buffers)
        barrier()
        t0 = time.perf_counter()
        for _ in range(a.e2e_steps):
            hw.step()
        torch.cuda.synchronize()
        dt = torch.tensor([(time.perf_counter() - t0) / a.e2e_steps], dtype=torch.float64, device=device)
        if world > 1:
            dist.all_reduce(dt, op=dist.ReduceOp.MAX)
        e2e = {"value": a.tokens * world / float(dt.item()), "unit": UNIT, "h2d_bytes_per_step": hw.h2d,
               "d2h_bytes_per_step": hw.d2h, "ms_per_step": float(dt.item()) * 1e3, "steps": a.e2e_steps,
               "api": "mo_calib_qdq_host_batch (include/modelopt_b200.h): collect + fake-quant of the step's 448 pinned host "
                      "tensors in one pipelined call, one upload and one download per tensor"}
        del hw

    cpu = None
    if rank == 0 and world == 1 and not a.no_cpu_baseline:
        cpu = cpu_baseline(a)

    if rank == 0:
        print(json.dumps({
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": a.steps,
            "warmup": max(a.warmup, 3), "ms_per_step": ms, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
            "config": {"workload": workload_name(a), "preset": a.workload, "tokens_per_gpu_per_step": a.tokens,
                       "parallelism": f"dp{world}", "l2": "inputs larger than L2 (each tensor >= 134 MB, reuse "
                       "distance >= 1.7 GB); no flush needed"},
            "roofline": roofline, "cpu_baseline": cpu, "e2e": e2e, "gpu_launches": launches,
            "clocks": clk.summary(),
        }), flush=True)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
