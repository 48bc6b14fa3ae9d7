"""Model-level calibration through the drop-in API: `modelopt_b200.quantize(model, cfg, forward_loop)` on a random-init
Llama (BASELINE.json configs[1], [2], [4]) — what SURVEY.md 8(d) calls the model-level measurement.

The work is the reference's contract (model_quant.py:147 `quantize`, model_calib.py:311-498 `max_calibrate`,
nn/modules/tensor_quantizer.py:1119-1221 `TensorQuantizer.forward`): insert quantizers, run `forward_loop` over the
calibration set (512 sequences x 2048 synthetic tokens, split evenly over the data-parallel ranks = STRONG scaling,
micro-batch 8), finish.  Reported beside it: the plain bf16 forward over the same micro-batches (the floor: calibration
"must add < a few %"), the post-calibration fake-quant forward, and the host cost per TensorQuantizer call.

Every micro-batch's token ids are copied from pinned host memory inside the timed region and a checksum of its logits is
copied back, so the numbers are end to end.

Stand-alone:  python tools/model_level.py [--model 8b|70b|tiny] [--preset nvfp4|fp8] [--seqs 512] [--micro-batch 8]
(under torchrun for N > 1).  bench.py imports `run()`.
"""

from __future__ import annotations

import argparse
import copy
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "tensorrt-model-optimizer_b200")]

import torch  # noqa: E402

# SURVEY.md section 8 config table
MODELS = {
    "8b": dict(hidden_size=4096, intermediate_size=14336, num_hidden_layers=32, num_attention_heads=32,
               num_key_value_heads=8, vocab_size=128256),
    "70b": dict(hidden_size=8192, intermediate_size=28672, num_hidden_layers=80, num_attention_heads=64,
                num_key_value_heads=8, vocab_size=128256),
    "tiny": dict(hidden_size=256, intermediate_size=512, num_hidden_layers=2, num_attention_heads=8,
                 num_key_value_heads=2, vocab_size=1024),
}
NAMES = {"8b": "Llama-3-8B", "70b": "Llama-3-70B", "tiny": "Llama-tiny (test)"}
PRESETS = {"nvfp4": "NVFP4_DEFAULT_CFG", "fp8": "FP8_DEFAULT_CFG", "int8": "INT8_DEFAULT_CFG"}
SEQ_LEN = 2048


def build_model(name: str, device, seed: int = 0, layers: int | None = None):
    """Random-init HF Llama from its config only (no checkpoint), bf16, on `device`."""
    from transformers import LlamaConfig, LlamaForCausalLM

    kw = dict(MODELS[name])
    if layers:
        kw["num_hidden_layers"] = layers
    cfg = LlamaConfig(max_position_embeddings=8192, rope_theta=500000.0, rms_norm_eps=1e-5, tie_word_embeddings=False, **kw)
    torch.manual_seed(seed)
    with torch.device(device):
        model = LlamaForCausalLM._from_config(cfg, dtype=torch.bfloat16)
    return model.eval()


def n_params(model):
    return sum(p.numel() for p in model.parameters())


class Feeder:
    """The rank's share of the calibration set in pinned host memory; `loop(model)` is the user's forward_loop."""

    def __init__(self, vocab, n_seq, micro_batch, seq_len, rank, device):
        g = torch.Generator().manual_seed(1234 + rank)
        self.ids = torch.randint(0, vocab, (n_seq, seq_len), generator=g).pin_memory()
        self.mb, self.device = micro_batch, device
        self.n_mb = (n_seq + micro_batch - 1) // micro_batch
        self.out = torch.zeros(self.n_mb, 8, dtype=torch.float32).pin_memory()
        self.h2d = self.d2h = 0

    def loop(self, model, limit=None):
        with torch.no_grad():
            for i in range(self.n_mb if limit is None else min(limit, self.n_mb)):
                ids = self.ids[i * self.mb:(i + 1) * self.mb].to(self.device, non_blocking=True)
                logits = model(input_ids=ids, use_cache=False).logits
                self.out[i].copy_(logits[0, -1, :8].float(), non_blocking=True)  # the step's result, read on the host
                self.h2d += ids.numel() * ids.element_size()
                self.d2h += 32
                del logits


class CollectiveLog:
    """Counts the torch.distributed.all_reduce calls the product issues inside a `with` block (op, calls, bytes): the
    evidence that the data-parallel exchanges (amax MAX, histogram MAX / SUM) ran over the process group."""

    def __init__(self):
        self.ops = {}

    def __enter__(self):
        import torch.distributed as dist

        self._dist, self._orig = dist, dist.all_reduce

        def logged(tensor, op=dist.ReduceOp.SUM, *a, **k):
            name = str(op).split(".")[-1].upper()
            e = self.ops.setdefault(name, {"calls": 0, "bytes": 0})
            e["calls"] += 1
            e["bytes"] += tensor.numel() * tensor.element_size()
            return self._orig(tensor, op, *a, **k)

        dist.all_reduce = logged
        return self

    def __exit__(self, *exc):
        self._dist.all_reduce = self._orig

    def summary(self, world):
        dist = self._dist
        on = dist.is_available() and dist.is_initialized()
        return {"backend": dist.get_backend() if on else None, "world_size": dist.get_world_size() if on else world,
                "all_reduce": self.ops}


def _timed(fn, device, dist_world):
    """Device time (CUDA events on the current stream) and wall time of fn(), both max over ranks; barrier + sync
    on both sides."""
    import torch.distributed as dist

    def barrier():
        if dist_world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    t0 = time.perf_counter()
    e0.record()
    r = fn()
    e1.record()
    torch.cuda.synchronize()
    wall = time.perf_counter() - t0
    barrier()
    t = torch.tensor([e0.elapsed_time(e1) * 1e-3, wall], dtype=torch.float64, device=device)
    if dist_world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    dev_s, wall_s = t.tolist()
    return r, dev_s, wall_s


def host_us_per_quantizer_call(model, device, reps=20):
    """Host (Python + ctypes) time of one enabled TensorQuantizer.forward in the calibrated model: a forward over 8 tokens
    keeps the GPU work negligible, so wall time / calls is the per-call host cost (launch latency included)."""
    from modelopt_b200.nn.tensor_quantizer import TensorQuantizer

    ids = torch.randint(0, 1000, (1, 8), device=device)
    calls = [0]
    hooks = [m.register_forward_pre_hook(lambda *_: calls.__setitem__(0, calls[0] + 1))
             for m in model.modules() if isinstance(m, TensorQuantizer) and m.is_enabled]
    with torch.no_grad():
        model(input_ids=ids, use_cache=False)
    for h in hooks:
        h.remove()
    n_calls = calls[0]

    def run():
        with torch.no_grad():
            for _ in range(reps):
                model(input_ids=ids, use_cache=False)
        torch.cuda.synchronize()

    run()
    t0 = time.perf_counter()
    run()
    return (time.perf_counter() - t0) / reps, n_calls


def awq_layer(device, rank=0, world=1, n_micro_batches=2, micro_batch=8, seq_len=SEQ_LEN):
    """BASELINE.json configs[3] on ONE full Llama-3-8B decoder layer: `quantize(layer_model, INT4_AWQ_CFG, forward_loop)`
    = awq_lite (model_calib.py:1395-1721): cache pass (act_scale per linear) + search pass (per linear and micro-batch:
    11 x (x / s, block-128 INT4 fake-quant of W * s, GEMM + fused loss) + the reference GEMM).  Every rank does the same
    work (replicas); the time is the max over ranks."""
    import modelopt_b200 as mtq
    from modelopt_b200 import config as qcfg
    from modelopt_b200 import ops

    model = build_model("8b", device, seed=0, layers=1).model  # LlamaModel: embedding + 1 decoder layer + norm, no lm_head
    g = torch.Generator().manual_seed(4321 + rank)
    ids = [torch.randint(0, model.config.vocab_size, (micro_batch, seq_len), generator=g).to(device)
           for _ in range(n_micro_batches)]
    passes = []

    def loop(m):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        with torch.no_grad():
            for b in ids:
                m(input_ids=b, use_cache=False)
        e1.record()
        passes.append((e0, e1))

    with torch.no_grad():
        for b in ids[:1]:
            model(input_ids=b, use_cache=False)  # warm-up of the plain layer
    launches0 = ops.LAUNCHES[0]
    _, dev_s, wall_s = _timed(lambda: mtq.quantize(model, qcfg.INT4_AWQ_CFG, loop), device, world)
    launches = ops.LAUNCHES[0] - launches0
    assert len(passes) == 2, "awq_lite runs forward_loop twice (cache, search)"
    cache_ms, search_ms = (a.elapsed_time(b) for a, b in passes)
    from modelopt_b200.utils import is_quantized_linear

    lin = [m for m in model.modules() if is_quantized_linear(m) and m.weight_quantizer.is_enabled]
    macs = sum(m._parameters["weight"].numel() for m in lin)
    T = micro_batch * seq_len
    n_alpha = 11
    search_flops = n_alpha * 2.0 * T * macs * n_micro_batches  # the candidate GEMMs (loss epilogue, output never written)
    t = torch.tensor([search_ms], dtype=torch.float64, device=device)
    if world > 1:
        torch.distributed.all_reduce(t, op=torch.distributed.ReduceOp.MAX)
    search_ms = float(t.item())
    pqs = [m.input_quantizer._pre_quant_scale for m in lin if hasattr(m.input_quantizer, "_pre_quant_scale")]
    assert len(pqs) == len(lin) and all(bool(torch.isfinite(p.float()).all()) for p in pqs), "awq_lite left no pre_quant_scale"
    return {
        "api": "modelopt_b200.quantize(model, INT4_AWQ_CFG, forward_loop)  [awq_lite, alpha_step 0.1, block-128 INT4]",
        "model": "one Llama-3-8B decoder layer (7 linears, %.1f M weights), random-init bf16" % (macs / 1e6),
        "tokens": T * n_micro_batches, "micro_batches": n_micro_batches,
        "quantize_wall_s": wall_s, "quantize_device_s": dev_s, "cache_pass_ms": cache_ms, "search_pass_ms": search_ms,
        "search_gemm_flops": search_flops,
        "search_tflops": search_flops / (search_ms * 1e-3) / 1e12,
        "tokens_per_s_per_layer": T * n_micro_batches / wall_s,
        "linears": len(lin), "gpu_launches": launches,
    }


def run(model_name="8b", presets=("nvfp4",), n_seq_total=512, micro_batch=8, seq_len=SEQ_LEN, device=None, rank=0, world=1,
        layers=None, post_micro_batches=4, plain_micro_batches=16, check=True):
    """Model-level measurement on this rank's GPU; returns {preset: dict} (rank 0's view, times = max over ranks).
    The plain forward is measured once and shared by the presets; the model is rebuilt (same seed) per preset."""
    import gc

    import modelopt_b200 as mtq
    from modelopt_b200 import config as qcfg
    from modelopt_b200 import ops
    from modelopt_b200.nn.tensor_quantizer import TensorQuantizer

    device = device or torch.device("cuda", torch.cuda.current_device())
    assert n_seq_total % world == 0, "the calibration set must split evenly over the ranks"
    n_seq = n_seq_total // world
    out = {}
    plain = None
    for preset in presets:
        t0 = time.perf_counter()
        model = build_model(model_name, device, seed=0, layers=layers)
        torch.cuda.synchronize()
        build_s = time.perf_counter() - t0
        feeder = Feeder(model.config.vocab_size, n_seq, micro_batch, seq_len, rank, device)
        tokens_all = n_seq * seq_len * world
        ids8 = torch.randint(0, 1000, (1, 8), device=device)

        def tiny(reps=20, model=model, ids8=ids8):
            with torch.no_grad():
                for _ in range(reps):
                    model(input_ids=ids8, use_cache=False)
            torch.cuda.synchronize()

        if plain is None:
            # ---- plain bf16 forward over the same micro-batches (the floor) ----
            feeder.loop(model, limit=2)  # warm-up: cuBLAS handles, SDPA kernel choice, allocator pools
            n_plain = feeder.n_mb if plain_micro_batches is None else min(plain_micro_batches, feeder.n_mb)
            _, plain_s, _ = _timed(lambda: feeder.loop(model, limit=n_plain), device, world)
            tiny()
            t0 = time.perf_counter()
            tiny()
            plain = {"s_per_mb": plain_s / n_plain, "n": n_plain, "host_s": (time.perf_counter() - t0) / 20}
        else:
            feeder.loop(model, limit=1)
        plain_tok_s = micro_batch * seq_len * world / plain["s_per_mb"]
        plain_equiv_s = plain["s_per_mb"] * feeder.n_mb  # plain forward over the rank's whole share

        # ---- the user's call: quantize(model, cfg, forward_loop) ----
        launches0 = ops.LAUNCHES[0]
        feeder.h2d = feeder.d2h = 0
        torch.cuda.reset_peak_memory_stats(device)
        if preset == "hist":
            # histogram calibration of every linear input (INT8 per tensor, entropy search): the reference's explicit flow
            # enable_stats_collection -> forward_loop -> finish_stats_collection(method="entropy") (model_calib.py:1128-1167,
            # calib/histogram.py); data-parallel: the batch maximum all-reduced(MAX) inside collect, counts all-reduced(SUM)
            from modelopt_b200 import model_calib

            hist_cfg = copy.deepcopy(qcfg.INT8_DEFAULT_CFG["quant_cfg"])
            for e in hist_cfg:
                if e["quantizer_name"] == "*weight_quantizer":
                    e.pop("cfg", None)
                    e["enable"] = False
                elif e["quantizer_name"] == "*input_quantizer":
                    e["cfg"] = {"num_bits": 8, "axis": None, "calibrator": "histogram"}

            def hist_calibrate():
                mtq.replace_quant_module(model)
                mtq.set_quantizer_by_cfg(model, hist_cfg)
                model_calib.enable_stats_collection(model)
                feeder.loop(model)
                model_calib.finish_stats_collection(model, method="entropy")

            with CollectiveLog() as clog:
                _, quant_s, quant_wall = _timed(hist_calibrate, device, world)
        else:
            cfg = getattr(qcfg, PRESETS[preset])
            with CollectiveLog() as clog:
                _, quant_s, quant_wall = _timed(lambda: mtq.quantize(model, cfg, feeder.loop), device, world)
        collectives = clog.summary(world)
        # _timed's own max-over-ranks reduction of the two clocks is not the product's
        if world > 1 and "MAX" in collectives["all_reduce"]:
            e = collectives["all_reduce"]["MAX"]
            e["calls"] -= 1
            e["bytes"] -= 16
            if e["calls"] == 0:
                del collectives["all_reduce"]["MAX"]
        launches = ops.LAUNCHES[0] - launches0
        h2d, d2h = feeder.h2d, feeder.d2h
        quantizers = [m for m in model.modules() if isinstance(m, TensorQuantizer)]
        enabled = [q for q in quantizers if q.is_enabled]
        with_amax = [q for q in enabled if getattr(q, "_amax", None) is not None]
        if check:
            assert with_amax and len(with_amax) == len(enabled), "calibration left an enabled quantizer without amax"
            am = torch.stack([q._amax.float().max() for q in with_amax])
            assert bool(torch.isfinite(am).all()) and bool((am > 0).all()), "non-finite or zero amax after calibration"

        # ---- post-calibration forward (fake-quant mode) ----
        feeder.loop(model, limit=1)  # fills the weight fake-quant caches (the reference re-quantizes per forward)
        n_post = min(post_micro_batches, feeder.n_mb)
        _, post_s, _ = _timed(lambda: feeder.loop(model, limit=n_post), device, world)
        post_tok_s = n_post * micro_batch * seq_len * world / post_s
        q_host_s, n_calls = host_us_per_quantizer_call(model, device)
        peak_mem = torch.cuda.max_memory_allocated(device)
        out[preset] = {
            "api": ("enable_stats_collection -> forward_loop -> finish_stats_collection(method='entropy'), "
                    "HistogramCalibrator (INT8 per tensor) on every quantized linear's input" if preset == "hist"
                    else "modelopt_b200.quantize(model, %s, forward_loop)" % PRESETS[preset]),
            "model": f"{NAMES[model_name]} random-init bf16, {n_params(model) / 1e9:.2f} B parameters"
                     + (f" ({layers} layers)" if layers else ""),
            "preset": PRESETS.get(preset, "INT8 activations, histogram calibrator, entropy"),
            "calibration_set": f"{n_seq_total} sequences x {seq_len} tokens, micro-batch {micro_batch}, "
                               f"{n_seq} sequences per rank (strong scaling over {world} ranks)",
            "tokens": tokens_all,
            "tokens_per_s": tokens_all / quant_wall,  # wall clock: the call's host work (conversion, finish) is in it
            "quantize_wall_s": quant_wall,
            "quantize_device_s": quant_s,
            "forward_only_tokens_per_s": plain_tok_s,
            "forward_only_s": plain_equiv_s,
            "forward_only_micro_batches_timed": plain["n"],
            "overhead_pct": 100.0 * (quant_wall - plain_equiv_s) / plain_equiv_s,
            "fake_quant_forward_tokens_per_s": post_tok_s,
            "fake_quant_forward_vs_plain_pct": 100.0 * (plain_tok_s / post_tok_s - 1.0),
            "quantizers": {"inserted": len(quantizers), "enabled": len(enabled), "calibrated": len(with_amax)},
            "quantizer_calls_per_forward": n_calls,
            # None when the 8-token probe took the fused-prologue linear (per-tensor FP8 / INT8: no quantizer call at all)
            "host_us_per_quantizer_call": 1e6 * (q_host_s - plain["host_s"]) / n_calls if n_calls else None,
            "host_ms_per_forward": {"plain": plain["host_s"] * 1e3, "quantized": q_host_s * 1e3},
            "gpu_launches_in_quantize": launches,
            "h2d_bytes": h2d, "d2h_bytes": d2h,
            "peak_mem_gb": peak_mem / 1e9,
            "build_s": build_s,
            "n_gpus": world,
            "collectives_in_quantize": collectives,
            "scaling": "strong",
            "plain_gemm_backend": ops.PLAIN_GEMM["backend"],
        }
        del model, feeder, quantizers, enabled, with_amax, tiny  # `tiny` holds the model in a default argument
        gc.collect()
        torch.cuda.empty_cache()
    return out


def main():
    p = argparse.ArgumentParser()
    p.add_argument("--model", default="8b", choices=list(MODELS))
    p.add_argument("--preset", default="nvfp4", help="comma list of " + "|".join(PRESETS) + ", or awq (one decoder layer)")
    p.add_argument("--seqs", type=int, default=512)
    p.add_argument("--micro-batch", type=int, default=8)
    p.add_argument("--seq-len", type=int, default=SEQ_LEN)
    p.add_argument("--layers", type=int, default=None)
    a = p.parse_args()
    rank, world = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
    local = int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(local)
    device = torch.device("cuda", local)
    if world > 1:
        import torch.distributed as dist

        dist.init_process_group("nccl", device_id=device)
    if a.preset == "awq":
        r = awq_layer(device, rank, world, micro_batch=a.micro_batch, seq_len=a.seq_len)
    else:
        r = run(a.model, tuple(a.preset.split(",")), a.seqs, a.micro_batch, a.seq_len, device, rank, world, a.layers)
    if rank == 0:
        print(json.dumps(r), flush=True)
    if world > 1:
        torch.distributed.destroy_process_group()


if __name__ == "__main__":
    main()
