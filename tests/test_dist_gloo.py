"""CPU, world size 2 (gloo): the data-parallel exchange of calibration statistics — one all-reduce(MAX) over the amax
arena, MAX + SUM over histograms.  Host logic only: the statistics are written by hand instead of by
kernels.  Mirrors the reference's tests/unit/torch/quantization/test_dist.py:27-47 (2 gloo ranks, amax ==
all_reduce(MAX)(amax))."""

import os
import socket
import sys

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, fn_name):
    sys.path[:0] = [ROOT, os.path.join(ROOT, "tensorrt-model-optimizer_b200"), os.path.join(ROOT, "tests")]
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        globals()[fn_name](rank, world)
    finally:
        dist.destroy_process_group()


def _spawn(fn_name, world=2):
    mp.spawn(_worker, args=(world, _free_port(), fn_name), nprocs=world, join=True)


# ---------------------------------------------------------------------------------------------------------
def _arena_sync(rank, world):
    from modelopt_b200.arena import AmaxArena

    a = AmaxArena("cpu", capacity=8)  # small on purpose: forces a growth re-allocation
    s0, s1, s2 = a.alloc(1), a.alloc(5), a.alloc(3)
    a.amax(s0)[:] = torch.tensor([1.5 + rank])
    a.amax(s1)[:] = torch.arange(5, dtype=torch.float32) * (1 if rank == 0 else -1) % 7
    a.amax(s2)[:] = torch.tensor([0.0, float("nan") if rank == 1 else 3.0, float("inf") if rank == 0 else 2.0])
    want = []
    for s in (s0, s1, s2):
        mine = a.amax(s).clone()
        gathered = [torch.empty_like(mine) for _ in range(world)]
        dist.all_gather(gathered, mine)
        want.append(torch.stack(gathered).max(0).values)  # torch.max propagates NaN like the reference
    a.sync()
    for s, w in zip((s0, s1, s2), want):
        got = a.amax(s)
        assert torch.equal(torch.isnan(got), torch.isnan(w)) and torch.equal(got.nan_to_num(), w.nan_to_num()), (got, w)
    with pytest.raises(AssertionError, match="nan"):
        a.check()


def _quantize_sync(rank, world):
    """max_calibrate's exchange with a stub backend standing in for the kernels: every rank 'collects' a different
    amax by writing its calibrator state directly; after the sync every rank holds the max."""
    import modelopt_b200 as mtq
    from modelopt_b200.arena import AmaxArenaPool
    from torch import nn

    m = nn.Sequential(nn.Linear(8, 8), nn.Linear(8, 4))
    mtq.replace_quant_module(m)
    mtq.set_quantizer_by_cfg(m, mtq.INT8_DEFAULT_CFG["quant_cfg"])
    pool = AmaxArenaPool()
    mtq.enable_stats_collection(m, pool)
    qs = [q for q in m.modules() if isinstance(q, mtq.TensorQuantizer) and q._if_calib]
    assert len(qs) == 4
    arena = pool.get("cpu")
    for i, q in enumerate(qs):  # what MaxCalibrator.collect would have left behind
        cal = q._calibrator
        n = 8 if q.axis == 0 and i == 1 else (4 if q.axis == 0 else 1)
        cal._slot, cal._slot_arena = arena.alloc(n), arena
        cal._shape, cal._dtype = ((n, 1) if q.axis == 0 else ()), torch.float32
        arena.amax(cal._slot)[:] = torch.arange(1, n + 1, dtype=torch.float32) * (rank + 1) + i
    pool.sync()
    pool.check()
    mtq.finish_stats_collection(m)
    for i, q in enumerate(qs):
        n = q.amax.numel()
        want = torch.arange(1, n + 1, dtype=torch.float32) * world + i
        assert torch.equal(q.amax.flatten(), want), (i, q.amax)
        assert q._if_quant and not q._if_calib


def _disabled_amax_sync(rank, world):
    """A quantizer that already holds an amax but is disabled when max_calibrate runs (AWQ's patched forwards leave the
    input quantizers like that) still takes part in the all-reduce(MAX), as in the reference (model_calib.py:117-130)."""
    import modelopt_b200 as mtq
    from modelopt_b200.model_calib import max_calibrate
    from torch import nn

    m = nn.Sequential(nn.Linear(8, 8))
    mtq.replace_quant_module(m)
    mtq.set_quantizer_by_cfg(m, mtq.INT8_DEFAULT_CFG["quant_cfg"])
    iq, wq = m[0].input_quantizer, m[0].weight_quantizer
    iq.amax = torch.tensor(2.0 + rank)
    iq.disable()
    wq.disable()  # nothing left for the arena; no kernel runs
    max_calibrate(m, lambda mod: None, distributed_sync=True)
    assert float(iq.amax) == 2.0 + world - 1 and iq._disabled
    iq.amax = torch.tensor(5.0 - rank)
    max_calibrate(m, lambda mod: None, distributed_sync=False)
    assert float(iq.amax) == 5.0 - rank


def _hist_sync(rank, world):
    import modelopt_b200 as mtq
    from modelopt_b200.calib import HistogramCalibrator
    from torch import nn

    m = nn.Sequential(nn.Linear(8, 8))
    mtq.replace_quant_module(m)
    mtq.set_quantizer_by_cfg(m, [{"quantizer_name": "*", "enable": False},
                                 {"quantizer_name": "*input_quantizer", "cfg": {"num_bits": 8, "calibrator": "histogram"}}])
    cal = m[0].input_quantizer._calibrator
    assert isinstance(cal, HistogramCalibrator)
    nb = 2048 if rank == 0 else 2500  # rank 1's range grew
    cal._counts = torch.arange(nb, dtype=torch.int64) % 7 + rank
    cal._calib_bin_edges = torch.arange(0, nb + 1, dtype=torch.float32) * 0.5
    cal._num_bins = nb
    mine = cal._counts.clone()
    mtq.sync_histograms(m)
    assert cal._counts.numel() == 2500 and cal._calib_bin_edges.numel() == 2501
    other_nb = 2500 if rank == 0 else 2048
    other = torch.arange(other_nb, dtype=torch.int64) % 7 + (1 - rank)
    want = torch.zeros(2500, dtype=torch.int64)
    want[:nb] += mine
    want[:other_nb] += other
    assert torch.equal(cal._counts, want)
    assert float(cal._calib_bin_edges[-1]) == 1250.0


def _hessian_sync(rank, world):
    """local-Hessian accumulators: after the sync every rank holds the Hessian of ALL ranks' tokens."""
    from modelopt_b200.model_calib import _LocalHessianAccumulator, sync_local_hessians

    g = torch.Generator().manual_seed(5)
    xs = [torch.randn(24 + 8 * r, 64, generator=g) for r in range(world)]  # ragged token counts per rank
    accs = [_LocalHessianAccumulator(32, 64, 16), _LocalHessianAccumulator(32, 64, 32), _LocalHessianAccumulator(8, 60, 16)]
    for a in accs:
        a.accumulate(xs[rank][:, :a.cin])
        a.normalized_hessian()  # cached before the sync on purpose: the sync must drop it
    sync_local_hessians(accs)
    allx = torch.cat(xs)
    for a in accs[:2]:
        xt = allx.T.reshape(a.num_blocks_per_cin, a.block_size, -1)
        want = xt @ xt.transpose(-1, -2) / allx.shape[0]
        assert a.num_samples == allx.shape[0]
        torch.testing.assert_close(a.normalized_hessian(), want, rtol=1e-5, atol=1e-5)
    assert not accs[2].is_enabled and accs[2].hessian_per_block is None
    got = accs[0].normalized_hessian().clone()
    gathered = [torch.empty_like(got) for _ in range(world)]
    dist.all_gather(gathered, got)
    assert all(torch.equal(t, gathered[0]) for t in gathered)  # bit-identical on every rank


@pytest.mark.parametrize("fn", ["_arena_sync", "_quantize_sync", "_hist_sync", "_hessian_sync", "_disabled_amax_sync"])
def test_world_size_2(fn):
    _spawn(fn)
