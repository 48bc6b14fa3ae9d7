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
    """sync_histograms: ranks that share the bin width (what the all-reduce(MAX) of the first batch's maximum inside collect
    guarantees) merge - the widest rank's bin count / range, exact counts summed in one flat buffer, a narrower rank's
    missing bins counting as zeros; ranks whose widths differ are refused.  The device state is written by hand here (the
    kernels need a GPU)."""
    import modelopt_b200 as mtq
    from modelopt_b200.calib import HistogramCalibrator
    from torch import nn

    m = nn.Sequential(nn.Linear(8, 8), nn.Linear(8, 8))
    mtq.replace_quant_module(m)
    mtq.set_quantizer_by_cfg(m, [{"quantizer_name": "*", "enable": False},
                                 {"quantizer_name": "*input_quantizer", "cfg": {"num_bits": 8, "calibrator": "histogram"}}])
    mtq.enable_stats_collection(m)
    cals = [m[0].input_quantizer._calibrator, m[1].input_quantizer._calibrator]
    assert all(isinstance(c, HistogramCalibrator) and c.dp_group is None for c in cals)  # the default (world) group
    assert all(c.dp_sync == "finish" for c in cals)

    def f32_bits(v):
        return int(torch.tensor([v], dtype=torch.float32).view(torch.int32)[0])

    def fill(cal, nb, k, r):
        cal._state = torch.zeros(16, dtype=torch.int32)
        cal._state[0], cal._state[1], cal._state[2] = nb, f32_bits(nb * 0.5), f32_bits(0.5)
        cal._state[3] = 0 if nb == 2048 else f32_bits(nb * 0.5 + 0.25)
        cal._state[5] = f32_bits(1024.0)
        cal._buf = torch.zeros(4096, dtype=torch.int64)
        cal._buf[:nb] = torch.arange(nb, dtype=torch.int64) % (7 + k) + r
        cal._synced = cal._snap_valid = False
        cal._edges_cache = None

    # (1) every rank ended with the same range
    for k, (cal, nb) in enumerate(zip(cals, (2048, 2500))):
        fill(cal, nb, k, rank)
    mtq.sync_histograms(m)
    for k, (cal, nb) in enumerate(zip(cals, (2048, 2500))):
        want = sum(torch.arange(nb, dtype=torch.int64) % (7 + k) + r for r in range(world))
        assert cal._synced and torch.equal(cal._counts, want)
        assert cal._calib_bin_edges.numel() == nb + 1
    before = [c._counts.clone() for c in cals]
    mtq.sync_histograms(m)  # already summed: a second call must not sum again
    assert all(torch.equal(a, c._counts) for a, c in zip(before, cals))
    # (2) the second calibrator's range grew further on the higher rank: the merge takes the widest range, the narrower
    # rank contributes zeros above its own last bin
    nbs = [2500 + 100 * r for r in range(world)]
    fill(cals[0], 2048, 0, rank)
    fill(cals[1], nbs[rank], 1, rank)
    mtq.sync_histograms(m)
    want = torch.zeros(nbs[-1], dtype=torch.int64)
    for r in range(world):
        want[:nbs[r]] += torch.arange(nbs[r], dtype=torch.int64) % 8 + r
    assert cals[1]._num_bins == nbs[-1] and torch.equal(cals[1]._counts, want)
    assert cals[1]._calib_bin_edges.numel() == nbs[-1] + 1
    assert float(cals[1]._calib_bin_edges[-1]) == nbs[-1] * 0.5
    # (3) ranks with different bin widths (no exchange at the first batch): refused, not silently mis-added
    fill(cals[0], 2048, 0, rank)
    fill(cals[1], 2500, 1, rank)
    cals[1]._state[2] = f32_bits(0.5 + 0.25 * rank)
    with pytest.raises(RuntimeError, match="disagree"):
        mtq.sync_histograms(m)
    # (4) a calibrator that froze on one rank (range beyond the counts buffer) is reported on every rank
    fill(cals[0], 2048, 0, rank)
    fill(cals[1], 2500, 1, rank)
    if rank == world - 1:
        cals[1]._state[4] = 1
    with pytest.raises(RuntimeError, match="headroom"):
        mtq.sync_histograms(m)


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


def _calibrating_model():
    import modelopt_b200 as mtq
    from modelopt_b200.arena import AmaxArenaPool
    from torch import nn

    m = nn.Sequential(nn.Linear(8, 8), nn.Linear(8, 4))
    mtq.replace_quant_module(m)
    mtq.set_quantizer_by_cfg(m, mtq.FP8_DEFAULT_CFG["quant_cfg"])  # per-tensor everywhere: 4 scalar amax
    pool = AmaxArenaPool()
    mtq.enable_stats_collection(m, pool)
    qs = [q for q in m.modules() if isinstance(q, mtq.TensorQuantizer) and q._if_calib]
    assert len(qs) == 4
    return m, pool, qs


def _put(cal, arena, value):
    """What MaxCalibrator.collect leaves behind: a scalar running amax in `arena` (or private when arena is None)."""
    if arena is None:
        cal._bits = torch.tensor([value], dtype=torch.float32).view(torch.int32)
    else:
        cal._slot, cal._slot_arena = arena.alloc(1), arena
        arena.amax(cal._slot)[:] = value
    cal._shape, cal._dtype = (), torch.float32


def _stray_calibrators_sync(rank, world):
    """Calibrators whose state lives OUTSIDE this call's pool (they collected before it existed: AWQ's stats passes, a second
    calibrate() without reset_amax) still take part in the exchange, and only pool-backed ones count as pool-validated."""
    from modelopt_b200.arena import AmaxArena
    from modelopt_b200.model_calib import max_calibrate, sync_max_calibrators

    m, pool, qs = _calibrating_model()
    old = AmaxArena("cpu")
    _put(qs[0]._calibrator, pool.get("cpu"), 1.0 + rank)
    _put(qs[1]._calibrator, None, 2.0 + 2 * rank)       # private bits
    _put(qs[2]._calibrator, old, 3.0 - rank)            # an older call's arena
    _put(qs[3]._calibrator, pool.get("cpu"), 4.0 + rank)
    validated = sync_max_calibrators(m, pool, None)
    assert validated == {id(qs[0]._calibrator), id(qs[3]._calibrator)}
    got = [float(q._calibrator._calib_amax) for q in qs]
    assert got == [1.0 + world - 1, 2.0 + 2 * (world - 1), 3.0, 4.0 + world - 1], got
    # the whole flow: a second max_calibrate (fresh pool) on calibrators that keep their old state
    for i, q in enumerate(qs):
        b = q._calibrator._slot_arena.bits(q._calibrator._slot) if q._calibrator._slot is not None else q._calibrator._bits
        b.view(torch.float32)[:] = 10.0 * (i + 1) + rank
    for mod in m:
        mod.__dict__["_weight_collect_key"] = None
    import modelopt_b200.model_calib as mc

    orig = mc.weight_only_quantize
    mc.weight_only_quantize = lambda model: None  # no kernels on the CPU box
    try:
        max_calibrate(m, lambda mod: None, distributed_sync=True)
    finally:
        mc.weight_only_quantize = orig
    assert [float(q.amax) for q in qs] == [10.0 * (i + 1) + world - 1 for i in range(4)]


def _uneven_layout_sync(rank, world):
    """One rank never fed a quantizer (empty MoE expert, uneven forward_loop): the arena layouts differ, the ranks notice
    and exchange quantizer by quantizer; the rank without data adopts the group's amax instead of hanging in the collective."""
    from modelopt_b200.model_calib import sync_max_calibrators

    m, pool, qs = _calibrating_model()
    arena = pool.get("cpu")
    for i, q in enumerate(qs):
        if rank == 1 and i == 1:
            continue  # no data on this rank
        _put(q._calibrator, arena, (i + 1) * (1.0 + rank))
    sync_max_calibrators(m, pool, None)
    got = [float(q._calibrator._calib_amax) for q in qs]
    assert got == [1.0 * world, 2.0, 3.0 * world, 4.0 * world], got


@pytest.mark.parametrize("fn", ["_arena_sync", "_quantize_sync", "_hist_sync", "_hessian_sync", "_disabled_amax_sync",
                                "_stray_calibrators_sync", "_uneven_layout_sync"])
def test_world_size_2(fn):
    _spawn(fn)
