"""GPU: `mo_hist_entropy_static` (the entry point that allocates nothing: its scratch is a fixed file-scope device array
inside the library) against `mo_hist_entropy` with a caller-provided scratch buffer, through the C ABI, on two calls in a
row that share the array.  The entry point has no counterpart in the reference.
This file is collected last on purpose (`pytest -x` stops at the first failure)."""

import ctypes

import pytest
import torch

from modelopt_b200 import _lib

pytestmark = pytest.mark.gpu


def _counts(nbins, seed):
    """A histogram shaped like |activations|: a half-normal body, a long thin tail and some empty bins."""
    g = torch.Generator().manual_seed(seed)
    x = torch.randn(400_000, generator=g).abs() * (nbins / 6.0)
    x[:64] = torch.rand(64, generator=g) * nbins  # outliers out to the last bin
    c = torch.bincount(x.clamp(max=nbins - 1).long(), minlength=nbins)
    return c.to(torch.int64).cuda()


def _existing(lib, counts, bits, unsigned, stride, start_bin):
    n = counts.numel()
    nbytes = lib.mo_hist_entropy_scratch_bytes(n, stride, start_bin)
    scratch = torch.empty(nbytes, dtype=torch.uint8, device="cuda")
    idx = torch.full((1,), -7, dtype=torch.int32, device="cuda")
    rc = lib.mo_hist_entropy(counts.data_ptr(), n, bits, unsigned, stride, start_bin, scratch.data_ptr(), idx.data_ptr(),
                             torch.cuda.current_stream().cuda_stream)
    assert rc == 0, lib.mo_last_error()
    return int(idx.item())


def test_static_scratch_entry_point_matches_and_allocates_nothing():
    lib = _lib.load()
    assert lib.mo_hist_entropy_static_max_bins() == 65536
    stream = torch.cuda.current_stream().cuda_stream
    # call 1: a large histogram fills most of the array; call 2: a smaller one with other parameters runs over the
    # first call's leftovers (the array keeps them); call 3 repeats call 1
    cases = [(12288, 8, 0, 1, 128, 1), (2048, 4, 1, 2, 64, 2), (12288, 8, 0, 1, 128, 1)]
    inputs = [(_counts(nb, seed), nb, bits, uns, stride, start) for nb, bits, uns, stride, start, seed in cases]
    want = [_existing(lib, c, bits, uns, stride, start) for c, _, bits, uns, stride, start in inputs]
    outs = [torch.full((1,), -7, dtype=torch.int32, device="cuda") for _ in inputs]
    torch.cuda.synchronize()
    # warm-up call: the first use looks up the array's device address (once per device); it allocates nothing either
    rc = lib.mo_hist_entropy_static(inputs[1][0].data_ptr(), 2048, 4, 1, 2, 64, outs[1].data_ptr(), stream)
    assert rc == 0, lib.mo_last_error()
    torch.cuda.synchronize()
    free0, _ = torch.cuda.mem_get_info()
    torch_alloc0 = torch.cuda.memory_allocated()
    for (c, nb, bits, uns, stride, start), out in zip(inputs, outs):
        rc = lib.mo_hist_entropy_static(c.data_ptr(), nb, bits, uns, stride, start, out.data_ptr(), stream)
        assert rc == 0, lib.mo_last_error()
    torch.cuda.synchronize()
    assert torch.cuda.mem_get_info()[0] == free0, "the static entry point changed the device's free memory"
    assert torch.cuda.memory_allocated() == torch_alloc0
    got = [int(o.item()) for o in outs]
    assert got == want  # integer indices: exact
    assert want[0] == want[2] and 0 <= want[0]


def test_static_scratch_entry_point_rejects_sizes_beyond_the_array():
    lib = _lib.load()
    n = 65536 + 1
    counts = torch.ones(n, dtype=torch.int64, device="cuda")
    idx = torch.zeros(1, dtype=torch.int32, device="cuda")
    rc = lib.mo_hist_entropy_static(counts.data_ptr(), n, 8, 0, 1, 128, idx.data_ptr(), None)
    assert rc == -3  # MO_ERR_ARG
    msg = lib.mo_last_error().decode()
    assert "65536" in msg and "mo_hist_entropy" in msg
    # at the limit it runs, and agrees with the existing entry point
    counts = _counts(65536, 3)
    want = _existing(lib, counts, 8, 0, 64, 128)
    rc = lib.mo_hist_entropy_static(counts.data_ptr(), 65536, 8, 0, 64, 128, idx.data_ptr(),
                                    torch.cuda.current_stream().cuda_stream)
    assert rc == 0, lib.mo_last_error()
    assert int(idx.item()) == want


def test_ops_wrapper_uses_the_callers_output():
    from modelopt_b200 import ops

    counts = _counts(4096, 5)
    out = torch.full((1,), -1, dtype=torch.int32, device="cuda")
    r = ops.hist_entropy_index_static(counts, 8, False, out)
    assert r is out
    assert int(out.item()) == int(ops.hist_entropy_index(counts, 8, False).item())
