"""smw_sumcheck_round_coeffs_noalloc (include/smw.h): the allocation-free variant of smw_sumcheck_round_coeffs — fixed 64 KiB device
scratch inside the library, device pointers in and out, no cudaMalloc / cudaFree.  No counterpart in the reference;
checked bit for bit (integer work: tolerance 0, as everywhere in this suite) against the existing
entry point, through the C-ABI, on two calls in a row that share the scratch.  Collected last on purpose (pytest -x)."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _both(ctx, smw, layer):
    import torch
    want = smw.sumcheck_round_coeffs(ctx, layer)                                  # existing entry point (allocates its partials)
    dev = torch.from_numpy(layer.view(np.int64).copy()).cuda()
    out = torch.full((8,), -1, dtype=torch.int64, device="cuda")
    torch.cuda.synchronize()                                                      # the context runs on its own stream: inputs must have landed
    smw.sumcheck_round_coeffs_noalloc(ctx, dev, out)
    ctx.sync()
    got = out.cpu().numpy().view(np.uint64).reshape(2, 4)
    return want, got


def test_two_calls_in_a_row_match_the_allocating_entry_point(ctx, smw, O, rnd):
    # first call: 1024 blocks of partials (the whole scratch); second: a short layer that uses 3 blocks — stale partials must not leak in
    for seed, n in ((31, 1 << 20), (32, 1400), (33, 2)):
        layer = rnd(seed, n)
        want, got = _both(ctx, smw, layer)
        assert np.array_equal(got[0], want[0]) and np.array_equal(got[1], want[1]), "n = %d" % n
    c0, c1 = O.sumcheck_round_coeffs(rnd(32, 1400))                               # and the oracle, once (channel/src/lib.rs:406-416)
    want, got = _both(ctx, smw, rnd(32, 1400))
    assert np.array_equal(got[0], c0) and np.array_equal(got[1], c1)


def test_rejects_sizes_beyond_the_fixed_scratch_and_bad_arguments(ctx, smw):
    import torch
    buf = torch.zeros((4, 4), dtype=torch.int64, device="cuda")
    out = torch.zeros((8,), dtype=torch.int64, device="cuda")
    L = ctx.L
    call = lambda p, n, o: L.smw_sumcheck_round_coeffs_noalloc(ctx.h, C.c_void_p(p), C.c_size_t(n), C.c_void_p(o))
    assert call(buf.data_ptr(), (1 << 26) + 2, out.data_ptr()) == 8               # SMW_ERR_INVALID_ARG before anything is launched or read
    assert b"2^26" in L.smw_last_error(ctx.h)
    assert call(buf.data_ptr(), 3, out.data_ptr()) == 8                           # odd length
    assert call(buf.data_ptr(), 0, out.data_ptr()) == 8
    assert call(None, 4, out.data_ptr()) == 8
    assert call(buf.data_ptr(), 4, None) == 8
    assert call(buf.data_ptr(), 4, out.data_ptr()) == 0                           # and still works afterwards
    ctx.sync()
