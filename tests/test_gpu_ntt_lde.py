"""GPU parity: NTT / inverse NTT / coset LDE (fft/src/lib.rs:6-32 + SURVEY.md A.7) against the CPU oracle, bit-exact.

The reference holds no NTT vector ("parity unpinned" for this kernel): the oracle's NTT is pinned on the
definition evals[i] = P(w^i) (tests/test_oracle_anchors.py::test_ntt_against_naive_dft) and the GPU is compared with it
here, plus the reference's own all-ones round trip (fft/src/lib.rs:40-54) and size-independent properties at 2^20."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def eq(a, b):
    return np.array_equal(np.asarray(a, dtype=np.uint64), np.asarray(b, dtype=np.uint64))


@pytest.mark.parametrize("log_n", [0, 1, 2, 5, 9, 10, 11, 13, 16])
def test_fft_ifft_match_oracle(ctx, smw, O, rnd, log_n):
    c = rnd(300 + log_n, 1 << log_n)
    ev = smw.fft(ctx, c)
    assert eq(ev, O.ntt(c) if log_n else c)
    assert eq(smw.ifft(ctx, ev), c)
    assert eq(smw.ifft(ctx, c), O.ntt(c, inverse=True) if log_n else c)


def test_fft_small_is_the_dft_definition(ctx, smw, O, rnd):
    c = rnd(77, 64)
    assert eq(smw.fft(ctx, c), O.dft_naive(c))


def test_reference_all_ones_round_trip(ctx, smw, O):
    """fft/src/lib.rs:40-54: n = 8, all ones, fft then ifft gives the input back"""
    ones = O.to_mont([1] * 8)
    ev = smw.fft(ctx, ones)
    assert eq(smw.ifft(ctx, ev), ones)
    assert O.from_mont(ev) == [8, 0, 0, 0, 0, 0, 0, 0]


def test_fft_many_columns(ctx, smw, O, rnd):
    cols = rnd(41, 5 << 12).reshape(5, 1 << 12, 4)
    got = smw.fft(ctx, cols)
    for c in range(5):
        assert eq(got[c], O.ntt(cols[c]))


@pytest.mark.parametrize("log_n,log_b,ncols", [(0, 2, 1), (1, 1, 2), (4, 0, 1), (6, 3, 3), (10, 2, 2), (13, 2, 2), (16, 1, 1)])
def test_lde_matches_oracle(ctx, smw, O, rnd, log_n, log_b, ncols):
    cols = rnd(500 + log_n * 7 + log_b, ncols << log_n).reshape(ncols, 1 << log_n, 4)
    shift = O.to_mont([5])[0]
    got = smw.lde(ctx, cols, log_b, shift)
    assert got.shape == (ncols, 1 << (log_n + log_b), 4)
    for c in range(ncols):
        assert eq(got[c], O.lde(cols[c], log_b, shift))


def test_lde_other_shift_and_unit_shift(ctx, smw, O, rnd):
    col = rnd(9, 1 << 8)
    for sh in (1, 7, 0xDEADBEEF):
        shift = O.to_mont([sh])[0]
        assert eq(smw.lde(ctx, col, 2, shift), O.lde(col, 2, shift))
    # shift 1: every 4th point of the extension is the original evaluation
    ext = smw.lde(ctx, col, 2, O.to_mont([1])[0])
    assert eq(ext[::4], col)


def test_lde_commit_matches_oracle_per_column(ctx, smw, O, rnd):
    cols = rnd(2024, 3 << 9).reshape(3, 1 << 9, 4)
    shift = O.to_mont([5])[0]
    roots = smw.lde_commit(ctx, cols, 2, shift, label0=40)
    for c in range(3):
        assert eq(roots[c], O.commitment_commit(O.lde(cols[c], 2, shift), 40 + c).root())


def test_cfg1_full_size_matches_oracle(ctx, smw, O):
    """BASELINE configs[0] / SURVEY 8d cfg1 in full: 8 columns x 2^16 rows of the bench generator (seed 0x5EED0001), blow-up 4, coset
    shift 5, one arity-16 commitment-default tree per extended column with tree_label = column index — every root equals the
    oracle's, computed here on the host cores (about ten seconds)."""
    from bench import gen_columns
    cols = gen_columns(0x5EED0001, 0, 8, 1 << 16)
    shift = O.to_mont([5])[0]
    roots = smw.lde_commit(ctx, cols, 2, shift, label0=0)
    for c in range(8):
        assert eq(roots[c], O.commitment_commit(O.lde(cols[c], 2, shift), c).root()), "cfg1 column %d" % c


def test_full_size_properties_2p20(ctx, smw, O, rnd):
    """BASELINE cfg2 column size (2^20 rows, blow-up 4), checked through size-independent properties:
    ifft(fft(x)) = x; linearity fft(a + b) = fft(a) + fft(b); LDE with shift 1 interpolates the column; and the oracle on
    the same column (one 2^22-point CPU NTT, a few seconds)."""
    n = 1 << 20
    a, b = rnd(1, n), rnd(2, n)
    fa = smw.fft(ctx, a)
    assert eq(smw.ifft(ctx, fa), a)
    fb = smw.fft(ctx, b)
    assert eq(smw.fft(ctx, O.fr_add(a, b)), O.fr_add(fa, fb))
    one = O.to_mont([1])[0]
    assert eq(smw.lde(ctx, a, 2, one)[::4], a)
    shift = O.to_mont([5])[0]
    assert eq(smw.lde(ctx, a, 2, shift), O.lde(a, 2, shift))


def test_bad_sizes(ctx, smw, O, rnd):
    with pytest.raises(smw.SmwError):
        smw.lde(ctx, rnd(3, 4), 31, O.to_mont([5])[0])     # 2^2 << 31 exceeds the two-adicity 32


def test_lde_commit_streams_in_chunks(ctx, smw, O, rnd, monkeypatch):
    """smw_lde_commit streams the columns in chunks (double-buffered H2D for host input): forcing 1-, 2- and 3-column chunks
    gives the same roots as one chunk, for host and for device-resident input"""
    import torch
    cols = rnd(77, 5 << 8).reshape(5, 1 << 8, 4)
    shift = O.to_mont([5])[0]
    want = np.stack([O.commitment_commit(O.lde(cols[c], 2, shift), 100 + c).root() for c in range(5)])
    dev = torch.from_numpy(cols.view(np.int64)).cuda()
    for ncols_per_chunk in (1, 2, 3, 64):
        monkeypatch.setenv("SMW_LDE_CHUNK_BYTES", str(ncols_per_chunk * (1 << 10) * 32))
        assert eq(smw.lde_commit(ctx, cols, 2, shift, label0=100), want)
        assert eq(smw.lde_commit(ctx, dev, 2, shift, label0=100), want)
