"""GPU parity of the cross-GPU four-step NTT building blocks (smw_scale_powers, smw_dft_strided, smw_ntt) and of
shard.ntt_cyclic_to_chunks + shard.fri_roots_sharded_chunks: one GPU plays every rank, the all-to-all
and all-gather are numpy shuffles, results must equal the CPU oracle's NTT (fft/src/lib.rs:6-12 convention) and FRI roots."""
import numpy as np
import pytest

from stark_mlwe_b200 import shard

pytestmark = pytest.mark.gpu
SEED_Z = 0xDEEFBAAD


def test_scale_powers_and_strided_dft_match_oracle(ctx, smw, O, rnd):
    import sys, os
    sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
    from oracle_shard_backend import OracleNttBackend
    ob = OracleNttBackend()
    x = rnd(7, 3000)
    g = rnd(8, 1)[0]
    assert np.array_equal(smw.scale_powers(ctx, x.copy(), g), ob.scale_powers(x, g))
    for log_g, stride, count in ((1, 40, 40), (2, 64, 50), (3, 128, 128), (4, 16, 16)):
        a = rnd(20 + log_g, (1 << log_g) * stride)
        assert np.array_equal(smw.dft_strided(ctx, a.copy(), log_g, stride, count), ob.dft_strided(a, log_g, stride, count))


def run_distributed(ctx, x, log_n, G):
    """every rank's chunks, with the all-to-all done by hand"""
    be = shard.GpuNttBackend(ctx)
    sent = {}

    class Stop(Exception):
        pass
    for r in range(G):
        def grab(p, r=r):
            sent[r] = np.array(p)
            raise Stop()
        try:
            shard.ntt_cyclic_to_chunks(be, x[r::G].copy(), log_n, r, G, all_to_all=grab)
        except Stop:
            pass
    return [shard.ntt_cyclic_to_chunks(be, x[r::G].copy(), log_n, r, G, all_to_all=lambda p, r=r: np.stack([sent[a][r] for a in range(G)]))
            for r in range(G)]


@pytest.mark.parametrize("log_n,G", [(10, 2), (12, 4), (12, 8), (9, 1)])
def test_distributed_ntt_matches_oracle(ctx, O, rnd, log_n, G):
    N = 1 << log_n
    M = N // G
    x = rnd(100 + log_n + G, N)
    X = O.ntt(x)
    if G == 1:
        got = shard.ntt_cyclic_to_chunks(shard.GpuNttBackend(ctx), x.copy(), log_n, 0, 1)
        assert np.array_equal(got.reshape(-1, 4), X)
        return
    out = run_distributed(ctx, x, log_n, G)
    for r in range(G):
        for k1 in range(G):
            lo = M * k1 + r * (M // G)
            assert np.array_equal(out[r][k1], X[lo:lo + M // G]), "rank %d chunk %d" % (r, k1)


@pytest.mark.parametrize("log_n,G,schedule", [(10, 2, [2] * 10), (12, 4, [16, 16, 8])])
def test_distributed_ntt_then_chunked_fri(ctx, O, rnd, log_n, G, schedule):
    N = 1 << log_n
    M = N // G
    x = rnd(200 + log_n, N)
    out = run_distributed(ctx, x, log_n, G)
    want = O.FriProverState(O.ntt(x), schedule, SEED_Z).roots()
    be = shard.GpuFriBackend(ctx)
    firsts = lambda r: [M * k1 + r * (M // G) for k1 in range(G)]      # noqa: E731
    # per-rank shards once, then a hand-made all-gather in rank-major order
    shs = {r: [be.shard(out[r][c], N, firsts(r)[c], schedule, SEED_Z) for c in range(G)] for r in range(G)}
    K = shs[0][0].num_local_layers

    class Rank(shard.GpuFriBackend):
        def __init__(self, r):
            super().__init__(ctx)
            self.r, self.i = r, 0

        def shard(self, *a):
            s = shs[self.r][self.i]
            self.i += 1
            return s
    for r in range(G):
        calls = [0]

        def gather(_mine):
            i = calls[0]
            calls[0] += 1
            if i == 0:
                return np.array([[f, 0, 0, 0] for q in range(G) for f in firsts(q)], dtype=np.uint64)
            l = i - 1
            if l < K:
                return np.concatenate([shs[q][c].cap(l)[0] for q in range(G) for c in range(G)])
            return np.concatenate([shs[q][c].tail() for q in range(G) for c in range(G)])
        got = shard.fri_roots_sharded_chunks(Rank(r), [out[r][c] for c in range(G)], firsts(r), N, G, schedule, SEED_Z, gather=gather)
        assert np.array_equal(got, want), "rank %d" % r
