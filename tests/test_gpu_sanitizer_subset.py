"""The subset run under `compute-sanitizer --tool memcheck` and `--tool racecheck` (tools/gpu_run.sh sanitizer; logs in profiles/):
small instances that drive every kernel family once — the shared-memory interpreter (run_ops52 / run_ops), the cooperative
permutation with its broadcast buffer and redux sums (coop_perm), the NTT passes, folds, the multi-run batch kernels.
Each result is still checked against the oracle, so the suite is also a quick parity smoke."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def test_sanitizer_subset(ctx, smw, O, rnd):
    # thread-per-node batch kernel (t = 17, > 4096 nodes) followed by cooperative tops
    leaves = rnd(1, 1 << 17)
    assert np.array_equal(smw.MerkleCommitment(smw.MerkleConfig.with_default_params(ctx, 5)).commit(leaves)[0], O.commitment_commit(leaves, 5).root())
    # t = 9: pair leaves + arity-2 levels, multi-run batch kernels and cooperative small rounds in one FRI build; hashed leaves (t = 17)
    for log_n0, sched in ((13, [2] * 13), (12, [16, 16, 8]), (10, [32, 4])):
        f0 = rnd(2 + log_n0, 1 << log_n0)
        assert np.array_equal(smw.fri_build_transcript(ctx, f0, smw.FriProverParams(sched, 0xDEEFBAAD)).roots(), O.FriProverState(f0, sched, 0xDEEFBAAD).roots())
    # slices + packed exchange buffer + tops
    f0 = rnd(3, 1 << 10)
    got = smw.fri_roots_sharded(ctx, None, f0.reshape(4, 256, 4), [0, 256, 512, 768], 1 << 10, smw.FriProverParams([2] * 10, 7))
    assert np.array_equal(got, O.FriProverState(f0, [2] * 10, 7).roots())
    # cooperative and batched permutations, both widths
    for t, batch in ((17, 5), (9, 7), (17, 9000), (9, 2500)):
        st = rnd(10 + t, batch * t).reshape(batch, t, 4)
        assert np.array_equal(smw.permute(ctx, st, smw.poseidon_params_for_width(ctx, t)), O.permute(O.Params.for_width(t), st))
    # transcript: long serial sponge on one warp, challenges
    xs = rnd(4, 200)
    assert np.array_equal(smw.tr_hash_fields_tagged(ctx, b"ALI/A", xs), O.tr_hash_fields_tagged(b"ALI/A", xs))
    # NTT / LDE passes, Merkle-folded sum-check loop
    col = rnd(5, 1 << 12)
    assert np.array_equal(smw.fft(ctx, col), O.ntt(col)) and np.array_equal(smw.ifft(ctx, col), O.ntt(col, inverse=True))
    g = O.to_mont([5])[0]
    assert np.array_equal(smw.lde(ctx, col.reshape(1, -1, 4), 2, g)[0], O.lde(col, 2, g))
    table = rnd(6, 1 << 9)
    assert all(np.array_equal(a, b) for a, b in zip(smw.mf_commit_phase(ctx, table, 9), O.mf_commit_phase(table, 9)))
