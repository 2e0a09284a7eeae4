"""GPU parity of the legacy trees: MerkleTree::new_legacy, new_pairs_legacy, verify_many, verify_pairs_legacy
(merkle/src/lib.rs:196-223, 448-470, 487-584, 704-721) — fixed t = 17, chunks of poseidon::RATE = 16, every node
hash_with_ds(chunk, ds_tag).  Every level and every true/false decision against the CPU oracle."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("n", [1, 2, 16, 17, 300, 4097])
def test_new_legacy_matches_oracle(ctx, smw, O, rnd, n):
    leaves, tag = rnd(600 + n, n), rnd(7, 1)[0]
    t, o = smw.MerkleTree.new_legacy(ctx, leaves, tag), O.MerkleTree.new_legacy(leaves, tag)
    assert t.arity() == 16 and t.height() + 1 == o.nlevels()
    for l in range(o.nlevels()):
        assert np.array_equal(t.level(l), o.level(l)), l
    assert np.array_equal(t.root(), o.root())
    # another parameter set (the bench seed, poseidon/benches/poseidon.rs:7) and the hash it is built from
    p, op = smw.generate_params_t17_x5(ctx, b"POSEIDON-T17-X5"), O.Params.from_seed(b"POSEIDON-T17-X5", 17, 64)
    assert np.array_equal(smw.MerkleTree.new_legacy(ctx, leaves, tag, p).root(), O.MerkleTree.new_legacy(leaves, tag, op).root())
    if n <= 16:
        assert np.array_equal(t.root() if n > 1 else smw.hash_with_ds(ctx, leaves, tag, smw.commitment_default_params(ctx)), O.hash_with_ds_legacy(O.Params.commitment_default(), leaves, tag))


def test_legacy_open_verify_round_trip_and_negatives(ctx, smw, O, rnd):
    n = 1000
    leaves, tag = rnd(61, n), rnd(8, 1)[0]
    t, o = smw.MerkleTree.new_legacy(ctx, leaves, tag), O.MerkleTree.new_legacy(leaves, tag)
    idx = [0, 15, 16, 500, 999, 500]
    pr, opr = t.open_many(idx), o.open_many(idx)
    opr.arity = 16
    assert pr.indices == opr.indices and all(np.array_equal(a, b) for a, b in zip(pr.siblings, opr.siblings))
    vals = leaves[idx]
    assert smw.verify_many(ctx, t.root(), idx, vals, pr, tag) and O.verify_many_legacy(o.root(), idx, vals, pr, tag)
    bad = vals.copy(); bad[1, 3] ^= np.uint64(1)
    assert not smw.verify_many(ctx, t.root(), idx, bad, pr, tag) and not O.verify_many_legacy(o.root(), idx, bad, opr, tag)
    assert not smw.verify_many(ctx, t.root(), idx, vals, pr, rnd(9, 1)[0])                  # another ds_tag
    assert not smw.verify_many(ctx, t.root(), idx[:-2], vals[:-2], pr, tag)                   # proof.indices != requested set
    assert not smw.verify_many(ctx, t.root(), [], vals[:0], pr, tag)
    # a DS-aware tree over the same leaves has a different root: the two sponges are domain-separated
    assert not np.array_equal(t.root(), smw.MerkleCommitment(smw.MerkleConfig.with_default_params(ctx, 0)).commit(leaves)[0])


@pytest.mark.parametrize("n", [1, 33, 700])
def test_new_pairs_legacy_and_verify_pairs_legacy(ctx, smw, O, rnd, n):
    f, cp, tag = rnd(71, n), rnd(72, n), rnd(73, 1)[0]
    t, o = smw.MerkleTree.new_pairs_legacy(ctx, f, cp, tag), O.MerkleTree.new_pairs_legacy(f, cp, tag)
    for l in range(o.nlevels()):
        assert np.array_equal(t.level(l), o.level(l)), l
    idx = sorted({0, n - 1, n // 2})
    pr = t.open_many(idx)
    pairs = np.stack([np.stack([f[i], cp[i]]) for i in idx])
    assert smw.verify_pairs_legacy(ctx, t.root(), idx, pairs, pr, tag) == O.verify_pairs_legacy(o.root(), idx, pairs, pr, tag)
    assert smw.verify_pairs_legacy(ctx, t.root(), idx, pairs, pr, tag) or n == 1          # a one-leaf tree has no levels: the root is the leaf digest
    bad = pairs.copy(); bad[0, 1, 0] ^= np.uint64(1)
    assert not smw.verify_pairs_legacy(ctx, t.root(), idx, bad, pr, tag)
    with pytest.raises(smw.SmwError) as e:
        smw.MerkleTree.new_pairs_legacy(ctx, f, cp[:-1] if n > 1 else rnd(1, 2), tag)
    assert e.value.code == 4
