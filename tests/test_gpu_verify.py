"""GPU parity: the multiproof verifiers (smw_verify_many_ds / smw_verify_pairs_ds) against the oracle's verify_many_ds and the
reference's own verifier tests (merkle/src/lib.rs:934-963 round trip, :966-1010 multi-level, :1053-1084 pairs, :1108-1137
tamper negatives).  Every decision (true / false) must equal the oracle's on the same proof, honest or malformed."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

LABEL = 0xC0FFEE


def _tree(ctx, smw, O, rnd, arity, n, seed):
    leaves = rnd(seed, n)
    cfg = smw.MerkleChannelCfg.new(ctx, arity).with_tree_label(LABEL)
    return leaves, smw.MerkleTree.new(leaves, cfg), cfg.params, O.Params.for_arity(arity)


def _both(smw, O, root, idx, vals, pr, label, gp, op, arity):
    g = smw.verify_many_ds(root, idx, vals, pr, label, gp)
    o = O.verify_many_ds(root, idx, vals, pr, label, op, arity)
    assert g == o, "device verifier and oracle disagree"
    return g


@pytest.mark.parametrize("arity,n", [(16, 1 << 12), (16, 17), (16, 1), (8, 777), (2, 64), (4, 100), (32, 1000), (64, 4097), (128, 300)])
def test_verify_many_ds_round_trip_and_tamper(ctx, smw, O, rnd, arity, n):
    leaves, tree, gp, op = _tree(ctx, smw, O, rnd, arity, n, 7000 + arity + n)
    root = tree.root()
    idx = sorted({0, n - 1, n // 2, n // 3, min(n // 3 + 1, n - 1), 5 % n})
    pr = tree.open_many(idx)
    vals = leaves[idx]
    assert _both(smw, O, root, idx, vals, pr, LABEL, gp, op, arity)
    # unsorted request with a duplicate: the verifier sorts and de-duplicates (merkle/src/lib.rs:598-600)
    perm = idx[::-1] + [idx[0]]
    assert _both(smw, O, root, perm, leaves[perm], pr, LABEL, gp, op, arity)
    if n == 1:
        return
    bad = vals.copy(); bad[-1, 1] ^= np.uint64(1)
    assert not _both(smw, O, root, idx, bad, pr, LABEL, gp, op, arity)                 # wrong opened value
    assert not _both(smw, O, root, idx, vals, pr, LABEL + 1, gp, op, arity)            # wrong tree label (DS)
    broot = root.copy(); broot[0] ^= np.uint64(2)
    assert not _both(smw, O, broot, idx, vals, pr, LABEL, gp, op, arity)               # wrong root
    if pr.nlev and pr.siblings[0].shape[0]:
        sib = [x.copy() for x in pr.siblings]; sib[0][0, 0] ^= np.uint64(1)
        assert not _both(smw, O, root, idx, vals, smw.MerkleProof.from_fields(pr.indices, sib, pr.group_sizes, arity), LABEL, gp, op, arity)
        short = [x.copy() for x in pr.siblings]; short[0] = short[0][:-1]             # one sibling missing
        assert not _both(smw, O, root, idx, vals, smw.MerkleProof.from_fields(pr.indices, short, pr.group_sizes, arity), LABEL, gp, op, arity)
        extra = [x.copy() for x in pr.siblings]; extra[-1] = np.concatenate([extra[-1], leaves[:1]])   # one too many
        assert not _both(smw, O, root, idx, vals, smw.MerkleProof.from_fields(pr.indices, extra, pr.group_sizes, arity), LABEL, gp, op, arity)
    grp = [g.copy() for g in pr.group_sizes]; grp[0][0] = 0
    assert not _both(smw, O, root, idx, vals, smw.MerkleProof.from_fields(pr.indices, pr.siblings, grp, arity), LABEL, gp, op, arity)
    grp = [g.copy() for g in pr.group_sizes]; grp[0] = grp[0][:-1] if grp[0].shape[0] > 1 else np.concatenate([grp[0], grp[0]])
    assert not _both(smw, O, root, idx, vals, smw.MerkleProof.from_fields(pr.indices, pr.siblings, grp, arity), LABEL, gp, op, arity)
    # proof for other indices than the ones claimed
    other = [i for i in range(n) if i not in idx][:1]
    if other:
        assert not _both(smw, O, root, other, leaves[other], pr, LABEL, gp, op, arity)
    # empty request, length mismatch
    assert not smw.verify_many_ds(root, [], leaves[:0], pr, LABEL, gp)
    assert not smw.verify_many_ds(root, idx, vals[:-1], pr, LABEL, gp)
    # parameters of the wrong width for the proof's arity (merkle/src/lib.rs:609-617)
    wrong = smw.poseidon_params_for_width(ctx, 9 if gp.t != 9 else 17)
    assert not smw.verify_many_ds(root, idx, vals, pr, LABEL, wrong)


def test_verify_all_leaves_and_single_leaf(ctx, smw, O, rnd):
    """opening every leaf leaves no siblings at all; opening one leaf is the classic path"""
    leaves, tree, gp, op = _tree(ctx, smw, O, rnd, 16, 300, 4242)
    idx = list(range(300))
    pr = tree.open_many(idx)
    assert sum(x.shape[0] for x in pr.siblings) == 0
    assert _both(smw, O, tree.root(), idx, leaves, pr, LABEL, gp, op, 16)
    pr1 = tree.open_many([299])
    assert _both(smw, O, tree.root(), [299], leaves[[299]], pr1, LABEL, gp, op, 16)
    assert not _both(smw, O, tree.root(), [298], leaves[[298]], pr1, LABEL, gp, op, 16)


@pytest.mark.parametrize("arity,n", [(2, 256), (4, 100), (8, 64), (2, 1 << 12)])
def test_verify_pairs_ds(ctx, smw, O, rnd, arity, n):
    """new_pairs -> open -> verify_pairs_ds (merkle/src/lib.rs:1053-1084): leaf digests are recomputed from (f, cp) with LEAF_LEVEL_DS"""
    f, cp = rnd(100 + n, n), rnd(200 + n, n)
    cfg = smw.MerkleChannelCfg.new(ctx, arity).with_tree_label(LABEL)
    tree = smw.MerkleTree.new_pairs(f, cp, cfg)
    op = O.Params.for_arity(arity)
    otree = O.MerkleTree.new_pairs(f, cp, arity, LABEL, op)
    assert np.array_equal(tree.root(), otree.root())
    idx = sorted({0, 1, n - 1, n // 2, n // 5})
    pr = tree.open_many(idx)
    pairs = np.stack([f[idx], cp[idx]], axis=1)
    assert smw.verify_pairs_ds(tree.root(), idx, pairs, pr, LABEL, cfg.params)
    # the same statement through the oracle: leaf digests = level 0 of the oracle's tree
    assert O.verify_many_ds(tree.root(), idx, otree.level(0)[idx], pr, LABEL, op, arity)
    perm = idx[::-1]
    assert smw.verify_pairs_ds(tree.root(), perm, np.stack([f[perm], cp[perm]], axis=1), pr, LABEL, cfg.params)
    bad = pairs.copy(); bad[2, 1, 0] ^= np.uint64(1)                               # a wrong cp value
    assert not smw.verify_pairs_ds(tree.root(), idx, bad, pr, LABEL, cfg.params)
    assert not smw.verify_pairs_ds(tree.root(), idx, pairs, pr, LABEL ^ 1, cfg.params)
    swapped = pairs[:, ::-1].copy()                                                 # (cp, f) instead of (f, cp)
    assert not smw.verify_pairs_ds(tree.root(), idx, swapped, pr, LABEL, cfg.params)
    # a pair tree is not a plain tree over f
    assert not smw.verify_many_ds(tree.root(), idx, f[idx], pr, LABEL, cfg.params)


def test_verify_openings_of_fri_layers(ctx, smw, O, rnd):
    """the batched openings fri_prove_queries emits verify against the layer roots (deep_ali/src/fri.rs:700-745's Merkle checks)"""
    n0 = 1 << 12
    st = smw.fri_build_transcript(ctx, rnd(55, n0), smw.FriProverParams([16, 16, 8], 0xDEEFBAAD))
    for l in range(3):
        lay = st.layers[l]
        tree = lay.tree
        n = lay.n
        idx = sorted({0, n - 1, n // 2, 7 % n})
        pr = tree.open_many(idx)
        leaves = tree.level(0)
        gp = smw.poseidon_params_for_arity(ctx, pr.arity)
        assert smw.verify_many_ds(tree.root(), idx, leaves[idx], pr, l, gp)
        assert not smw.verify_many_ds(tree.root(), idx, leaves[idx], pr, l + 1, gp)


def test_verify_deserialised_proof(ctx, smw, O, rnd):
    """prover side: open -> bincode bytes (canonical field encoding from the device conversion); verifier side: bytes -> proof ->
    device verifier.  The bytes equal what the same proof serialises to through the oracle's conversions."""
    from stark_mlwe_b200 import wire
    leaves, tree, gp, op = _tree(ctx, smw, O, rnd, 16, 3000, 991)
    idx = [0, 1, 17, 2999]
    pr = tree.open_many(idx)
    blob = wire.merkle_proof_to_bincode(pr, ctx.to_canonical)
    assert blob == wire.merkle_proof_to_bincode(pr, O.from_mont_c) and len(blob) == wire.merkle_proof_size_bincode(pr)
    back = wire.merkle_proof_from_bincode(blob, ctx.from_canonical, smw.MerkleProof.from_fields)
    assert smw.verify_many_ds(tree.root(), idx, leaves[idx], back, LABEL, gp)
    flipped = bytearray(blob); flipped[8 + 8 * 4 + 8 + 8 + 8] ^= 1              # lowest byte of the first sibling
    bad = wire.merkle_proof_from_bincode(bytes(flipped), ctx.from_canonical, smw.MerkleProof.from_fields)
    assert not smw.verify_many_ds(tree.root(), idx, leaves[idx], bad, LABEL, gp)
