"""The tensor-core node hash (csrc/poseidon_tc.cuh: linear layers of the t = 17 permutation as INT8 GEMMs on tcgen05, S-boxes and
Montgomery reductions on the FP64 pipe) against the CPU oracle, level by level: MerkleTree::new with arity 16
(merkle/src/lib.rs:164-179, poseidon/src/lib.rs:219-312).  Bit-exact."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


@pytest.fixture()
def tc_ctx(ctx):
    ctx.set_tc_min_nodes(1)          # every full arity-16 level takes the tensor-core pipeline
    yield ctx
    ctx.set_tc_min_nodes(16384)      # the library's default


@pytest.mark.parametrize("n,label", [(16, 0), (16 * 5, 3), (16 * 128, 7), (16 * 300, 0), (1 << 16, 11)])
def test_tree_levels_match_oracle(tc_ctx, smw, O, rnd, n, label):
    leaves = rnd(4000 + n, n)
    leaves[0] = 0
    leaves[1] = O.ints_to_limbs([(O.P - 1) * O.R % O.P])[0]
    cfg = smw.MerkleChannelCfg.new(tc_ctx, 16).with_tree_label(label)
    launches0 = tc_ctx.launch_count()
    got = smw.MerkleTree.new(leaves, cfg)
    assert tc_ctx.launch_count() - launches0 > 40          # the staged pipeline ran (one level of the plain path is a single launch)
    want = O.MerkleTree.new(leaves, 16, label)
    assert got.height() + 1 == want.nlevels()
    for l in range(1, want.nlevels()):
        assert np.array_equal(got.level(l), want.level(l)), "level %d" % l
    assert np.array_equal(got.root(), want.root())


def test_commit_columns_matches_oracle(tc_ctx, smw, O, rnd):
    """the cfg2 shape in small: several columns hashed by one launch per level (labels label0 + column)"""
    n, n_cols = 1 << 12, 5
    cols = rnd(77, n * n_cols).reshape(n_cols, n, 4)
    p = smw.poseidon_params_for_arity(tc_ctx, 16)
    roots = smw.merkle_commit_columns(tc_ctx, cols, 16, p, label0=9, label_stride=1)
    for c in range(n_cols):
        assert np.array_equal(roots[c], O.MerkleTree.new(cols[c], 16, 9 + c).root())


def test_same_digests_as_the_interpreter(ctx, smw, rnd):
    leaves = rnd(5, 1 << 14)
    cfg = smw.MerkleChannelCfg.new(ctx, 16)
    try:
        ctx.set_tc_min_nodes(0)                  # never: thread-per-node interpreter / cooperative kernels
        a = smw.MerkleTree.new(leaves, cfg).root()
        ctx.set_tc_min_nodes(1)
        b = smw.MerkleTree.new(leaves, cfg).root()
    finally:
        ctx.set_tc_min_nodes(16384)
    assert np.array_equal(a, b)


@pytest.mark.parametrize("n", [1, 127, 129, 5000])
def test_hashed_leaves_match_oracle(tc_ctx, smw, O, rnd, n):
    """hash_leaf_pair (deep_ali/src/fri.rs:38-44): one transcript-parameter permutation per leaf, head kernel from the 17-lane template"""
    f, s = rnd(600 + n, n), rnd(700 + n, n)
    f[0] = 0
    got = smw.hash_leaf_pair(tc_ctx, f, s)
    want = O.hash_leaf_pair(f, s)
    assert np.array_equal(got, want)


@pytest.mark.parametrize("n", [16 * 40, 16 * 300 + 7])
def test_pair_leaf_tree_arity16_matches_oracle(tc_ctx, smw, O, rnd, n):
    """MerkleTree::new_pairs with arity 16 (merkle/src/lib.rs:382-400): pair leaves and (full) node levels on the staged pipeline"""
    f, cp = rnd(800 + n, n), rnd(900 + n, n)
    cfg = smw.MerkleChannelCfg.new(tc_ctx, 16).with_tree_label(5)
    got = smw.MerkleTree.new_pairs(f, cp, cfg)
    want = O.MerkleTree.new_pairs(f, cp, 16, 5)
    for l in range(want.nlevels()):
        assert np.array_equal(got.level(l), want.level(l)), "level %d" % l


def test_fri_paper_schedule_matches_oracle_with_tc_everywhere(tc_ctx, smw, O, rnd):
    """fri_build_transcript (deep_ali/src/fri.rs:231-312) on [16, 16, 8]: hashed leaves, pair leaves and tree levels of every layer"""
    n0 = 1 << 12
    f0 = rnd(4242, n0)
    st = smw.fri_build_transcript(tc_ctx, f0, smw.FriProverParams([16, 16, 8], 0xDEEFBAAD))
    assert np.array_equal(st.roots(), O.FriProverState(f0, [16, 16, 8], 0xDEEFBAAD).roots())
    sharded = smw.fri_roots_sharded(tc_ctx, None, f0.reshape(4, n0 // 4, 4), [i * (n0 // 4) for i in range(4)], n0, smw.FriProverParams([16, 16, 8], 0xDEEFBAAD))
    assert np.array_equal(sharded, st.roots())


# ---- t = 9 (arity-2 trees: the fold-2 FRI layers of BASELINE configs[2] and the Merkle-folded sum-check) on the same pipeline
@pytest.mark.parametrize("n,label", [(2, 0), (2 * 129, 3), (600, 7), (1 << 13, 11)])
def test_arity2_tree_levels_match_oracle(tc_ctx, smw, O, rnd, n, label):
    """MerkleTree::new with arity 2 (merkle/src/lib.rs:164-179): one t = 9 permutation per node (4 DS fields + 2 children + padding one)"""
    leaves = rnd(5000 + n, n)
    leaves[0] = 0
    leaves[1] = O.ints_to_limbs([(O.P - 1) * O.R % O.P])[0]
    cfg = smw.MerkleChannelCfg.new(tc_ctx, 2).with_tree_label(label)
    launches0 = tc_ctx.launch_count()
    got = smw.MerkleTree.new(leaves, cfg)
    assert tc_ctx.launch_count() - launches0 >= 29         # the staged pipeline ran: 29 launches per batch of one-permutation t = 9 hashes
    want = O.MerkleTree.new(leaves, 2, label)
    assert got.height() + 1 == want.nlevels()
    for l in range(1, want.nlevels()):
        assert np.array_equal(got.level(l), want.level(l)), "level %d" % l


@pytest.mark.parametrize("n", [2 * 64, 2 * 300 + 1, 1 << 12])
def test_pair_leaf_tree_arity2_matches_oracle(tc_ctx, smw, O, rnd, n):
    """MerkleTree::new_pairs with arity 2 (merkle/src/lib.rs:382-400): t = 9 pair leaves and node levels on the staged pipeline"""
    f, cp = rnd(1800 + n, n), rnd(1900 + n, n)
    cfg = smw.MerkleChannelCfg.new(tc_ctx, 2).with_tree_label(6)
    got = smw.MerkleTree.new_pairs(f, cp, cfg)
    want = O.MerkleTree.new_pairs(f, cp, 2, 6)
    for l in range(want.nlevels()):
        assert np.array_equal(got.level(l), want.level(l)), "level %d" % l


def test_arity2_same_digests_as_the_interpreter(ctx, smw, rnd):
    leaves = rnd(6, 1 << 12)
    cfg = smw.MerkleChannelCfg.new(ctx, 2)
    try:
        ctx.set_tc_min_nodes(0)
        a = smw.MerkleTree.new(leaves, cfg).root()
        ctx.set_tc_min_nodes(1)
        b = smw.MerkleTree.new(leaves, cfg).root()
    finally:
        ctx.set_tc_min_nodes(16384)
    assert np.array_equal(a, b)


def test_fri_fold2_schedule_matches_oracle_with_tc_everywhere(tc_ctx, smw, O, rnd):
    """fri_build_transcript (deep_ali/src/fri.rs:231-312) with fold factor 2 down to one element: every layer an arity-2 pair-leaf tree"""
    n0 = 1 << 11
    f0 = rnd(4343, n0)
    sched = [2] * 11
    st = smw.fri_build_transcript(tc_ctx, f0, smw.FriProverParams(sched, 0xDEEFBAAD))
    assert np.array_equal(st.roots(), O.FriProverState(f0, sched, 0xDEEFBAAD).roots())
    sharded = smw.fri_roots_sharded(tc_ctx, None, f0.reshape(4, n0 // 4, 4), [i * (n0 // 4) for i in range(4)], n0, smw.FriProverParams(sched, 0xDEEFBAAD))
    assert np.array_equal(sharded, st.roots())


def test_custom_round_count_stays_on_the_interpreter(tc_ctx, smw, O, rnd):
    """a parameter set whose rp is not a multiple of the group size (here t = 17, rp = 60; groups of 8) cannot take the staged pipeline: the level
    launches fall back to the thread-per-node kernels instead of failing, and the roots still equal the oracle's"""
    n, n_cols = 16 * 64, 2
    cols = rnd(91, n * n_cols).reshape(n_cols, n, 4)
    p = smw.params_from_seed(tc_ctx, b"TC-FALLBACK-SEED", 17, 60)
    assert tc_ctx.tc_node_counts(p)["fp64_instr"] == 0
    roots = smw.merkle_commit_columns(tc_ctx, cols, 16, p, label0=2, label_stride=1)
    po = O.Params.from_seed(b"TC-FALLBACK-SEED", 17, 60)
    for c in range(n_cols):
        assert np.array_equal(roots[c], O.MerkleTree.new(cols[c], 16, 2 + c, po).root())
