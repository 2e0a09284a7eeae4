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
