"""GPU parity: Poseidon constants, permutation, sponges and Merkle trees of libsmw.so against the CPU oracle
(bit-exact; integer work).  Mirrors the reference's own tests where they exist (poseidon/src/lib.rs:438-481,
merkle/src/lib.rs:913-1183, commitment/src/lib.rs:122-137)."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def eq(a, b):
    return np.array_equal(np.asarray(a, dtype=np.uint64), np.asarray(b, dtype=np.uint64))


@pytest.mark.parametrize("which", ["commitment", "transcript", "w9", "w17", "bench"])
def test_constants_match_oracle(ctx, smw, O, which):
    """derive_mds / derive_rc_full / derive_rc_partial (poseidon/src/lib.rs:176-216) run on the device (BLAKE3 -> Fr)"""
    if which == "commitment":
        g, o = smw.commitment_default_params(ctx), O.Params.commitment_default()
    elif which == "transcript":
        g, o = smw.transcript_default_params(ctx), O.Params.transcript_default()
    elif which == "w9":
        g, o = smw.poseidon_params_for_width(ctx, 9), O.Params.for_width(9)
    elif which == "w17":
        g, o = smw.poseidon_params_for_width(ctx, 17), O.Params.for_width(17)
    else:   # the seed the reference's benches use (merkle/benches/merkle.rs:36)
        g, o = smw.generate_params_t17_x5(ctx, b"POSEIDON-T17-X5"), O.Params.from_seed(b"POSEIDON-T17-X5")
    for a, b in zip(g.export(), o.export()):
        assert eq(a, b)


def test_anchor_A1_and_A13(ctx, smw, O):
    p = smw.commitment_default_params(ctx)
    assert O.fr_hex(p.export()[0][0, 0]) == "0x148f5dcc6124c3773271264effe5afe6f2fff544d74a67ccfa20f23fa13afcd3"
    out = smw.permute(ctx, O.to_mont(range(1, 18)).reshape(1, 17, 4), p)
    assert O.fr_hex(out[0, 0]) == "0x29ec0039891f2402ddd7312d5dcfad61a3bd9f7c125de9286b6460e6d8088d8a"
    assert O.fr_hex(out[0, 16]) == "0x325f385ebdd97e2a2ef7e298c56984b829713e53199b7aeacc04cd036e0f0e13"


@pytest.mark.parametrize("t,batch", [(17, 1), (17, 333), (9, 1), (9, 700), (17, 9000), (9, 2500)])   # <= 8192 / 2048 states: the cooperative kernel; above: one thread per state
@pytest.mark.parametrize("dense", [False, True])
def test_permute_matches_oracle(ctx, smw, O, rnd, t, batch, dense):
    """the production path (sparse partial rounds, LU in-place mat-vec, lazy reduction) and the literal round structure
    both equal the reference's dense permutation"""
    if dense and batch > 400:
        batch = 300
    if batch > 5000:
        O.set_threads(O.max_threads())
    g, o = smw.poseidon_params_for_width(ctx, t), O.Params.for_width(t)
    st = rnd(1000 + t + batch, batch * t).reshape(batch, t, 4)
    st[0, :, :] = 0                                    # all-zero state
    if batch > 1:
        st[1] = O.ints_to_limbs([(O.P - 1) * O.R % O.P] * t)   # p - 1 in every lane
    got = smw.permute(ctx, st, g, dense=dense)
    assert eq(got, O.permute(o, st))


def test_static_dynamic_roundtrip_t17(ctx, smw, O, rnd):
    """poseidon/src/lib.rs:443-460: params from a seed, handed back as caller tables, give the same permutation"""
    a = smw.generate_params_t17_x5(ctx, b"POSEIDON-T17-X5-SEED")
    b = smw.params_from_tables(ctx, *a.export())
    st = rnd(5, 17 * 4).reshape(4, 17, 4)
    assert eq(smw.permute(ctx, st, a), smw.permute(ctx, st, b))
    assert eq(smw.permute(ctx, st, a), O.permute(O.Params.commitment_default(), st))


def test_unsupported_width_and_arity(ctx, smw):
    with pytest.raises(smw.SmwError) as e:
        smw.poseidon_params_for_width(ctx, 10)          # panics at poseidon/src/lib.rs:127
    assert e.value.name == "SMW_ERR_UNSUPPORTED_WIDTH"
    with pytest.raises(smw.SmwError) as e:
        smw.poseidon_params_for_arity(ctx, 129)         # panics at poseidon/src/lib.rs:163
    assert e.value.name == "SMW_ERR_UNSUPPORTED_ARITY"


def test_hash_with_ds_dynamic_anchors_and_random(ctx, smw, O, rnd):
    p17, p9 = smw.poseidon_params_for_width(ctx, 17), smw.poseidon_params_for_width(ctx, 9)
    assert O.fr_hex(smw.hash_with_ds_dynamic(ctx, O.to_mont([16, 0, 3, 42]), O.to_mont(range(1, 17)), p17)) == "0x1db81a1aeeca1af12104fabd427983bc78f9a93740ca708cc7b5935028587d32"
    assert O.fr_hex(smw.hash_with_ds_dynamic(ctx, O.to_mont([8, 2, 5, 7]), O.to_mont(range(11, 19)), p9)) == "0x094ff8105c4a51766bf2019269147483aef185b18066735fff2d7ae7d2509d4b"
    assert O.fr_hex(smw.hash_with_ds_dynamic(ctx, O.to_mont([2, 0xFFFFFFFF, 0, 0]), O.to_mont([5, 6]), p9)) == "0x2bcba62578291522220b4ea229ec8b6088999a72be71614ae684bac5293180cb"
    o17, o9 = O.Params.for_width(17), O.Params.for_width(9)
    for n in [0, 1, 11, 12, 13, 28, 44, 45]:            # block boundaries of rate 16 (4 DS + n + pad)
        ds, x = rnd(n + 1, 4), rnd(50 + n, max(n, 1))[:n]
        assert eq(smw.hash_with_ds_dynamic(ctx, ds, x, p17), O.hash_with_ds_dynamic(o17, ds, x))
    for n in [0, 2, 3, 4, 8, 11, 12]:
        ds, x = rnd(n + 7, 4), rnd(90 + n, max(n, 1))[:n]
        assert eq(smw.hash_with_ds_dynamic(ctx, ds, x, p9), O.hash_with_ds_dynamic(o9, ds, x))


def test_legacy_hash_with_ds(ctx, smw, O, rnd):
    g, o = smw.commitment_default_params(ctx), O.Params.commitment_default()
    for n in [1, 16, 17, 40]:
        x, tag = rnd(n, n), rnd(7, 1)[0]
        assert eq(smw.hash_with_ds(ctx, x, tag, g), O.hash_with_ds_legacy(o, x, tag))


@pytest.mark.parametrize("arity,n", [(16, 32), (16, 1), (16, 2), (16, 17), (16, 256), (16, 257), (16, 4097), (8, 64), (8, 100), (4, 64), (2, 64), (2, 33), (3, 10)])
def test_merkle_new_matches_oracle(ctx, smw, O, rnd, arity, n):
    """MerkleTree::new (merkle/src/lib.rs:147-193): every level, ragged last chunks, 1-leaf tree"""
    leaves = rnd(arity * 1000 + n, n)
    label = 1234 + n
    g = smw.MerkleTree.new(leaves, smw.MerkleChannelCfg.new(ctx, arity).with_tree_label(label))
    o = O.MerkleTree.new(leaves, arity, label)
    assert g.height() + 1 == o.nlevels()
    for l in range(o.nlevels()):
        assert eq(g.level(l), o.level(l)), "level %d" % l
    assert eq(g.root(), o.root())


def test_merkle_anchors(ctx, smw, O):
    t = smw.MerkleTree.new(O.to_mont(range(1, 33)), smw.MerkleChannelCfg.new(ctx, 16).with_tree_label(1234))
    assert O.fr_hex(t.root()) == "0x04d08926c0819dad4f79908de48ddd13f2e5ca558b64c58b23a118dd70416b3a"
    scheme = smw.MerkleCommitment(smw.MerkleConfig.with_default_params(ctx, 123))
    root, aux = scheme.commit(O.to_mont(range(1, 65)))
    assert O.fr_hex(root) == "0x0706b07581b02fdb520541942fbc85d3100d818466452a4b84c7a31a89efca37"
    root, _ = smw.MerkleCommitment(smw.MerkleConfig.with_default_params(ctx, 2025)).commit(O.to_mont(range(1, 56)))
    assert O.fr_hex(root) == "0x2b848674cbb439cfd5e4869693b1295a2041163a19252c86ff1d59330fe11f24"


@pytest.mark.parametrize("arity,n", [(2, 16), (2, 1), (4, 64), (8, 24), (16, 48)])
def test_merkle_new_pairs_matches_oracle(ctx, smw, O, rnd, arity, n):
    """MerkleTree::new_pairs (merkle/src/lib.rs:392-445), leaf DS level = u32::MAX"""
    f, cp = rnd(arity + n, n), rnd(arity + n + 99, n)
    g = smw.MerkleTree.new_pairs(f, cp, smw.MerkleChannelCfg.new(ctx, arity).with_tree_label(5))
    o = O.MerkleTree.new_pairs(f, cp, arity, 5)
    for l in range(o.nlevels()):
        assert eq(g.level(l), o.level(l))


def test_merkle_error_behaviour(ctx, smw, rnd):
    cfg = smw.MerkleChannelCfg.new(ctx, 16)
    with pytest.raises(smw.SmwError) as e:
        smw.MerkleTree.new(np.zeros((0, 4), np.uint64), cfg)                     # assert!(!leaves.is_empty()) merkle/src/lib.rs:148
    assert e.value.name == "SMW_ERR_EMPTY"
    with pytest.raises(smw.SmwError) as e:                                       # arity/width assert merkle/src/lib.rs:161
        smw.MerkleTree.new(rnd(1, 8), smw.MerkleChannelCfg.with_params(ctx, 8, smw.poseidon_params_for_width(ctx, 17)))
    assert e.value.name == "SMW_ERR_ARITY_WIDTH"
    with pytest.raises(smw.SmwError) as e:                                       # length mismatch merkle/src/lib.rs:393
        smw.MerkleTree.new_pairs(rnd(1, 8), rnd(2, 7), smw.MerkleChannelCfg.new(ctx, 2))
    assert e.value.name == "SMW_ERR_LENGTH_MISMATCH"


def test_ds_hygiene(ctx, smw, rnd):
    """merkle/src/lib.rs:966-1010, 1053-1084: tree label, arity and leaf position all separate roots"""
    leaves = rnd(77, 64)
    r = lambda ar, lab, lv=leaves: smw.MerkleTree.new(lv, smw.MerkleChannelCfg.new(ctx, ar).with_tree_label(lab)).root()
    assert not eq(r(16, 1), r(16, 2))
    assert not eq(r(16, 1), r(8, 1))
    swapped = leaves.copy(); swapped[[0, 1]] = swapped[[1, 0]]
    assert not eq(r(16, 1), r(16, 1, swapped))
    assert eq(r(16, 1), r(16, 1))


def test_commit_columns_matches_per_column_trees(ctx, smw, O, rnd):
    cols = rnd(4242, 5 * 300).reshape(5, 300, 4)
    p = smw.commitment_default_params(ctx)
    roots = smw.merkle_commit_columns(ctx, cols, 16, p, label0=10, label_stride=3)
    for c in range(5):
        assert eq(roots[c], O.MerkleTree.new(cols[c], 16, 10 + 3 * c, O.Params.commitment_default()).root())


def test_unsupported_round_counts_and_arity_one_are_rejected(ctx, smw, O, rnd):
    """the sparse two-round blocks need an even rp, the lazy-reduction bound rp <= 96: anything else is an error,
    not a wrong digest; arity 1 over more than one leaf never terminates in the reference (merkle/src/lib.rs:164) -> error"""
    for rp in (61, 63, 98, 200):
        with pytest.raises(smw.SmwError) as e:
            smw.params_from_seed(ctx, b"POSEIDON-T17-X5-SEED", 17, rp)
        assert e.value.code == 8
    p = smw.params_from_seed(ctx, b"SOME-OTHER-SEED", 9, 2)                 # small even rp: sparse path == dense path == oracle
    st = rnd(31, 9 * 5).reshape(5, 9, 4)
    want = O.permute(O.Params.from_seed(b"SOME-OTHER-SEED", 9, 2), st)
    assert np.array_equal(smw.permute(ctx, st, p), want) and np.array_equal(smw.permute(ctx, st, p, dense=True), want)
    cfg = smw.MerkleChannelCfg.new(ctx, 1)
    with pytest.raises(smw.SmwError) as e:
        smw.MerkleTree.new(rnd(1, 4), cfg)
    assert e.value.code == 7
    assert np.array_equal(smw.MerkleTree.new(rnd(1, 1), cfg).root(), rnd(1, 1)[0])


@pytest.mark.parametrize("arity,n_cols,n", [(16, 3, 1 << 12), (16, 70, 256), (2, 5, 1 << 11), (8, 2, 4097), (4, 1, 3000)])
def test_commit_columns_across_kernel_boundary(ctx, smw, O, rnd, arity, n_cols, n):
    """column batches whose lower levels take the thread-per-node kernel and whose tops take the cooperative one (<= 4096 / 2048
    nodes per launch): every root against the oracle"""
    cols = rnd(arity * 100 + n_cols, n_cols * n).reshape(n_cols, n, 4)
    p = smw.poseidon_params_for_arity(ctx, arity)
    got = smw.merkle_commit_columns(ctx, cols, arity, p, label0=11, label_stride=3)
    op = O.Params.for_arity(arity)
    for c in range(n_cols):
        assert np.array_equal(got[c], O.MerkleTree.new(cols[c], arity, 11 + 3 * c, op).root()), c
