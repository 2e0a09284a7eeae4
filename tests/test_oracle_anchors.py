"""Pins the CPU oracle (oracle/oracle.c) before anything is compared against it.

 * test_proof_bytes_kat: the reference's own published data, crates/channel/benchmarkdata.csv:2-9 column
   `proof_bytes`, reproduced on the bench's seeded inputs (crates/channel/benches/end_to_end.rs:187-260).
 * the Appendix-B anchors of SURVEY.md (derived values; A9b is the public ChaCha12 zero-key vector).
 * BLAKE3 against the independent `blake3` Python package.
"""
import numpy as np
import pytest

CSV = {11: 39592, 12: 52000, 13: 60968, 14: 72936, 15: 87736, 16: 101976, 17: 119952, 18: 140032}
A12 = {11: "0x0b425ae39aa0516a419705d539933ffc0cc05eca4d32ae47d79d229f3b651c46",
       12: "0x21b2c56ae883f913414ad3c3712e4a36530bdc2b87cf414dafef231368db8318",
       13: "0x25a4b735b205dd46ee6e2e04a72b6a9a6b0dd01063e412b460b58c7dff1dfe00"}


def kat_inputs(O, k_target):
    """a, s, e, t for the "paper" preset at k (end_to_end.rs:214,243-253)"""
    seed = 1337
    for k in range(11, k_target + 1):
        seed = (seed * 1103515245 + 12345) & 0xFFFFFFFFFFFFFFFF
    g = O.StdRng.seed_from_u64(seed)
    n0 = 1 << k_target
    return g.fr_rand(n0), g.fr_rand(n0), g.fr_rand(n0), g.fr_rand(n0)


@pytest.mark.parametrize("k", [11, 12, 13])
def test_proof_bytes_kat(O, k):
    a, s, e, t = kat_inputs(O, k)
    size, roots = O.deep_fri_prove_size(a, s, e, t, [16, 16, 8], 32, 0xDEEFBAAD)
    assert size == CSV[k]
    assert O.fr_hex(roots[0]) == A12[k]


@pytest.mark.slow
@pytest.mark.parametrize("k", [14, 15, 16, 17, 18])
def test_proof_bytes_kat_slow(O, k):
    a, s, e, t = kat_inputs(O, k)
    size, _ = O.deep_fri_prove_size(a, s, e, t, [16, 16, 8], 32, 0xDEEFBAAD)
    assert size == CSV[k]


def test_blake3_against_package(O):
    b3 = pytest.importorskip("blake3")
    for n in [0, 1, 31, 55, 57, 58, 63, 64, 65, 127, 128, 129, 500, 1024]:
        d = bytes((i * 7 + 3) & 255 for i in range(n))
        assert O.blake3(d) == b3.blake3(d).digest()


def test_anchor_constants(O):
    assert O.fr_hex(O.Params.commitment_default().export()[0][0, 0]) == "0x148f5dcc6124c3773271264effe5afe6f2fff544d74a67ccfa20f23fa13afcd3"
    assert O.fr_hex(O.Params.transcript_default().export()[0][0, 0]) == "0x009cf7c66ca352ebad4a177d79582c222590c08e8e3f1183188c31e648951c41"
    assert O.fr_hex(O.Params.for_width(9).export()[0][0, 0]) == "0x31969836352d3ae3b0def18064eb5a75d42a97f0b95b74934f978fe4b1a65c6f"


def test_anchor_permute_and_sponges(O):
    pc = O.Params.commitment_default()
    out = O.permute(pc, O.to_mont(range(1, 18)).reshape(1, 17, 4))
    assert O.fr_hex(out[0, 0]) == "0x29ec0039891f2402ddd7312d5dcfad61a3bd9f7c125de9286b6460e6d8088d8a"
    assert O.fr_hex(out[0, 16]) == "0x325f385ebdd97e2a2ef7e298c56984b829713e53199b7aeacc04cd036e0f0e13"
    p17, p9 = O.Params.for_width(17), O.Params.for_width(9)
    assert O.fr_hex(O.hash_with_ds_dynamic(p17, O.to_mont([16, 0, 3, 42]), O.to_mont(range(1, 17)))) == "0x1db81a1aeeca1af12104fabd427983bc78f9a93740ca708cc7b5935028587d32"
    assert O.fr_hex(O.hash_with_ds_dynamic(p9, O.to_mont([8, 2, 5, 7]), O.to_mont(range(11, 19)))) == "0x094ff8105c4a51766bf2019269147483aef185b18066735fff2d7ae7d2509d4b"
    assert O.fr_hex(O.hash_with_ds_dynamic(p9, O.to_mont([2, 0xFFFFFFFF, 0, 0]), O.to_mont([5, 6]))) == "0x2bcba62578291522220b4ea229ec8b6088999a72be71614ae684bac5293180cb"


def test_anchor_merkle(O):
    assert O.fr_hex(O.MerkleTree.new(O.to_mont(range(1, 33)), 16, 1234).root()) == "0x04d08926c0819dad4f79908de48ddd13f2e5ca558b64c58b23a118dd70416b3a"
    assert O.fr_hex(O.commitment_commit(O.to_mont(range(1, 65)), 123).root()) == "0x0706b07581b02fdb520541942fbc85d3100d818466452a4b84c7a31a89efca37"
    assert O.fr_hex(O.commitment_commit(O.to_mont(range(1, 56)), 2025).root()) == "0x2b848674cbb439cfd5e4869693b1295a2041163a19252c86ff1d59330fe11f24"


def test_anchor_transcript_rng_fri(O):
    assert O.fr_hex(O.hash_leaf_pair(O.to_mont([1]), O.to_mont([2]))[0]) == "0x3042f58a799105a2a4a424423410439a63abeb1f12cbe66f8dc9af4ab164963a"
    assert O.fr_hex(O.tr_hash_fields_tagged(b"FRI/seed", O.to_mont([1, 2, 3]))) == "0x1bb77b74e853c0d1c38fe81adb3c6167e226cd3535e2a0b778102bf64c2efbdb"
    assert [O.from_mont(O.fri_sample_z_ell(0xDEEFBAAD, l, n))[0] for l, n in [(0, 2048), (1, 128), (2, 8)]] == [0xFC0CF1A3D2ADB98E, 0xD6DC1C5AFDD24324, 0x3D365436CC083420]
    z3 = O.to_mont([3])[0]
    assert O.from_mont(O.fri_fold_layer(O.to_mont(range(1, 9)), z3, 2)) == [7, 15, 23, 31]
    assert O.from_mont(O.fri_fold_layer(O.to_mont(range(1, 9)), z3, 4)) == [142, 302]
    g = O.StdRng.seed_from_u64(42)
    assert (g.next_u64(), g.next_u64()) == (0x86CC7763222724A2, 0x8AF00A133FAD517D)
    assert O.StdRng.from_seed(bytes(32)).next_u64() == 0x53F955076A9AF49B      # public ChaCha12 zero-key vector
    assert O.fr_hex(O.StdRng.seed_from_u64(42).fr_rand(1)[0]) == "0x0f32b310b727b973e9de40601a65754795922e0822e28a899b91a748125f08fa"
    assert O.fr_hex(O.root_of_unity(11)) == "0x3a3b5103f6d0105d125b967f811966337bc1e694ef9d23f5fab7afba742345b6"
    assert O.fr_hex(O.root_of_unity(20)) == "0x19e9df871f6c4b14a3615751d1f43923b5f9ce8d08f92d2629c2afb67fb34869"


def test_anchor_fri_build(O):
    S = O.FriProverState(O.to_mont(range(1, 17)), [2, 2, 2, 2], 7)
    assert [O.fr_hex(r) for r in S.roots()] == [
        "0x09768afbdf9b1c98ec891f9c8862333148ab89003a4c40623e604bbbc83b1050", "0x2b85f833eb35727e4947dad28fb67bf0274829bca4aec750d8ddaad99a5a1787",
        "0x0c5a495617eb90c92e2684eaccf969d985290d28f657017d234f4d1d11a3a4ea", "0x3a9b4751de84b4813ece37f5b3c0610d8276544b1a6761fe3f8e303be6a33d13",
        "0x1f54028c997135f7c8ec7e8bc9cec713a247c9f85f02703b2ca9dfe1a91f82a7"]
    assert O.fr_hex(S.f(4)[0]) == "0x1317b89f738d009c514ba19dae847856b46bb6d6e9168c2fc64c2741d08335d0"
    assert O.fr_hex(S.prove_size(1)[1]) == "0x0f38e923194e6d0d34c0a0f390dffb90b1a8cbb528bb3597d984bdf841441eee"


def test_ntt_against_naive_dft(O):
    """parity unpinned by the reference (no NTT vector exists there): pin the oracle's NTT on the definition evals[i] = P(w^i)"""
    for log_n in [1, 3, 6]:
        c = O.StdRng.seed_from_u64(100 + log_n).fr_rand(1 << log_n)
        ev = O.ntt(c)
        assert np.array_equal(ev, O.dft_naive(c))
        assert np.array_equal(O.ntt(ev, inverse=True), c)
    ones = O.to_mont([1] * 8)       # the reference's only FFT test: all-ones round trip (fft/src/lib.rs:40-54)
    assert np.array_equal(O.ntt(O.ntt(ones), inverse=True), ones)


def test_ntt_and_lde_against_horner_at_full_sizes(O):
    """The definition pin at BASELINE sizes: evals[i] = P(w^i) by Horner's rule (poly/src/lib.rs `evaluate`) at random i, n = 2^16 and
    2^20; the inverse transform returns the coefficients; the coset LDE at index i is P(g w_N^i) for P = ifft(column)."""
    for log_n in (16, 20):
        n = 1 << log_n
        c = O.StdRng.seed_from_u64(4000 + log_n).fr_rand(n)
        ev = O.ntt(c)
        w = O.root_of_unity(log_n)
        rng = np.random.default_rng(log_n)
        for i in [0, 1, n - 1, n // 2] + [int(v) for v in rng.integers(0, n, 6)]:
            assert np.array_equal(ev[i], O.horner(c, O.fr_pow(w, i))), (log_n, i)
        assert np.array_equal(O.ntt(ev, inverse=True), c)
    log_n, lb = 14, 2
    col = O.StdRng.seed_from_u64(77).fr_rand(1 << log_n)
    g = O.to_mont([5])[0]
    ext = O.lde(col, lb, g)
    coef = O.ntt(col, inverse=True)
    wN = O.root_of_unity(log_n + lb)
    for i in (0, 1, 4, 12345, (1 << (log_n + lb)) - 1):
        x = O.fr_mul(g.reshape(1, 4), O.fr_pow(wN, i).reshape(1, 4))[0]
        assert np.array_equal(ext[i], O.horner(coef, x)), i
    assert np.array_equal(ext[::4], O.lde(col, 0, g))          # every 4th point of the blow-up-4 coset is the blow-up-1 coset
