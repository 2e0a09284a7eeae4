"""GPU parity at the sizes BASELINE.json's configs name, against goldens frozen from the pinned CPU oracle
(tests/golden/full_size_golden.json, written by tests/golden/make_golden_full.py; the oracle itself is pinned by the reference's
published proof_bytes, tests/test_oracle_anchors.py) and against the DEFINITION of the transform (Horner evaluation).

    cfg2   one full column 2^20 -> 2^22 (blow-up 4, shift 5): commitment root                  commitment/src/lib.rs:85-90, merkle/src/lib.rs:147-193
    cfg3   2^22-element polynomial, fold factor 2 to one element and the published [16,16,8]:     deep_ali/src/fri.rs:231-312
           every layer root + roots_seed
    cfg4   one column at blow-up 8 (2^22 -> 2^25, and 2^24 -> 2^27 when the golden holds it)
    NTT    2^22 / 2^24 points against P(w^i) evaluated by the oracle's Horner loop                  fft/src/lib.rs:6-32 (ark-poly convention)
"""
import json
import os

import numpy as np
import pytest

from bench import gen_columns, SHIFT

pytestmark = pytest.mark.gpu
GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "full_size_golden.json")
SEED_Z = 0xDEEFBAAD


def gold(section):
    if not os.path.exists(GOLD):
        pytest.skip("tests/golden/full_size_golden.json missing")
    d = json.load(open(GOLD))
    if section not in d:
        pytest.skip("golden section %s not generated" % section)
    return d[section]


def hexes(O, arr):
    return [O.fr_hex(r) for r in np.asarray(arr).reshape(-1, 4)]


@pytest.mark.parametrize("section", ["cfg3_fold2", "cfg3_paper"])
def test_cfg3_fri_roots_match_golden(ctx, smw, O, section):
    g = gold(section)
    f0 = gen_columns(0x5EED0003, 0, 1, g["n0"])[0]
    st = smw.fri_build_transcript(ctx, f0, smw.FriProverParams(g["schedule"], g["seed_z"]))
    roots = st.roots()
    assert hexes(O, roots) == g["roots"]                                     # every layer, including the 2^22-leaf ones
    assert [lay.arity for lay in st.layers] == g["arity"]
    assert O.fr_hex(smw.tr_hash_fields_tagged(ctx, b"FRI/seed", roots)) == g["roots_seed"]
    if g["final_f"]:
        assert O.fr_hex(st.layers[-1].f[0]) == g["final_f"]


@pytest.mark.parametrize("section", ["cfg2", "cfg4_slice", "cfg4"])
def test_full_column_lde_commit_root_matches_golden(ctx, smw, O, section):
    g = gold(section)
    shift = O.to_mont([SHIFT])[0]
    for col, want in g["roots"].items():
        cols = gen_columns(g["seed"], int(col), 1, 1 << g["log_n"])
        got = smw.lde_commit(ctx, cols, g["log_blowup"], shift, label0=int(col))
        assert O.fr_hex(got[0]) == want, (section, col)


@pytest.mark.parametrize("log_n", [22, 24])
def test_ntt_matches_definition_by_horner(ctx, smw, O, log_n):
    """evals[i] = P(w^i): the GPU transform of 2^22 / 2^24 coefficients against the oracle's Horner evaluation at a handful of i,
    and the inverse transform returns the coefficients"""
    n = 1 << log_n
    c = gen_columns(0x5EED0005, 0, 1, n)[0]
    ev = smw.fft(ctx, c)
    w = O.root_of_unity(log_n)
    rng = np.random.default_rng(log_n)
    for i in [0, 1, n - 1] + [int(v) for v in rng.integers(0, n, 3)]:
        assert np.array_equal(ev[i], O.horner(c, O.fr_pow(w, i))), (log_n, i)
    assert np.array_equal(smw.ifft(ctx, ev), c)


def test_lde_matches_definition_by_horner_2p20(ctx, smw, O):
    log_n, lb = 20, 2
    col = gen_columns(0x5EED0002, 7, 1, 1 << log_n)[0]
    g = O.to_mont([SHIFT])[0]
    ext = smw.lde(ctx, col.reshape(1, -1, 4), lb, g)[0]
    coef = smw.ifft(ctx, col)
    wN = O.root_of_unity(log_n + lb)
    for i in (0, 1, 3, 2 ** 21 + 12345, (1 << (log_n + lb)) - 1):
        x = O.fr_mul(g.reshape(1, 4), O.fr_pow(wN, i).reshape(1, 4))[0]
        assert np.array_equal(ext[i], O.horner(coef, x)), i
