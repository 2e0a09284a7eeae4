"""CPU: the oracle still reproduces the committed golden fixtures (tests/golden/, written by make_golden.py), and the
reference's published proof_bytes table is the one the KAT test uses."""
import json
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
G = json.load(open(os.path.join(HERE, "golden", "oracle_golden.json")))
PB = json.load(open(os.path.join(HERE, "golden", "proof_bytes.json")))


def test_proof_bytes_table_is_the_reference_csv():
    from test_oracle_anchors import CSV
    assert {int(k): v for k, v in PB["proof_bytes"].items()} == CSV


def test_oracle_reproduces_lde_commit_roots(O):
    from bench import gen_columns, SEED, SHIFT
    shift = O.to_mont([SHIFT])[0]
    for log_n in ("4", "8", "12"):
        cols = gen_columns(SEED, 0, 2, 1 << int(log_n))
        got = [O.fr_hex(O.commitment_commit(O.lde(cols[c], 2, shift), c).root()) for c in range(2)]
        assert got == G["lde_commit_roots_blowup4_shift5"][log_n]


def test_oracle_reproduces_fri_mf_permute(O):
    from bench import gen_columns
    for name in ("paper_k11", "fold2_k12"):
        e = G["fri"][name]
        f0 = gen_columns(0x5EED0003, 0, 1, 1 << e["k"])[0]
        st = O.FriProverState(f0, e["schedule"], 0xDEEFBAAD)
        assert [O.fr_hex(r) for r in st.roots()] == e["roots"]
        assert [O.fr_hex(st.z(l)) for l in range(len(e["schedule"]))] == e["z"]
    m = G["mf"]
    roots, rs, fin = O.mf_commit_phase(gen_columns(0x5EED0006, 0, 1, 1 << m["k"])[0], m["ds_tag"])
    assert [O.fr_hex(r) for r in roots] == m["roots"] and [O.fr_hex(r) for r in rs] == m["r"] and O.fr_hex(fin) == m["final"]
    for t in (9, 17, 33):
        out = O.permute(O.Params.for_width(t), O.to_mont(range(1, t + 1)).reshape(1, t, 4))
        assert [O.fr_hex(out[0, 0]), O.fr_hex(out[0, t - 1])] == G["permute_1_to_t"][str(t)]
