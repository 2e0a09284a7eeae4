"""GPU against the committed golden fixtures (tests/golden/oracle_golden.json): no oracle call in these tests, so they
also cover sizes (2^16-row columns, k = 14/15 FRI, t = 65/129) that would be slow to recompute on the CPU per run."""
import json
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
HERE = os.path.dirname(os.path.abspath(__file__))
G = json.load(open(os.path.join(HERE, "golden", "oracle_golden.json")))


def test_lde_commit_roots(ctx, smw, O):
    from bench import gen_columns, SEED, SHIFT
    shift = ctx.from_canonical(np.array([[SHIFT, 0, 0, 0]], dtype=np.uint64))[0]
    for log_n, want in G["lde_commit_roots_blowup4_shift5"].items():
        cols = gen_columns(SEED, 0, 2, 1 << int(log_n))
        roots = smw.lde_commit(ctx, cols, 2, shift, label0=0)
        assert [O.fr_hex(r) for r in roots] == want


def test_fri_roots_and_z(ctx, smw, O):
    from bench import gen_columns
    for name, e in G["fri"].items():
        f0 = gen_columns(0x5EED0003, 0, 1, 1 << e["k"])[0]
        st = smw.fri_build_transcript(ctx, f0, smw.FriProverParams(e["schedule"], 0xDEEFBAAD))
        assert [O.fr_hex(r) for r in st.roots()] == e["roots"], name
        assert [O.fr_hex(z) for z in st.z_layers] == e["z"], name


def test_mf_commit_phase(ctx, smw, O):
    from bench import gen_columns
    m = G["mf"]
    roots, rs, fin = smw.mf_commit_phase(ctx, gen_columns(0x5EED0006, 0, 1, 1 << m["k"])[0], m["ds_tag"])
    assert [O.fr_hex(r) for r in roots] == m["roots"] and [O.fr_hex(r) for r in rs] == m["r"] and O.fr_hex(fin) == m["final"]


@pytest.mark.parametrize("t", [9, 17, 33, 65, 129])
def test_permute_all_widths(ctx, smw, O, t):
    p = smw.poseidon_params_for_width(ctx, t)
    st = O.to_mont(range(1, t + 1)).reshape(1, t, 4)
    for dense in (False, True):
        out = smw.permute(ctx, st, p, dense=dense)
        assert [O.fr_hex(out[0, 0]), O.fr_hex(out[0, t - 1])] == G["permute_1_to_t"][str(t)]
