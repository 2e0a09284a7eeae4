"""The reference's own published data reproduced by the GPU path (no oracle in the loop for the number itself):
crates/channel/benchmarkdata.csv:2-9, column proof_bytes, k = 11..18 (every published row), on the bench's seeded inputs
(channel/benches/end_to_end.rs:187-260) through deep_ali_build_f0 -> fri_build_transcript -> fri_prove_queries ->
deep_fri_proof_size_bytes (deep_ali/src/fri.rs:535-641, 355-466, 775-805) — every sponge, Poseidon constant, DS label,
fold, DEEP merge, ChaCha12 draw and multiproof opening runs in libsmw.so."""
import numpy as np
import pytest

from test_oracle_anchors import CSV, A12, kat_inputs

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("k", [11, 12, 13, 14, 15, 16, 17, 18])
def test_proof_bytes_kat_on_gpu(ctx, smw, O, k):
    a, s, e, t = kat_inputs(O, k)              # the oracle only replays rand's StdRng to produce the INPUTS
    st, proofs = smw.deep_fri_prove(ctx, a, s, e, t, [16, 16, 8], 32, 0xDEEFBAAD)
    assert proofs.size_bytes() == CSV[k]
    if k in A12:
        assert O.fr_hex(st.roots()[0]) == A12[k]


def test_query_phase_pieces_match_oracle(ctx, smw, O, rnd):
    a, s, e, t = kat_inputs(O, 11)
    f0, z = smw.deep_ali_build_f0(ctx, a, s, e, t)
    f0_o, z_o = O.deep_ali_build_f0(a, s, e, t)
    assert np.array_equal(z, z_o) and np.array_equal(f0, f0_o)
    seeds = rnd(77, 50)
    want = [O.StdRng.from_seed(int(O.from_mont(seeds[i:i + 1])[0]).to_bytes(32, "little")).next_u64() & 1023 for i in range(50)]
    assert smw.index_from_seed(ctx, seeds, 1024) == want
