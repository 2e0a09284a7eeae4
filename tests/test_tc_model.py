"""The algebra behind the tensor-core hash pipeline (csrc/poseidon_tc.cuh), checked on the CPU with Python integers against the oracle's
permutation (poseidon/src/lib.rs:219-272): the grouping of the partial rounds (dots e0^T Mh^k v, gamma corrections, one (t + G) x (t + G)
matrix per group) for the two widths and group sizes the library uses, and the byte GEMM with reduced constants plus the fold of the bits above
2^254.  No GPU: the kernels themselves are compared with the oracle in tests/test_gpu_poseidon_tc.py."""
import os
import sys

import pytest

sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tools"))
import tc_model as M          # noqa: E402
from oracle import oracle as O  # noqa: E402


@pytest.mark.parametrize("t,G,gemm", [(17, 8, "matvec"), (9, 6, "matvec"), (9, 6, "byte_gemm")])
def test_grouped_rounds_and_byte_gemm_match_the_oracle(t, G, gemm):
    prm = O.Params.for_width(t)
    assert prm.rp % G == 0
    st = O.StdRng.seed_from_u64(5 + t).fr_rand(t)
    want = O.from_mont(O.permute(prm, st.reshape(1, t, 4))[0])
    got = M.model(prm, O.from_mont(st), G, getattr(M, gemm))
    assert want == got
