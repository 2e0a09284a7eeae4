"""CPU: the control flow of stark_mlwe_b200/channel.py (ProverChannel, SumCheckMFProver / Verifier, MerkleProver) run over the
oracle kit alone — the prover's messages verify, tampered ones do not, and the prover's claim equals the plain sum of the
table (channel/src/lib.rs:626-631), which the module derives as 2 c0 + c1."""
import os
import sys

import numpy as np
import pytest

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
from oracle_kit import OracleKit, run_prover   # noqa: E402


@pytest.mark.parametrize("k,q", [(4, 2), (6, 3)])
def test_mf_round_trip_over_oracle(O, rnd, k, q):
    from stark_mlwe_b200 import channel as ch
    kit = OracleKit(O)
    table = rnd(1234 + k, 1 << k)
    g = run_prover(ch, kit, table, q)
    total = sum(O.from_mont(table)) % O.P
    assert O.from_mont(g["claim"].reshape(1, 4))[0] == total
    vchan = ch.VerifierChannel(kit.transcript(b"MF-TEST"))
    ver = ch.SumCheckMFVerifier(ch.SumCheckMFConfig(q), ch.MerkleChannelCfg(kit, tree_label=0xABCD), vchan, g["root0"], k)
    ver.receive_initial_root(g["root0"])
    ver.recv_claim(g["claim"])
    s_prev = g["claim"]
    for i, (c0, c1, r_i, next_root, op) in enumerate(g["rounds"]):
        ver.start_round(i, s_prev, c0, c1)
        prev_root = ver.current_root().copy()
        assert np.array_equal(ver.derive_round_challenge(i), r_i)
        ver.recv_next_root(next_root)
        half = 1 << (k - i - 1)
        q_target, seen, j = min(max(q, 1), half), set(), 0
        while len(seen) < q_target and j < max(q_target * 16, 16):
            limbs = kit.to_canonical(vchan.challenge_scalar(b"sumcheck-mf/q" + int(i).to_bytes(8, "little") + int(j).to_bytes(8, "little")))
            seen.add((int(limbs[0]) ^ int(limbs[1]) ^ int(limbs[2]) ^ int(limbs[3])) % half)
            j += 1
        assert sorted(seen) == op.next_indices
        assert op.cur_indices == [x for jj in op.next_indices for x in (2 * jj, 2 * jj + 1)]
        assert ver.verify_fold_openings(op.cur_indices, op.cur_values, op.cur_proof, op.next_indices, op.next_values, op.next_proof, r_i, prev_root, next_root)
        bad = op.next_values.copy(); bad[0, 0] ^= np.uint64(1)
        assert not ver.verify_fold_openings(op.cur_indices, op.cur_values, op.cur_proof, op.next_indices, bad, op.next_proof, r_i, prev_root, next_root)
        assert not ver.verify_fold_openings(op.cur_indices[:-1], op.cur_values[:-1], op.cur_proof, op.next_indices, op.next_values, op.next_proof, r_i, prev_root, next_root)
        vchan.recv_opening(op.cur_indices, op.cur_values, op.cur_proof)
        vchan.recv_opening(op.next_indices, op.next_values, op.next_proof)
        s_prev = ver.compute_s_next(c0, c1, r_i)
    ver.finalize_and_check(g["final"], s_prev)
    assert np.array_equal(vchan.challenge_scalar(b"tail"), g["tail"])
