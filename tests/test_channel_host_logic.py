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


@pytest.mark.parametrize("k,q", [(4, 2), (5, 3)])
def test_prove_mf_verify_mf_and_wire_round_trip_over_oracle(O, rnd, k, q):
    """prove_mf -> ProofMF -> bincode -> ProofMF -> verify_mf (channel/src/lib.rs:945-979, 1130-1240) over the oracle kit; tampered
    and malformed proofs are rejected"""
    from stark_mlwe_b200 import channel as ch, wire
    from stark_mlwe_b200.api import MerkleProof
    kit = OracleKit(O)
    table = rnd(4321 + k, 1 << k)
    vk = ch.build_vk_mf(kit, k, O.to_mont([0xABCD])[0], q)
    assert vk.tree_label == 0xABCD and vk.to_bincode()[:15] == (7).to_bytes(8, "little") + b"default"
    proof = ch.prove_mf(kit, vk, table)
    assert len(proof.rounds) == k and ch.verify_mf(kit, vk, proof)
    blob = wire.proof_mf_to_bincode(proof, O.from_mont_c)
    back = wire.proof_mf_from_bincode(blob, O.to_mont_c, MerkleProof.from_fields)
    assert wire.proof_mf_to_bincode(back, O.from_mont_c) == blob and ch.verify_mf(kit, vk, back)
    # layout: initial_root = u64 32 + canonical little-endian bytes, then the round count
    assert blob[:8] == (32).to_bytes(8, "little") and int.from_bytes(blob[8:40], "little") == O.from_mont(proof.initial_root.reshape(1, 4))[0]
    assert int.from_bytes(blob[40:48], "little") == k
    back.rounds[1].next_values[0, 0] ^= np.uint64(1)
    assert not ch.verify_mf(kit, vk, back)
    back = wire.proof_mf_from_bincode(blob, O.to_mont_c, MerkleProof.from_fields)
    back.rounds[0].c1 = back.rounds[0].c0
    assert not ch.verify_mf(kit, vk, back)
    assert not ch.verify_mf(kit, wire.VK(0xABCD, k, None), proof)                 # wrong VK variant
    for bad in (blob[:-1], blob + b"\0", blob[:40] + (1 << 40).to_bytes(8, "little") + blob[48:]):
        with pytest.raises(wire.WireError):
            wire.proof_mf_from_bincode(bad, O.to_mont_c, MerkleProof.from_fields)
