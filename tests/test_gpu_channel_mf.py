"""GPU: the Merkle-folded sum-check prover / verifier of the channel crate (channel/src/lib.rs:13-60, 536-880) through
stark_mlwe_b200/channel.py.  The same control flow runs once over the device kit and once over the CPU oracle; every value the
prover emits (claim, c0, c1, r_i, roots, query sets, opened values, multiproofs, final evaluation) and the transcripts' final
challenge must agree bit for bit.  Then the device verifier replays the device prover's messages (the protocol loop of
channel/src/lib.rs:1370-1451) and rejects tampered openings."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def eq(a, b):
    return np.array_equal(np.asarray(a, dtype=np.uint64).reshape(-1), np.asarray(b, dtype=np.uint64).reshape(-1))


from oracle_kit import OracleKit, run_prover


@pytest.mark.parametrize("k,q", [(5, 2), (7, 3), (4, 40), (1, 1)])
def test_mf_prover_matches_oracle_and_verifies(ctx, O, rnd, k, q):
    from stark_mlwe_b200 import channel as ch
    table = rnd(900 + k, 1 << k)
    gk = ch.GpuKit(ctx)
    g, o = run_prover(ch, gk, table, q), run_prover(ch, OracleKit(O), table, q)
    assert eq(g["root0"], o["root0"]) and eq(g["claim"], o["claim"]) and eq(g["final"], o["final"]) and eq(g["tail"], o["tail"])
    assert len(g["rounds"]) == k
    for (gc0, gc1, gr, groot, gop), (oc0, oc1, orr, oroot, oop) in zip(g["rounds"], o["rounds"]):
        assert eq(gc0, oc0) and eq(gc1, oc1) and eq(gr, orr) and eq(groot, oroot)
        assert gop.cur_indices == oop.cur_indices and gop.next_indices == oop.next_indices
        assert eq(gop.cur_values, oop.cur_values) and eq(gop.next_values, oop.next_values)
        for a, b in ((gop.cur_proof, oop.cur_proof), (gop.next_proof, oop.next_proof)):
            assert a.indices == b.indices and a.arity == b.arity == 16 and len(a.siblings) == len(b.siblings)
            assert all(eq(x, y) for x, y in zip(a.siblings, b.siblings)) and all(eq(x, y) for x, y in zip(a.group_sizes, b.group_sizes))

    # the verifier's side of the protocol on the device
    vchan = ch.VerifierChannel(gk.transcript(b"MF-TEST"))
    ver = ch.SumCheckMFVerifier(ch.SumCheckMFConfig(q), ch.MerkleChannelCfg(gk, tree_label=0xABCD), vchan, g["root0"], k)
    ver.receive_initial_root(g["root0"])
    ver.recv_claim(g["claim"])
    s_prev = g["claim"]
    for i, (c0, c1, r_i, next_root, op) in enumerate(g["rounds"]):
        ver.start_round(i, s_prev, c0, c1)
        prev_root = ver.current_root().copy()
        assert eq(ver.derive_round_challenge(i), r_i)
        ver.recv_next_root(next_root)
        # the prover draws its queries between the digest and the openings: replay the same draws on the verifier's transcript
        half = 1 << (k - i - 1)
        q_target, seen, j = min(max(q, 1), half), set(), 0
        while len(seen) < q_target and j < max(q_target * 16, 16):
            limbs = gk.to_canonical(vchan.challenge_scalar(b"sumcheck-mf/q" + int(i).to_bytes(8, "little") + int(j).to_bytes(8, "little")))
            seen.add((int(limbs[0]) ^ int(limbs[1]) ^ int(limbs[2]) ^ int(limbs[3])) % half)
            j += 1
        assert sorted(seen) == op.next_indices or len(seen) < q_target
        assert ver.verify_fold_openings(op.cur_indices, op.cur_values, op.cur_proof, op.next_indices, op.next_values, op.next_proof, r_i, prev_root, next_root)
        bad = op.next_values.copy(); bad[0, 0] ^= np.uint64(1)
        assert not ver.verify_fold_openings(op.cur_indices, op.cur_values, op.cur_proof, op.next_indices, bad, op.next_proof, r_i, prev_root, next_root)
        bad = op.cur_values.copy(); bad[-1, 2] ^= np.uint64(1)
        assert not ver.verify_fold_openings(op.cur_indices, bad, op.cur_proof, op.next_indices, op.next_values, op.next_proof, r_i, prev_root, next_root)
        vchan.recv_opening(op.cur_indices, op.cur_values, op.cur_proof)
        vchan.recv_opening(op.next_indices, op.next_values, op.next_proof)
        s_prev = ver.compute_s_next(c0, c1, r_i)
    ver.finalize_and_check(g["final"], s_prev)
    assert eq(vchan.challenge_scalar(b"tail"), g["tail"])          # both transcripts absorbed the same conversation
    wrong = g["claim"].copy(); wrong[0] ^= np.uint64(1)
    with pytest.raises(AssertionError):                              # 2 c0 + c1 != s_prev (channel/src/lib.rs:797)
        ver.start_round(0, wrong, g["rounds"][0][0], g["rounds"][0][1])


def test_merkle_prover_commit_vector_and_open(ctx, O, rnd):
    """MerkleProver::commit_vector / open_indices (channel/src/lib.rs:172-190) against the oracle kit"""
    from stark_mlwe_b200 import channel as ch
    table = rnd(77, 300)
    outs = []
    for kit in (ch.GpuKit(ctx), OracleKit(O)):
        chan = ch.ProverChannel(kit.transcript(b"MP"))
        mp = ch.MerkleProver(chan, ch.MerkleChannelCfg(kit, ds_tag_fr=table[0]))
        root = mp.commit_vector(table)
        vals, proof = mp.open_indices([3, 17, 299], table)
        outs.append((root, vals, proof, chan.challenge_scalar(b"x"), mp.cfg.tree_label))
    (r0, v0, p0, c0, l0), (r1, v1, p1, c1, l1) = outs
    assert l0 == l1 and eq(r0, r1) and eq(v0, v1) and eq(c0, c1) and p0.indices == p1.indices
    assert all(eq(x, y) for x, y in zip(p0.siblings, p1.siblings))


@pytest.mark.parametrize("k,q", [(6, 3), (3, 8)])
def test_prove_mf_on_device_serialises_to_the_same_bytes_as_the_oracle(ctx, O, rnd, k, q):
    """prove_mf / verify_mf (channel/src/lib.rs:1130-1240) on the device; the ProofMF bincode bytes equal those of the same
    protocol run over the CPU oracle, and a deserialised proof verifies on the device"""
    from stark_mlwe_b200 import channel as ch, wire
    from stark_mlwe_b200.api import MerkleProof
    table = rnd(4400 + k, 1 << k)
    gk, ok = ch.GpuKit(ctx), OracleKit(O)
    tag = O.to_mont([2025])[0]
    vk = ch.build_vk_mf(gk, k, tag, q)
    pg, po = ch.prove_mf(gk, vk, table), ch.prove_mf(ok, ch.build_vk_mf(ok, k, tag, q), table)
    bg, bo = wire.proof_mf_to_bincode(pg, ctx.to_canonical), wire.proof_mf_to_bincode(po, O.from_mont_c)
    assert bg == bo
    back = wire.proof_mf_from_bincode(bg, ctx.from_canonical, MerkleProof.from_fields)
    assert ch.verify_mf(gk, vk, back) and ch.verify_mf(ok, vk, back)
    back.rounds[-1].cur_values[0, 1] ^= np.uint64(1)
    assert not ch.verify_mf(gk, vk, back)
