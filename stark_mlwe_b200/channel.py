"""Host-side mirror of the reference's channel crate for the Merkle-folded sum-check prover
(channel/src/lib.rs:13-60 ProverChannel, :120-145 MerkleChannelCfg, :536-760 SumCheckMFProver): same names, same call order, same
transcript traffic.  Every field operation, hash, commitment and opening is a call into the C-ABI library (`GpuKit`); what stays
here is the reference's own control flow — which bytes go into the transcript in which order, how query indices are drawn.
The logic is written against a small kit interface so that tests/ can run the identical control flow over the CPU oracle and
compare every emitted value; the product only ever constructs `GpuKit`."""
import numpy as np

from . import api

P_MOD = 0x40000000000000000000000000000000224698fc0994a8dd8c46eb2100000001


def _u64le(i):
    return int(i).to_bytes(8, "little")


class GpuKit:
    """the primitives of the control flow below, on the device"""

    def __init__(self, ctx, commit_params=None, transcript_params=None):
        self.ctx, self.commit_params, self.transcript_params = ctx, commit_params, transcript_params

    def transcript(self, label):
        return api.Transcript(self.ctx, label, self.transcript_params)

    def commit(self, values, tree_label):
        """MerkleCommitment::commit (commitment/src/lib.rs:85-90) -> (root, tree)"""
        cfg = api.MerkleConfig(self.ctx, tree_label, self.commit_params or api.commitment_default_params(self.ctx))
        root, aux = api.MerkleCommitment(cfg).commit(values)
        return root, aux.tree

    def round_coeffs(self, values):
        return api.sumcheck_round_coeffs(self.ctx, values)

    def round_challenge(self, i, root):
        return api.mf_round_challenge_from_root(self.ctx, i, root, self.transcript_params)

    def fold(self, values, r):
        return api.mf_fold(self.ctx, values, r)

    def verify(self, root, indices, values, proof, tree_label):
        """MerkleCommitment::verify (commitment/src/lib.rs:96-113) -> verify_many_ds on the device"""
        return api.verify_many_ds(root, indices, values, proof, tree_label, self.commit_params or api.commitment_default_params(self.ctx))

    def to_canonical(self, x):
        return self.ctx.to_canonical(np.asarray(x, dtype=np.uint64).reshape(1, 4))[0]

    def from_canonical(self, c):
        return self.ctx.from_canonical(np.asarray(c, dtype=np.uint64).reshape(1, 4))[0]


def _limbs_to_int(l):
    return int(l[0]) | (int(l[1]) << 64) | (int(l[2]) << 128) | (int(l[3]) << 192)


def _int_to_limbs(v):
    return np.array([(v >> (64 * k)) & 0xFFFFFFFFFFFFFFFF for k in range(4)], dtype=np.uint64)


class ProverChannel:
    """channel/src/lib.rs:13-60"""

    def __init__(self, transcript):
        self.tr = transcript

    def transcript_mut(self):
        return self.tr

    def send_digest(self, label, digest):
        self.tr.absorb_bytes(b"CHAN/SEND/DIGEST")
        self.tr.absorb_bytes(label)
        self.tr.absorb_field(digest)

    def challenge_scalar(self, label):
        return self.tr.challenge(label)

    def send_opening(self, indices, values, proof):
        tr = self.tr
        tr.absorb_bytes(b"CHAN/SEND/OPEN")
        for i in indices:
            tr.absorb_bytes(_u64le(i))
        for v in values:
            tr.absorb_field(v)
        tr.absorb_bytes(b"PROOF/ARITY")
        tr.absorb_bytes(_u64le(proof.arity))
        tr.absorb_bytes(b"PROOF/GROUP_SIZES")
        for lvl in proof.group_sizes:
            tr.absorb_bytes(_u64le(len(lvl)))
            for sz in lvl:
                tr.absorb_bytes(bytes([int(sz)]))
        tr.absorb_bytes(b"PROOF/SIBLINGS")
        for lvl in proof.siblings:
            tr.absorb_bytes(_u64le(len(lvl)))
            for s in lvl:
                tr.absorb_field(s)


class VerifierChannel(ProverChannel):
    """channel/src/lib.rs:62-116: the same absorptions under recv_* names"""
    recv_digest = ProverChannel.send_digest
    recv_opening = ProverChannel.send_opening


class MerkleChannelCfg:
    """channel/src/lib.rs:127-145: the Fr tag's low 64 bits are the tree label"""

    def __init__(self, kit, ds_tag_fr=None, tree_label=None):
        self.kit = kit
        self.tree_label = int(tree_label) if tree_label is not None else int(kit.to_canonical(ds_tag_fr)[0])


class MerkleProver:
    """channel/src/lib.rs:147-190: commit_vector / open_indices"""

    def __init__(self, chan, cfg):
        self.chan, self.cfg, self.root, self.tree = chan, cfg, None, None

    def commit_vector(self, leaves):
        self.root, self.tree = self.cfg.kit.commit(leaves, self.cfg.tree_label)
        self.chan.send_digest(b"commit/root", self.root)
        return self.root

    def open_indices(self, indices, table):
        values = np.asarray(table, dtype=np.uint64).reshape(-1, 4)[list(indices)]
        proof = self.tree.open_many(indices)
        self.chan.send_opening(indices, values, proof)
        return values, proof


class SumCheckMFConfig:
    def __init__(self, queries_per_round=2):
        self.queries_per_round = int(queries_per_round)


class MFFoldOpenings:
    """channel/src/lib.rs:582-589"""

    def __init__(self, cur_indices, cur_values, cur_proof, next_indices, next_values, next_proof):
        self.cur_indices, self.cur_values, self.cur_proof = cur_indices, cur_values, cur_proof
        self.next_indices, self.next_values, self.next_proof = next_indices, next_values, next_proof


class SumCheckMFProver:
    """channel/src/lib.rs:600-745.  `table` is the MLE evaluation table (2^k field elements)."""

    def __init__(self, cfg, merkle_cfg, chan, table):
        self.cfg, self.merkle_cfg, self.chan, self.kit = cfg, merkle_cfg, chan, merkle_cfg.kit
        self.values = np.ascontiguousarray(table, dtype=np.uint64).reshape(-1, 4)
        n = self.values.shape[0]
        assert n and n & (n - 1) == 0
        self.rounds = n.bit_length() - 1
        self.root, self.tree = self.kit.commit(self.values, merkle_cfg.tree_label)
        chan.send_digest(b"sumcheck-mf/root/0", self.root)

    def send_claim(self):
        # sum of the table = sum(a) + sum(b) = 2 c0 + c1 with the pair sums the device already computes
        c0, c1 = self.kit.round_coeffs(self.values) if self.values.shape[0] > 1 else (self.values[0], None)
        if c1 is None:
            s = self.values[0]
        else:
            v = (2 * _limbs_to_int(self.kit.to_canonical(c0)) + _limbs_to_int(self.kit.to_canonical(c1))) % P_MOD
            s = self.kit.from_canonical(_int_to_limbs(v))
        t = self.chan.transcript_mut()
        t.absorb_bytes(b"SUMCHECK/MF/CLAIM")
        t.absorb_field(s)
        return s

    def round(self, i):
        kit = self.kit
        c0, c1 = kit.round_coeffs(self.values)
        t = self.chan.transcript_mut()
        t.absorb_bytes(b"SUMCHECK/MF/ROUND")
        t.absorb_bytes(_u64le(i))
        t.absorb_bytes(b"COEFF/c0")
        t.absorb_field(c0)
        t.absorb_bytes(b"COEFF/c1")
        t.absorb_field(c1)
        r_i = kit.round_challenge(i, self.root)
        half = self.values.shape[0] // 2
        nxt = kit.fold(self.values, r_i)
        next_root, next_tree = kit.commit(nxt, self.merkle_cfg.tree_label)
        self.chan.send_digest(b"sumcheck-mf/root/next", next_root)

        # unique, sorted queries Q_i (channel/src/lib.rs:659-695)
        q_target = min(max(self.cfg.queries_per_round, 1), half)
        chosen, attempt, j = set(), 0, 0
        max_attempts = max(q_target * 16, 16)
        while len(chosen) < q_target and attempt < max_attempts:
            r = self.chan.challenge_scalar(b"sumcheck-mf/q" + _u64le(i) + _u64le(j))
            limbs = kit.to_canonical(r)
            acc = int(limbs[0]) ^ int(limbs[1]) ^ int(limbs[2]) ^ int(limbs[3])
            if half > 0:
                chosen.add(acc % half)
            j += 1
            attempt += 1
        if len(chosen) < q_target:
            for idx in range(half):
                chosen.add(idx)
                if len(chosen) == q_target:
                    break
        queries = sorted(chosen)

        cur_indices = [x for jj in queries for x in (2 * jj, 2 * jj + 1)]
        cur_values = self.values[cur_indices]
        cur_proof = self.tree.open_many(cur_indices)
        next_indices = list(queries)
        next_values = np.asarray(nxt, dtype=np.uint64).reshape(-1, 4)[next_indices]
        next_proof = next_tree.open_many(next_indices)
        self.chan.send_opening(cur_indices, cur_values, cur_proof)
        self.chan.send_opening(next_indices, next_values, next_proof)

        self.values, self.tree, self.root = np.asarray(nxt, dtype=np.uint64).reshape(-1, 4), next_tree, next_root
        return c0, c1, r_i, self.root, MFFoldOpenings(cur_indices, cur_values, cur_proof, next_indices, next_values, next_proof)

    def finalize_eval(self):
        assert self.values.shape[0] == 1
        val = self.values[0]
        t = self.chan.transcript_mut()
        t.absorb_bytes(b"SUMCHECK/MF/FINAL/EVAL")
        t.absorb_field(val)
        return val

    def current_root(self):
        return self.root

    def num_rounds(self):
        return self.rounds


class SumCheckMFVerifier:
    """channel/src/lib.rs:747-880.  Scalar checks (2 c0 + c1 == s_prev, s_next = c0 + c1 r) are exact integer arithmetic on the
    canonical values of a handful of scalars; the openings are verified on the device (kit.verify) and the fold relation of
    the opened pairs with the device fold."""

    def __init__(self, cfg, merkle_cfg, chan, initial_root, rounds):
        self.cfg, self.merkle_cfg, self.chan, self.kit = cfg, merkle_cfg, chan, merkle_cfg.kit
        self.cur_root, self.rounds = np.asarray(initial_root, dtype=np.uint64).reshape(4), int(rounds)

    def _int(self, x):
        return _limbs_to_int(self.kit.to_canonical(x))

    def _fr(self, v):
        return self.kit.from_canonical(_int_to_limbs(v % P_MOD))

    def receive_initial_root(self, root):
        self.chan.recv_digest(b"sumcheck-mf/root/0", root)
        self.cur_root = np.asarray(root, dtype=np.uint64).reshape(4)

    def recv_claim(self, s):
        t = self.chan.transcript_mut()
        t.absorb_bytes(b"SUMCHECK/MF/CLAIM")
        t.absorb_field(s)

    def start_round(self, i, s_prev, c0, c1):
        t = self.chan.transcript_mut()
        t.absorb_bytes(b"SUMCHECK/MF/ROUND")
        t.absorb_bytes(_u64le(i))
        t.absorb_bytes(b"COEFF/c0")
        t.absorb_field(c0)
        t.absorb_bytes(b"COEFF/c1")
        t.absorb_field(c1)
        if (2 * self._int(c0) + self._int(c1)) % P_MOD != self._int(s_prev):
            raise AssertionError("sum-check MF round consistency failed")      # assert_eq! at channel/src/lib.rs:797

    def derive_round_challenge(self, i):
        return self.kit.round_challenge(i, self.cur_root)

    def recv_next_root(self, next_root):
        self.chan.recv_digest(b"sumcheck-mf/root/next", next_root)
        self.cur_root = np.asarray(next_root, dtype=np.uint64).reshape(4)

    def compute_s_next(self, c0, c1, r_i):
        return self._fr(self._int(c0) + self._int(c1) * self._int(r_i))

    def verify_fold_openings(self, cur_indices, cur_values, cur_proof, next_indices, next_values, next_proof, r_i, prev_root, next_root):
        label = self.merkle_cfg.tree_label
        if not self.kit.verify(prev_root, cur_indices, cur_values, cur_proof, label):
            return False
        if not self.kit.verify(next_root, next_indices, next_values, next_proof, label):
            return False
        cur_values = np.asarray(cur_values, dtype=np.uint64).reshape(-1, 4)
        next_values = np.asarray(next_values, dtype=np.uint64).reshape(-1, 4)
        if len(cur_indices) != cur_values.shape[0] or len(next_indices) != next_values.shape[0]:
            return False
        pairs = {}
        for ix, val in zip(cur_indices, cur_values):
            pairs.setdefault(ix // 2, [None, None])[ix % 2] = val
        ab = []
        for j in next_indices:
            a, b = pairs.get(j, [None, None])
            if a is None or b is None:
                return False
            ab += [a, b]
        if not ab:
            return True
        folded = np.asarray(self.kit.fold(np.stack(ab), r_i), dtype=np.uint64).reshape(-1, 4)     # (1 - r) a + r b per pair
        return bool(np.array_equal(folded, next_values))

    def finalize_and_check(self, final_eval, s_k):
        t = self.chan.transcript_mut()
        t.absorb_bytes(b"SUMCHECK/MF/FINAL/EVAL")
        t.absorb_field(final_eval)
        if self._int(final_eval) != self._int(s_k):
            raise AssertionError("final MF sum-check evaluation mismatch")     # channel/src/lib.rs:868

    def current_root(self):
        return self.cur_root


# ------------------------------------------------------------------------------------------------ prove_mf / verify_mf
def build_vk_mf(kit, k, ds_tag_fr, queries_per_round):
    """channel/src/lib.rs:1034-1041: tree_label = low 64 bits of the canonical tag (fr_tag_to_u64, :120-124)"""
    from . import wire
    return wire.VK(int(kit.to_canonical(ds_tag_fr)[0]), k, queries_per_round)


def prove_mf(kit, vk, witness):
    """channel/src/lib.rs:1130-1174: transcript "E2E/MF", SumCheckMFProver over the witness table, one RoundMF per variable"""
    from . import wire
    assert vk.queries_per_round is not None, "wrong VK variant"
    chan = ProverChannel(kit.transcript(b"E2E/MF"))
    cfg = MerkleChannelCfg(kit, tree_label=vk.tree_label)
    sp = SumCheckMFProver(SumCheckMFConfig(vk.queries_per_round), cfg, chan, witness)
    initial_root = np.array(sp.current_root(), dtype=np.uint64)
    sp.send_claim()
    rounds = []
    for i in range(vk.k):
        c0, c1, _r, next_root, op = sp.round(i)
        rounds.append(wire.RoundMF(c0, c1, np.array(next_root, dtype=np.uint64), op.cur_indices, op.cur_values, op.cur_proof,
                                   op.next_indices, op.next_values, op.next_proof))
    return wire.ProofMF(initial_root, rounds, sp.finalize_eval())


def verify_mf(kit, vk, proof):
    """channel/src/lib.rs:1176-1240.  Like the reference, the claim is not carried by ProofMF: round 0 starts from 2 c0 + c1 and
    the final check is finalize_and_check(final_eval, running) (which asserts inside the verifier; a failed check returns False here)."""
    if vk.queries_per_round is None:
        return False
    chan = VerifierChannel(kit.transcript(b"E2E/MF"))
    cfg = MerkleChannelCfg(kit, tree_label=vk.tree_label)
    sv = SumCheckMFVerifier(SumCheckMFConfig(vk.queries_per_round), cfg, chan, proof.initial_root, vk.k)
    sv.receive_initial_root(proof.initial_root)
    running, prev_root = None, proof.initial_root
    for i, r in enumerate(proof.rounds):
        if running is None:
            v = (2 * _limbs_to_int(kit.to_canonical(r.c0)) + _limbs_to_int(kit.to_canonical(r.c1))) % P_MOD
            running = kit.from_canonical(_int_to_limbs(v))
        try:
            sv.start_round(i, running, r.c0, r.c1)
        except AssertionError:
            return False
        r_i = sv.derive_round_challenge(i)
        sv.recv_next_root(r.next_root)
        if not sv.verify_fold_openings(r.cur_indices, r.cur_values, r.cur_proof, r.next_indices, r.next_values, r.next_proof, r_i, prev_root, r.next_root):
            return False
        running = sv.compute_s_next(r.c0, r.c1, r_i)
        prev_root = r.next_root
    try:
        sv.finalize_and_check(proof.final_eval, running if running is not None else proof.final_eval)
    except AssertionError:
        return False
    return True
