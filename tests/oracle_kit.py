"""The kit interface of stark_mlwe_b200/channel.py over the CPU oracle (test infrastructure only)."""
import numpy as np


class OracleKit:
    """the kit interface of stark_mlwe_b200/channel.py over the CPU oracle (test infrastructure)"""

    def __init__(self, O):
        self.O = O
        self.cp = O.Params.commitment_default()

    def transcript(self, label):
        return self.O.Transcript(label)

    class _Tree:
        def __init__(self, t):
            self.t = t

        def open_many(self, idx):
            pr = self.t.open_many(idx)
            pr.arity = 16                              # MerkleProof.arity (merkle/src/lib.rs:142); the oracle's proof object leaves it to the caller
            return pr

    def commit(self, values, tree_label):
        tree = self.O.commitment_commit(values, tree_label, self.cp)
        return tree.root(), OracleKit._Tree(tree)

    def round_coeffs(self, values):
        return self.O.sumcheck_round_coeffs(values)

    def round_challenge(self, i, root):
        return self.O.mf_round_challenge_from_root(i, root)

    def fold(self, values, r):
        return self.O.mf_fold(values, r)

    def verify(self, root, indices, values, proof, tree_label):
        return self.O.verify_many_ds(root, indices, values, proof, tree_label, self.cp, 16)

    def to_canonical(self, x):
        return self.O.from_mont_c(np.asarray(x, dtype=np.uint64).reshape(1, 4))[0]

    def from_canonical(self, c):
        return self.O.to_mont_c(np.asarray(c, dtype=np.uint64).reshape(1, 4))[0]


def run_prover(ch, kit, table, q, tag=0xABCD):
    chan = ch.ProverChannel(kit.transcript(b"MF-TEST"))
    pr = ch.SumCheckMFProver(ch.SumCheckMFConfig(q), ch.MerkleChannelCfg(kit, tree_label=tag), chan, table)
    msgs = {"root0": pr.current_root().copy(), "claim": pr.send_claim(), "rounds": []}
    for i in range(pr.num_rounds()):
        msgs["rounds"].append(pr.round(i))
    msgs["final"] = pr.finalize_eval()
    msgs["tail"] = chan.challenge_scalar(b"tail")
    return msgs
