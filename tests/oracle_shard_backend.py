"""CPU stand-in for one rank's device work in the range-sharded FRI commit phase (test infrastructure only: it calls the
oracle).  Restates, over oracle primitives, what smw_fri_shard_build / smw_merkle_root_from_level / smw_fri_build_transcript_from
do on the GPU: local folds (deep_ali/src/fri.rs:85-102), leaf digests (fri.rs:38-44, merkle/src/lib.rs:382-400) and tree levels
with global DS positions (merkle/src/lib.rs:164-181)."""
import numpy as np

from oracle import oracle as O

U32MAX = 0xFFFFFFFF


def _node(params, arity, level, pos, label, children):
    return O.hash_with_ds_dynamic(params, O.to_mont([arity, level, pos, label]), children)


def _levels_up(params, arity, label, nodes, level, pos0, whole_only):
    """hash levels upward from `nodes` (DS level of their parents = `level`, global position of the first parent = pos0 / arity)"""
    nodes = np.asarray(nodes, dtype=np.uint64).reshape(-1, 4)
    while len(nodes) > 1 or (whole_only and len(nodes) >= arity):
        if whole_only and (arity < 2 or len(nodes) < arity or len(nodes) % arity):
            break
        if arity < 2:
            break
        nxt = [_node(params, arity, level, pos0 // arity + q, label, nodes[q * arity:(q + 1) * arity]) for q in range((len(nodes) + arity - 1) // arity)]
        nodes, level, pos0 = np.stack(nxt), level + 1, pos0 // arity
    return nodes, level


def _leaf_digests(f, s, m, n_global, label, first):
    arity = O.pick_arity_for_layer(n_global, m)
    params = O.Params.for_arity(arity)
    cp = np.zeros_like(f) if s is None else np.repeat(np.asarray(s).reshape(-1, 4), m, axis=0)[: len(f)]
    if arity >= 8:
        return O.hash_leaf_pair(f, cp), arity, params
    dig = [O.hash_with_ds_dynamic(params, O.to_mont([arity, U32MAX, first + i, label]), np.stack([f[i], cp[i]])) for i in range(len(f))]
    return np.stack(dig), arity, params


class OracleShard:
    def __init__(self, f0_slice, n0, first_index, schedule, seed_z):
        G = n0 // len(f0_slice)
        n = [n0]
        for m in schedule:
            n.append(n[-1] // m)
        ms = list(schedule) + [1]
        L, K = len(schedule), 0
        while K <= L and n[K] % G == 0 and (K == L or (n[K] // G) % ms[K] == 0):
            K += 1
        f = [np.asarray(f0_slice, dtype=np.uint64).reshape(-1, 4)]
        for l in range(min(K, L)):
            f.append(O.fri_fold_layer(f[l], O.fri_sample_z_ell(seed_z, l, n[l]), schedule[l]))
        self.num_local_layers, self._caps = K, []
        for l in range(K):
            first = first_index // (n0 // n[l])
            dig, arity, params = _leaf_digests(f[l], f[l + 1] if l < L else None, ms[l], n[l], l, first)
            nodes, level = _levels_up(params, arity, l, dig, 0, first, True)
            self._caps.append((nodes, level, arity))
        self._tail = f[K] if K <= L else np.zeros((0, 4), np.uint64)

    def cap(self, l):
        return self._caps[l]

    def tail(self):
        return self._tail


class OracleFriBackend:
    def shard(self, f0_slice, n0, first_index, schedule, seed_z):
        return OracleShard(f0_slice, n0, first_index, schedule, seed_z)

    def root_from_level(self, nodes, level, arity, label):
        top, _ = _levels_up(O.Params.for_arity(arity), arity, label, nodes, level, 0, False)
        return top[0]

    def build_from(self, f, schedule, seed_z, level0):
        f = [np.asarray(f, dtype=np.uint64).reshape(-1, 4)]
        for l, m in enumerate(schedule):
            f.append(O.fri_fold_layer(f[l], O.fri_sample_z_ell(seed_z, level0 + l, len(f[l])), m))
        ms, roots = list(schedule) + [1], []
        for l in range(len(f)):
            dig, arity, params = _leaf_digests(f[l], f[l + 1] if l + 1 < len(f) else None, ms[l], len(f[l]), level0 + l, 0)
            top, _ = _levels_up(params, arity, level0 + l, dig, 0, 0, False)
            roots.append(top[0])
        return roots


class OracleNttBackend:
    """CPU stand-in for the device steps of the distributed NTT (shard.ntt_cyclic_to_chunks): oracle NTT and field ops"""

    def root_of_unity(self, log_n):
        return O.root_of_unity(log_n)

    def ntt(self, a):
        return O.ntt(a)

    def scale_powers(self, a, g):
        a = np.asarray(a, dtype=np.uint64).reshape(-1, 4)
        out, p, g = a.copy(), O.root_of_unity(0).reshape(1, 4), np.asarray(g, dtype=np.uint64).reshape(1, 4)
        for i in range(len(a)):
            out[i] = O.fr_mul(a[i:i + 1], p)[0]
            p = O.fr_mul(p, g)
        return out

    def dft_strided(self, a, log_g, stride, count):
        G = 1 << log_g
        a = np.asarray(a, dtype=np.uint64).reshape(-1, 4)
        out = a.copy()
        wp = self.scale_powers(np.tile(O.root_of_unity(0).reshape(1, 4), (G, 1)), O.root_of_unity(log_g))
        for k in range(G):
            acc = np.zeros((count, 4), np.uint64)
            for q in range(G):
                acc = O.fr_add(acc, O.fr_mul(a[q * stride:q * stride + count], np.tile(wp[(q * k) % G].reshape(1, 4), (count, 1))))
            out[k * stride:k * stride + count] = acc
        return out


class OracleMfBackend:
    """one rank's work in the range-sharded Merkle-folded sum-check (shard.mf_commit_phase_sharded_host): arity-16 trees with the commitment
    default parameters and tree_label = ds_tag (commitment/src/lib.rs:85-90), r_i from (i, root_i), the pairwise fold"""

    def __init__(self):
        self.params = O.Params.commitment_default()

    def cap(self, cur, ds_tag, first):
        return _levels_up(self.params, 16, ds_tag, cur, 0, first, True)

    def root_from_level(self, nodes, level, ds_tag):
        top, _ = _levels_up(self.params, 16, ds_tag, nodes, level, 0, False)
        return top[0]

    def challenge(self, i, root):
        return O.mf_round_challenge_from_root(i, root)

    def fold(self, cur, r):
        return O.mf_fold(cur, r)
