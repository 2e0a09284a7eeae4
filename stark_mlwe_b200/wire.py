"""Wire format of `MerkleProof` (merkle/src/lib.rs:131-143) as the reference's serde derive + bincode 1.3 default options lay
it out (fixed-width little-endian integers, u64 length prefixes, `usize` as u64):

    indices      u64 n, n x u64
    siblings     u64 levels, per level: u64 count, per sibling: u64 32, 32 bytes   (SerFr -> serialize_bytes of the canonical
                                                                                    compressed field element, merkle/src/lib.rs:24-32)
    group_sizes  u64 levels, per level: u64 count, count x u8
    arity        u64

Field elements cross the ABI in Montgomery form; the canonical bytes come from the device conversion (`Context.to_canonical`,
field/src/lib.rs:206-215).  No reference build is available to pin these bytes (no cargo here): the layout follows the
published bincode 1.3 / serde_bytes rules and is checked by round trips and by size (`tests/`)."""
import struct

import numpy as np


def merkle_proof_to_bincode(proof, to_canonical):
    """to_canonical: (n, 4) Montgomery uint64 -> (n, 4) canonical uint64 (e.g. Context.to_canonical)"""
    out = [struct.pack("<Q", len(proof.indices))] + [struct.pack("<Q", int(i)) for i in proof.indices]
    out.append(struct.pack("<Q", len(proof.siblings)))
    for lvl in proof.siblings:
        lvl = np.ascontiguousarray(lvl, dtype=np.uint64).reshape(-1, 4)
        out.append(struct.pack("<Q", lvl.shape[0]))
        if lvl.shape[0]:
            canon = np.ascontiguousarray(to_canonical(lvl), dtype="<u8").reshape(-1, 4)
            for row in canon:
                out.append(struct.pack("<Q", 32))
                out.append(row.tobytes())
    out.append(struct.pack("<Q", len(proof.group_sizes)))
    for g in proof.group_sizes:
        g = np.ascontiguousarray(g, dtype=np.uint8).reshape(-1)
        out.append(struct.pack("<Q", g.shape[0]))
        out.append(g.tobytes())
    out.append(struct.pack("<Q", int(proof.arity)))
    return b"".join(out)


def merkle_proof_size_bincode(proof):
    n_sib = sum(int(np.asarray(x).reshape(-1, 4).shape[0]) for x in proof.siblings)
    n_grp = sum(int(np.asarray(g).reshape(-1).shape[0]) for g in proof.group_sizes)
    return 8 + 8 * len(proof.indices) + 8 + 8 * len(proof.siblings) + 40 * n_sib + 8 + 8 * len(proof.group_sizes) + n_grp + 8


class WireError(ValueError):
    pass


def merkle_proof_from_bincode(data, from_canonical, make_proof):
    """make_proof(indices, siblings, group_sizes, arity) -> proof object (e.g. api.MerkleProof.from_fields).  Rejects truncated
    input, trailing bytes, sibling blobs that are not 32 bytes and non-canonical field encodings (ark-serialize Validate::Yes)."""
    P = 0x40000000000000000000000000000000224698fc0994a8dd8c46eb2100000001
    pos = 0

    def u64():
        nonlocal pos
        if pos + 8 > len(data):
            raise WireError("truncated")
        v = struct.unpack_from("<Q", data, pos)[0]
        pos += 8
        return v

    def take(n):
        nonlocal pos
        if pos + n > len(data):
            raise WireError("truncated")
        b = data[pos:pos + n]
        pos += n
        return b

    n = u64()
    if n > len(data):
        raise WireError("length prefix exceeds input")
    indices = [u64() for _ in range(n)]
    siblings = []
    for _ in range(u64()):
        cnt = u64()
        if cnt * 40 > len(data):
            raise WireError("length prefix exceeds input")
        rows = np.zeros((cnt, 4), dtype=np.uint64)
        for k in range(cnt):
            if u64() != 32:
                raise WireError("field element is not 32 bytes")
            b = take(32)
            if int.from_bytes(b, "little") >= P:
                raise WireError("non-canonical field element")
            rows[k] = np.frombuffer(b, dtype="<u8")
        siblings.append(np.asarray(from_canonical(rows), dtype=np.uint64).reshape(-1, 4) if cnt else rows)
    group_sizes = []
    for _ in range(u64()):
        cnt = u64()
        group_sizes.append(np.frombuffer(take(cnt), dtype=np.uint8).copy())
    arity = u64()
    if pos != len(data):
        raise WireError("trailing bytes")
    return make_proof(indices, siblings, group_sizes, arity)


# ---------------------------------------------------------------------------------------------- channel crate proof structs
# channel/src/lib.rs:894-979: FBytes (#[serde(with = "serde_bytes")] Vec<u8> holding the 32 canonical compressed bytes),
# MerkleProofBytes { arity, group_sizes, indices, siblings }, RoundMF, ProofMF, VK — serde derive + bincode 1.3 defaults
# (struct fields in declaration order, u64 length prefixes, usize as u64, enum variant index as u32).
def _fbytes(rows_canon):
    return b"".join(struct.pack("<Q", 32) + np.ascontiguousarray(r, dtype="<u8").tobytes() for r in np.asarray(rows_canon, dtype=np.uint64).reshape(-1, 4))


def merkle_proof_bytes_to_bincode(proof, to_canonical):
    """MerkleProofBytes (channel/src/lib.rs:965-971): arity, group_sizes, indices, siblings — note the field order differs from MerkleProof"""
    out = [struct.pack("<Q", int(proof.arity)), struct.pack("<Q", len(proof.group_sizes))]
    for g in proof.group_sizes:
        g = np.ascontiguousarray(g, dtype=np.uint8).reshape(-1)
        out += [struct.pack("<Q", g.shape[0]), g.tobytes()]
    out += [struct.pack("<Q", len(proof.indices))] + [struct.pack("<Q", int(i)) for i in proof.indices]
    out.append(struct.pack("<Q", len(proof.siblings)))
    for lvl in proof.siblings:
        lvl = np.ascontiguousarray(lvl, dtype=np.uint64).reshape(-1, 4)
        out.append(struct.pack("<Q", lvl.shape[0]))
        if lvl.shape[0]:
            out.append(_fbytes(to_canonical(lvl)))
    return b"".join(out)


class _Reader:
    P = 0x40000000000000000000000000000000224698fc0994a8dd8c46eb2100000001

    def __init__(self, data):
        self.d, self.pos = data, 0

    def u64(self):
        if self.pos + 8 > len(self.d):
            raise WireError("truncated")
        v = struct.unpack_from("<Q", self.d, self.pos)[0]
        self.pos += 8
        return v

    def count(self, unit=1):
        n = self.u64()
        if n * unit > len(self.d) - self.pos:
            raise WireError("length prefix exceeds input")
        return n

    def take(self, n):
        if self.pos + n > len(self.d):
            raise WireError("truncated")
        b = self.d[self.pos:self.pos + n]
        self.pos += n
        return b

    def fbytes(self):
        if self.u64() != 32:
            raise WireError("field element is not 32 bytes")
        b = self.take(32)
        if int.from_bytes(b, "little") >= self.P:
            raise WireError("non-canonical field element")          # F::deserialize_compressed validates (channel/src/lib.rs:913-918)
        return np.frombuffer(b, dtype="<u8").astype(np.uint64)

    def fvec(self, from_canonical):
        n = self.count(40)
        rows = np.zeros((n, 4), dtype=np.uint64)
        for k in range(n):
            rows[k] = self.fbytes()
        return np.asarray(from_canonical(rows), dtype=np.uint64).reshape(-1, 4) if n else rows

    def merkle_proof_bytes(self, from_canonical, make_proof):
        arity = self.u64()
        group_sizes = [np.frombuffer(self.take(self.count()), dtype=np.uint8).copy() for _ in range(self.count(8))]
        indices = [self.u64() for _ in range(self.count(8))]
        siblings = [self.fvec(from_canonical) for _ in range(self.count(8))]
        return make_proof(indices, siblings, group_sizes, arity)


class RoundMF:
    """channel/src/lib.rs:952-962"""

    def __init__(self, c0, c1, next_root, cur_indices, cur_values, cur_proof, next_indices, next_values, next_proof):
        self.c0, self.c1, self.next_root = c0, c1, next_root
        self.cur_indices, self.cur_values, self.cur_proof = list(cur_indices), cur_values, cur_proof
        self.next_indices, self.next_values, self.next_proof = list(next_indices), next_values, next_proof


class ProofMF:
    """channel/src/lib.rs:945-949"""

    def __init__(self, initial_root, rounds, final_eval):
        self.initial_root, self.rounds, self.final_eval = initial_root, list(rounds), final_eval


def proof_mf_to_bincode(proof, to_canonical):
    def fb(x):
        return _fbytes(to_canonical(np.asarray(x, dtype=np.uint64).reshape(1, 4)))

    def fv(xs):
        xs = np.asarray(xs, dtype=np.uint64).reshape(-1, 4)
        return struct.pack("<Q", xs.shape[0]) + (_fbytes(to_canonical(xs)) if xs.shape[0] else b"")

    def iv(idx):
        return struct.pack("<Q", len(idx)) + b"".join(struct.pack("<Q", int(i)) for i in idx)
    out = [fb(proof.initial_root), struct.pack("<Q", len(proof.rounds))]
    for r in proof.rounds:
        out += [fb(r.c0), fb(r.c1), fb(r.next_root), iv(r.cur_indices), fv(r.cur_values), merkle_proof_bytes_to_bincode(r.cur_proof, to_canonical),
                iv(r.next_indices), fv(r.next_values), merkle_proof_bytes_to_bincode(r.next_proof, to_canonical)]
    out.append(fb(proof.final_eval))
    return b"".join(out)


def proof_mf_from_bincode(data, from_canonical, make_proof):
    rd = _Reader(data)

    def f1():
        return np.asarray(from_canonical(rd.fbytes().reshape(1, 4)), dtype=np.uint64).reshape(4)
    initial_root = f1()
    rounds = []
    for _ in range(rd.count(8)):
        c0, c1, next_root = f1(), f1(), f1()
        cur_indices = [rd.u64() for _ in range(rd.count(8))]
        cur_values = rd.fvec(from_canonical)
        cur_proof = rd.merkle_proof_bytes(from_canonical, make_proof)
        next_indices = [rd.u64() for _ in range(rd.count(8))]
        next_values = rd.fvec(from_canonical)
        next_proof = rd.merkle_proof_bytes(from_canonical, make_proof)
        rounds.append(RoundMF(c0, c1, next_root, cur_indices, cur_values, cur_proof, next_indices, next_values, next_proof))
    final_eval = f1()
    if rd.pos != len(data):
        raise WireError("trailing bytes")
    return ProofMF(initial_root, rounds, final_eval)


class VK:
    """channel/src/lib.rs:925-937: { poseidon_seed: String, tree_label: u64, k: usize, variant: VKVariant }"""

    def __init__(self, tree_label, k, queries_per_round=None, poseidon_seed="default"):
        self.poseidon_seed, self.tree_label, self.k, self.queries_per_round = poseidon_seed, int(tree_label), int(k), queries_per_round

    def to_bincode(self):
        s = self.poseidon_seed.encode()
        out = struct.pack("<Q", len(s)) + s + struct.pack("<QQ", self.tree_label, self.k)
        if self.queries_per_round is None:
            return out + struct.pack("<I", 0)                                     # VKVariant::SumCheckPlain
        return out + struct.pack("<IQ", 1, int(self.queries_per_round))           # VKVariant::SumCheckMF { queries_per_round }
