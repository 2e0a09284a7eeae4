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
