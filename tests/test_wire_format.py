"""CPU: bincode layout of MerkleProof (stark_mlwe_b200/wire.py) — round trip, size formula, canonical bytes, rejection of
malformed input — with the oracle's tree as the proof source and the oracle's conversions.  (GPU: tests/test_gpu_verify.py
feeds a deserialised proof to the device verifier.)"""
import struct

import numpy as np
import pytest


def test_merkle_proof_bincode_round_trip(O, rnd):
    from stark_mlwe_b200 import wire
    from stark_mlwe_b200.api import MerkleProof
    leaves = rnd(5150, 1000)
    tree = O.MerkleTree.new(leaves, 16, 77, O.Params.for_arity(16))
    pr = tree.open_many([3, 4, 500, 999])
    pr.arity = 16
    blob = wire.merkle_proof_to_bincode(pr, O.from_mont_c)
    assert len(blob) == wire.merkle_proof_size_bincode(pr)
    # header: 4 indices
    assert struct.unpack_from("<Q", blob, 0)[0] == 4 and struct.unpack_from("<4Q", blob, 8) == (3, 4, 500, 999)
    # first sibling of level 0 = canonical little-endian bytes of its value
    off = 8 + 32 + 8 + 8
    assert struct.unpack_from("<Q", blob, off)[0] == 32
    assert int.from_bytes(blob[off + 8:off + 40], "little") == O.from_mont(pr.siblings[0][:1])[0]
    back = wire.merkle_proof_from_bincode(blob, O.to_mont_c, MerkleProof.from_fields)
    assert back.indices == [3, 4, 500, 999] and back.arity == 16 and back.nlev == pr.nlev
    assert all(np.array_equal(a, b) for a, b in zip(back.siblings, pr.siblings))
    assert all(np.array_equal(a, b) for a, b in zip(back.group_sizes, pr.group_sizes))
    assert O.verify_many_ds(tree.root(), [3, 4, 500, 999], leaves[[3, 4, 500, 999]], back, 77, O.Params.for_arity(16), 16)
    with pytest.raises(wire.WireError):
        wire.merkle_proof_from_bincode(blob[:-1], O.to_mont_c, MerkleProof.from_fields)
    with pytest.raises(wire.WireError):
        wire.merkle_proof_from_bincode(blob + b"\0", O.to_mont_c, MerkleProof.from_fields)
    bad = bytearray(blob); bad[off + 8:off + 40] = b"\xff" * 32                    # >= p: ark-serialize Validate::Yes rejects it
    with pytest.raises(wire.WireError):
        wire.merkle_proof_from_bincode(bytes(bad), O.to_mont_c, MerkleProof.from_fields)
    bad = bytearray(blob); struct.pack_into("<Q", bad, off, 31)
    with pytest.raises(wire.WireError):
        wire.merkle_proof_from_bincode(bytes(bad), O.to_mont_c, MerkleProof.from_fields)
