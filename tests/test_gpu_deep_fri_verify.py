"""GPU: deep_fri_prove -> DeepFriProof -> deep_fri_verify round trip (deep_ali/src/fri.rs:604-762) and the negatives the
reference's verifier is built to catch: a changed payload value, a broken fold relation, a wrong root, a truncated proof.
The Merkle decisions underneath are checked against the oracle in tests/test_gpu_verify.py."""
import copy
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _bench_inputs(O, k, seed=1337):
    n0 = 1 << k
    g = O.StdRng.seed_from_u64(seed)
    return [g.fr_rand(n0) for _ in range(4)]


@pytest.mark.parametrize("schedule,k,r", [([16, 16, 8], 11, 32), ([2, 2, 2, 2], 6, 8), ([4, 8], 7, 5), ([16], 8, 4)])
def test_deep_fri_prove_verify_round_trip(ctx, smw, O, schedule, k, r):
    a, s, e, t = _bench_inputs(O, k)
    params = smw.DeepFriParams(schedule, r, 0xDEEFBAAD)
    proof = smw.deep_fri_prove_proof(ctx, a, s, e, t, params)
    assert len(proof.roots) == len(schedule) + 1 and len(proof.queries) == r and proof.n0 == 1 << k
    assert smw.deep_fri_verify(ctx, params, proof)

    # structure (fri.rs:645-647)
    assert not smw.deep_fri_verify(ctx, smw.DeepFriParams(schedule, r + 1, 0xDEEFBAAD), proof)
    assert not smw.deep_fri_verify(ctx, smw.DeepFriParams(schedule + [2], r, 0xDEEFBAAD), proof)

    # a changed opened value: the child multiproof of layer 0 no longer reaches the root
    bad = copy.deepcopy(proof)
    bad.queries[0].per_layer_payloads[0][0, 0] ^= np.uint64(1)
    assert not smw.deep_fri_verify(ctx, params, bad)

    # s_i != f_parent[b] while both Merkle openings stay valid is impossible for an honest state; break the relation on the
    # payload of a query whose (i, b) another query also opened first (or_insert keeps the first), else on query r-1
    bad = copy.deepcopy(proof)
    bad.queries[-1].per_layer_payloads[-1][2, 1] ^= np.uint64(4)
    assert not smw.deep_fri_verify(ctx, params, bad)

    # wrong root of a middle layer
    bad = copy.deepcopy(proof)
    bad.roots = bad.roots.copy(); bad.roots[1, 0] ^= np.uint64(1)
    assert not smw.deep_fri_verify(ctx, params, bad)

    # a sibling removed from a parent multiproof
    bad = copy.deepcopy(proof)
    pp = bad.layers[0].parent_proof
    lv = next((i for i, x in enumerate(pp.siblings) if x.shape[0]), None)
    if lv is not None:
        pp.siblings[lv] = pp.siblings[lv][:-1]
        assert not smw.deep_fri_verify(ctx, params, bad)

    # final layer: index must be 0 and the pair must open
    bad = copy.deepcopy(proof)
    bad.queries[0].final_index = 1
    assert not smw.deep_fri_verify(ctx, params, bad)
    bad = copy.deepcopy(proof)
    f, s_ = bad.queries[0].final_pair
    f = f.copy(); f[0] ^= np.uint64(1)
    bad.queries[0].final_pair = (f, s_)
    assert not smw.deep_fri_verify(ctx, params, bad)


def test_deep_fri_verify_rejects_proof_of_other_statement(ctx, smw, O):
    """a proof for inputs A does not verify once its roots are replaced by those of inputs B"""
    params = smw.DeepFriParams([16, 16, 8], 8, 0xDEEFBAAD)
    pa = smw.deep_fri_prove_proof(ctx, *_bench_inputs(O, 11, 1), params)
    pb = smw.deep_fri_prove_proof(ctx, *_bench_inputs(O, 11, 2), params)
    assert smw.deep_fri_verify(ctx, params, pa) and smw.deep_fri_verify(ctx, params, pb)
    mixed = copy.deepcopy(pa)
    mixed.roots = pb.roots
    assert not smw.deep_fri_verify(ctx, params, mixed)
