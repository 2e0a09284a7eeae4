"""GPU parity of the range-sharded FRI commit phase (csrc/capi.cu smw_fri_shard_build,
smw_merkle_root_from_level, smw_fri_build_transcript_from; host logic stark_mlwe_b200/shard.py): one GPU plays every rank in
turn, the all-gather is a concatenation, and the layer roots must equal (a) the CPU oracle's fri_build_transcript
(deep_ali/src/fri.rs:231-312) bit for bit at oracle sizes and (b) goldens of the pinned oracle at 2^18."""
import numpy as np
import pytest

from stark_mlwe_b200 import shard

pytestmark = pytest.mark.gpu
SEED_Z = 0xDEEFBAAD


def sharded_roots(ctx, f0, schedule, world):
    n0 = len(f0)
    k = n0 // world
    be = shard.GpuFriBackend(ctx)
    shs = [be.shard(f0[r * k:(r + 1) * k], n0, r * k, schedule, SEED_Z) for r in range(world)]
    K = shs[0].num_local_layers
    assert all(s.num_local_layers == K for s in shs)

    class Rank(shard.GpuFriBackend):
        def __init__(self, r):
            super().__init__(ctx)
            self.r = r

        def shard(self, *a):
            return shs[self.r]

    out = []
    for r in range(world):
        calls = [0]

        def gather(_mine):
            i = calls[0]
            calls[0] += 1
            parts = [shs[q].cap(i)[0] for q in range(world)] if i < K else [shs[q].tail() for q in range(world)]
            return np.concatenate(parts)
        out.append(shard.fri_roots_sharded(Rank(r), f0[r * k:(r + 1) * k], n0, r, world, schedule, SEED_Z, gather=gather))
    for r in range(1, world):
        assert np.array_equal(out[0], out[r]), "ranks disagree"
    return out[0], K


@pytest.mark.parametrize("n0,schedule,world", [
    (1 << 10, [2] * 10, 2), (1 << 10, [2] * 10, 8), (1 << 12, [16, 16, 8], 4), (1 << 12, [16, 16, 8], 8),
    (1 << 11, [4, 8, 2], 2), (1 << 9, [128], 4), (1 << 6, [2, 2], 1)])
def test_sharded_roots_match_oracle(ctx, O, rnd, n0, schedule, world):
    f0 = rnd(n0 + world, n0)
    got, K = sharded_roots(ctx, f0, schedule, world)
    assert K == shard.fri_shard_plan(n0, schedule, world)[0]
    assert np.array_equal(got, O.FriProverState(f0, schedule, SEED_Z).roots())


@pytest.mark.parametrize("name,schedule,world", [("fold2", [2] * 18, 8), ("paper", [16, 16, 8], 8), ("paper", [16, 16, 8], 2)])
def test_sharded_roots_match_golden_2p18(ctx, smw, O, rnd, name, schedule, world):
    """2^18 elements: the sharded roots against the pinned oracle's (tests/golden/full_size_golden.json, section shard18) and
    against the unsharded GPU path"""
    import json, os
    n0 = 1 << 18
    f0 = rnd(99, n0)
    got, _ = sharded_roots(ctx, f0, schedule, world)
    path = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "full_size_golden.json")
    gold = json.load(open(path)).get("shard18") if os.path.exists(path) else None
    if gold is None:
        pytest.skip("golden section shard18 not generated")
    assert [O.fr_hex(r) for r in got] == gold[name]["roots"]
    assert np.array_equal(got, smw.fri_build_transcript(ctx, f0, smw.FriProverParams(schedule, SEED_Z)).roots())


def test_build_from_level_offset(ctx, smw, O, rnd):
    """layers 2.. of an instance rebuilt from f_2 with level0 = 2 carry the same roots as the full build"""
    n0, schedule = 1 << 8, [2, 4, 2, 2]
    f0 = rnd(5, n0)
    full = smw.fri_build_transcript(ctx, f0, smw.FriProverParams(schedule, SEED_Z))
    f2 = full.layers[2].f
    part = smw.fri_build_transcript(ctx, f2, smw.FriProverParams(schedule[2:], SEED_Z), level0=2)
    assert np.array_equal(part.roots(), full.roots()[2:])
    assert np.array_equal(full.roots(), O.FriProverState(f0, schedule, SEED_Z).roots())
