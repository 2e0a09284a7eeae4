"""The multi-GPU entry points of the library (include/smw.h "multi-GPU data plane"; csrc/comm.cu, capi.cu) on ONE GPU:
with a NULL communicator a rank may still hold several slices, which exercises the whole slice logic (local folds, leaf
hashing with offset DS positions, caps, the packed exchange buffer, global slice order, tree tops, the replicated tail) —
only NCCL itself is left out, and that is covered by tools/multi_gpu_check.py under torchrun (run here when the box has >= 2 GPUs).
Layer roots must equal the CPU oracle's fri_build_transcript (deep_ali/src/fri.rs:231-312) bit for bit."""
import json
import os
import subprocess
import sys

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
SEED_Z = 0xDEEFBAAD
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.parametrize("n0,schedule,C,shuffle", [
    (1 << 10, [2] * 10, 8, True), (1 << 10, [2] * 10, 2, False), (1 << 12, [16, 16, 8], 4, True), (1 << 12, [16, 16, 8], 8, False),
    (1 << 11, [4, 8, 2], 2, True), (1 << 9, [128], 4, False), (1 << 6, [2, 2], 1, False), (1 << 14, [2] * 14, 16, True), (1 << 8, [], 4, True)])
def test_fri_roots_sharded_matches_oracle(ctx, smw, O, rnd, n0, schedule, C, shuffle):
    f0 = rnd(n0 + C, n0)
    k = n0 // C
    order = list(range(C))
    if shuffle:
        order = order[1::2] + order[0::2]                   # the rank holds its slices in any order; starts say where they belong
    chunks = np.stack([f0[s * k:(s + 1) * k] for s in order])
    got = smw.fri_roots_sharded(ctx, None, chunks, [s * k for s in order], n0, smw.FriProverParams(schedule, SEED_Z))
    assert np.array_equal(got, O.FriProverState(f0, schedule, SEED_Z).roots())


def test_fri_roots_sharded_rejects_bad_tilings(ctx, smw, rnd):
    f0 = rnd(1, 64)
    prm = smw.FriProverParams([2, 2], SEED_Z)
    with pytest.raises(smw.SmwError):
        smw.fri_roots_sharded(ctx, None, f0.reshape(2, 32, 4), [0, 0], 64, prm)          # the same slice twice
    with pytest.raises(smw.SmwError):
        smw.fri_roots_sharded(ctx, None, f0.reshape(2, 32, 4), [0, 16], 64, prm)         # start not a multiple of the slice length
    with pytest.raises(smw.SmwError):
        smw.fri_roots_sharded(ctx, None, f0[:48].reshape(3, 16, 4), [0, 16, 32], 64, prm)   # does not cover n0


@pytest.mark.parametrize("name,schedule,C", [("fold2", [2] * 18, 8), ("paper", [16, 16, 8], 8), ("paper", [16, 16, 8], 2)])
def test_fri_roots_sharded_golden_2p18(ctx, smw, O, rnd, name, schedule, C):
    path = os.path.join(ROOT, "tests", "golden", "full_size_golden.json")
    gold = json.load(open(path)).get("shard18") if os.path.exists(path) else None
    if gold is None:
        pytest.skip("golden section shard18 not generated")
    n0 = 1 << 18
    f0 = rnd(99, n0)
    got = smw.fri_roots_sharded(ctx, None, f0.reshape(C, n0 // C, 4), [s * (n0 // C) for s in range(C)], n0, smw.FriProverParams(schedule, SEED_Z))
    assert [O.fr_hex(r) for r in got] == gold[name]["roots"]


def test_single_rank_forms_equal_the_plain_calls(ctx, smw, O, rnd):
    import torch
    table = rnd(5, 1 << 9)
    a = smw.mf_commit_phase_sharded(ctx, None, table, 77)
    b = O.mf_commit_phase(table, 77)
    assert all(np.array_equal(x, y) for x, y in zip(a, b))
    x = rnd(6, 1 << 10)
    t = torch.from_numpy(x.view(np.int64).copy()).cuda()
    torch.cuda.synchronize()                 # the context runs on its own stream: the copy must have landed
    out = smw.ntt_cyclic_to_chunks(ctx, None, t, 10)
    ctx.sync()
    assert np.array_equal(out.cpu().numpy().view(np.uint64)[0], O.ntt(x))


def test_arena_pool_reuse_and_release(ctx, smw, O, rnd):
    """the large temporaries of the commit phase come from the context's arena pool: a second call reuses the blocks of the first, a call after
    smw_ctx_release_cached allocates afresh, a differently sized call in between takes other blocks - the roots never change"""
    n0, sched = 1 << 16, [16, 16, 8]           # 2 MiB layers: above the pool's 1 MiB threshold
    f0 = rnd(31337, n0)
    prm = smw.FriProverParams(sched, 0xDEEFBAAD)
    want = O.FriProverState(f0, sched, 0xDEEFBAAD).roots()
    chunks = f0.reshape(4, n0 // 4, 4)
    firsts = [i * (n0 // 4) for i in range(4)]
    for step in range(4):
        got = smw.fri_roots_sharded(ctx, None, chunks, firsts, n0, prm)
        assert np.array_equal(got, want), "call %d" % step
        if step == 1:
            ctx.release_cached()
        if step == 2:                           # another size in between
            g0 = rnd(4141, 1 << 15)
            st = smw.fri_build_transcript(ctx, g0, smw.FriProverParams([2] * 15, 7))
            assert np.array_equal(st.roots(), O.FriProverState(g0, [2] * 15, 7).roots())
            del st
    ctx.release_cached()


def test_multi_gpu_data_plane_under_torchrun():
    """>= 2 GPUs on the box: the NCCL path itself (all-gather of caps, all-to-all of the four-step NTT, per-round cap gathers of
    the sum-check) against the single-GPU results, two ranks"""
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs >= 2 GPUs (covered by tools/multi_gpu_check.py in the multi-GPU runs; profiles/)")
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr", "127.0.0.1",
                        "--master-port", "29533", os.path.join(ROOT, "tools", "multi_gpu_check.py")], capture_output=True, text=True, timeout=900)
    assert r.returncode == 0, r.stderr[-3000:]
    assert json.loads(r.stdout.strip().splitlines()[-1])["ok"]
