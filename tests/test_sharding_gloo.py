"""World-size-2 (and 3) CPU test of the multi-GPU host logic (stark_mlwe_b200/shard.py) over gloo: column partition, global
tree labels, and the root all-gather.  Each rank "commits" its columns with the CPU oracle standing in for the device
(test infrastructure); the gathered roots must equal the single-process result in column order."""
import os
import socket

import numpy as np
import pytest
import torch.distributed as dist
import torch.multiprocessing as mp

from stark_mlwe_b200 import shard


def _free_port():
    s = socket.socket(); s.bind(("127.0.0.1", 0)); p = s.getsockname()[1]; s.close(); return p


def _worker(rank, world, port, n_cols, log_n, q):
    os.environ["MASTER_ADDR"] = "127.0.0.1"; os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from oracle import oracle as O
    from bench import gen_columns, SEED, SHIFT
    O.lib(); O.set_threads(1)
    c0, c1 = shard.column_range(rank, world, n_cols)
    cols = gen_columns(SEED, c0, c1 - c0, 1 << log_n)
    shift = O.to_mont([SHIFT])[0]
    mine = np.zeros((c1 - c0, 4), dtype=np.uint64)
    for c in range(c1 - c0):
        mine[c] = O.commitment_commit(O.lde(cols[c], 2, shift), c0 + c).root()      # tree_label = global column index
    allr = shard.gather_roots(mine, n_cols)
    if rank == 0:
        q.put(allr)
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("world,n_cols", [(2, 4), (3, 5)])
def test_column_sharded_commit_matches_single_process(O, world, n_cols):
    from bench import gen_columns, SEED, SHIFT
    log_n = 4
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, n_cols, log_n, q)) for r in range(world)]
    for p in procs:
        p.start()
    got = q.get(timeout=120)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    cols = gen_columns(SEED, 0, n_cols, 1 << log_n)
    shift = O.to_mont([SHIFT])[0]
    want = np.stack([O.commitment_commit(O.lde(cols[c], 2, shift), c).root() for c in range(n_cols)])
    assert np.array_equal(got, want)


def test_ranges():
    assert [shard.column_range(r, 4, 10) for r in range(4)] == [(0, 3), (3, 6), (6, 8), (8, 10)]
    assert [shard.column_range(r, 8, 256) for r in range(8)] == [(32 * r, 32 * r + 32) for r in range(8)]
    assert [shard.column_range(r, 3, 2) for r in range(3)] == [(0, 1), (1, 2), (2, 2)]
    # a 2^22-leaf layer, fold 2 x arity 2 -> aligned to 4; ranges tile the layer exactly
    rs = [shard.leaf_range(r, 8, 1 << 22, 4) for r in range(8)]
    assert rs[0][0] == 0 and rs[-1][1] == 1 << 22 and all(a[1] == b[0] for a, b in zip(rs, rs[1:])) and all(a % 4 == 0 for a, _ in rs)
    assert shard.leaf_range(5, 8, 8, 4) == (8, 8)       # fewer blocks than ranks: empty tail ranges


# ----------------------------------------------------------------------------- one polynomial, range-sharded FRI commit phase
def _fri_worker(rank, world, port, n0, schedule, seed_z, q):
    os.environ["MASTER_ADDR"] = "127.0.0.1"; os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    import sys
    sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
    from oracle import oracle as O
    from oracle_shard_backend import OracleFriBackend
    O.lib(); O.set_threads(1)
    f0 = O.StdRng.seed_from_u64(n0).fr_rand(n0)                 # every rank derives the same polynomial, keeps its slice only
    k = n0 // world
    roots = shard.fri_roots_sharded(OracleFriBackend(), f0[rank * k:(rank + 1) * k], n0, rank, world, schedule, seed_z)
    q.put((rank, roots))
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("world,n0,schedule", [(2, 64, [2] * 6), (2, 512, [8, 8]), (4, 256, [4, 2, 2])])
def test_range_sharded_fri_roots_match_single_process(O, world, n0, schedule):
    seed_z = 0xDEEFBAAD
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_fri_worker, args=(r, world, port, n0, schedule, seed_z, q)) for r in range(world)]
    for p in procs:
        p.start()
    got = dict(q.get(timeout=240) for _ in range(world))
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    want = O.FriProverState(O.StdRng.seed_from_u64(n0).fr_rand(n0), schedule, seed_z).roots()
    for r in range(world):
        assert np.array_equal(got[r], want), "rank %d" % r


def test_fri_shard_plan():
    # cfg3 at 8 GPUs: 2^22, fold 2: slices of 2^19 fold locally down to one element per rank (layer 19); layers 19..22 are the tail
    K, n = shard.fri_shard_plan(1 << 22, [2] * 22, 8)
    assert K == 19 and n[19] == 8
    # [16, 16, 8] on 2^22 at 8 GPUs: 2^19 -> 2^15 -> 2^11 -> 2^8 per rank: every layer (and the final one) is local
    K, n = shard.fri_shard_plan(1 << 22, [16, 16, 8], 8)
    assert K == 4 and n[3] == 1 << 11
    assert shard.fri_shard_plan(64, [2] * 6, 1)[0] == 7


# ----------------------------------------------------------------------------- one NTT across ranks (four-step), then FRI on its chunks
def _ntt_worker(rank, world, port, log_n, schedule, q):
    os.environ["MASTER_ADDR"] = "127.0.0.1"; os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    import sys
    sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
    from oracle import oracle as O
    from oracle_shard_backend import OracleFriBackend, OracleNttBackend
    O.lib(); O.set_threads(1)
    N = 1 << log_n
    M = N // world
    x = O.StdRng.seed_from_u64(log_n).fr_rand(N)
    chunks = shard.ntt_cyclic_to_chunks(OracleNttBackend(), x[rank::world].copy(), log_n, rank, world)     # cyclic in
    firsts = [M * k1 + rank * (M // world) for k1 in range(world)]
    roots = shard.fri_roots_sharded_chunks(OracleFriBackend(), [chunks[c] for c in range(world)], firsts, N, world, schedule, 0xDEEFBAAD)
    q.put((rank, np.asarray(chunks), roots))
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("world,log_n,schedule", [(2, 6, [2] * 6), (2, 7, [8, 4])])
def test_distributed_ntt_then_sharded_fri(O, world, log_n, schedule):
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_ntt_worker, args=(r, world, port, log_n, schedule, q)) for r in range(world)]
    for p in procs:
        p.start()
    got = {}
    for _ in range(world):
        r, ch, roots = q.get(timeout=240)
        got[r] = (ch, roots)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    N = 1 << log_n
    M = N // world
    X = O.ntt(O.StdRng.seed_from_u64(log_n).fr_rand(N))                           # fft/src/lib.rs:6-12 convention, natural order
    want_roots = O.FriProverState(X, schedule, 0xDEEFBAAD).roots()
    for r in range(world):
        for k1 in range(world):
            lo = M * k1 + r * (M // world)
            assert np.array_equal(got[r][0][k1], X[lo:lo + M // world]), "rank %d chunk %d" % (r, k1)
        assert np.array_equal(got[r][1], want_roots)


# ----------------------------------------------------------------------------- Merkle-folded sum-check over a range-sharded table
def _mf_worker(rank, world, port, k, ds_tag, q):
    os.environ["MASTER_ADDR"] = "127.0.0.1"; os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    import sys
    sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
    from oracle import oracle as O
    from oracle_shard_backend import OracleMfBackend
    O.lib(); O.set_threads(1)
    n = 1 << k
    table = O.StdRng.seed_from_u64(7000 + k).fr_rand(n)         # every rank derives the same table, keeps its slice only
    m = n // world
    out = shard.mf_commit_phase_sharded_host(OracleMfBackend(), table[rank * m:(rank + 1) * m], ds_tag, rank, world)
    q.put((rank, out))
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("world,k", [(2, 1), (2, 6), (4, 9), (2, 10)])
def test_range_sharded_mf_commit_phase_matches_single_process(O, world, k):
    """roots of every folded layer, every r_i and the final evaluation equal the single-process commit loop (channel/src/lib.rs:600-657) on
    every rank; slices of 1 (k = 1 on two ranks), of fewer than 16 leaves, of whole subtrees (k = 9 on four ranks: 128 leaves per rank) and mixed"""
    ds_tag = 0x5EED
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_mf_worker, args=(r, world, port, k, ds_tag, q)) for r in range(world)]
    for p in procs:
        p.start()
    got = dict(q.get(timeout=240) for _ in range(world))
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    roots, rs, fin = O.mf_commit_phase(O.StdRng.seed_from_u64(7000 + k).fr_rand(1 << k), ds_tag)
    for r in range(world):
        g_roots, g_rs, g_fin = got[r]
        assert np.array_equal(g_roots, roots), "roots, rank %d" % r
        assert np.array_equal(g_rs, rs), "challenges, rank %d" % r
        assert np.array_equal(g_fin, fin), "final evaluation, rank %d" % r
