"""Parity of the in-library multi-GPU data plane (csrc/comm.cu, smw_fri_roots_sharded, smw_ntt_cyclic_to_chunks,
smw_mf_commit_phase_sharded) under torchrun with N >= 2 ranks: every sharded result must equal the single-GPU result computed by
the same rank (and the goldens of the pinned oracle where they exist).  Rank 0 prints one JSON line; any mismatch raises.

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29533 tools/multi_gpu_check.py
"""
import json
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from stark_mlwe_b200 import api as smw, shard  # noqa: E402
from bench import gen_columns  # noqa: E402

SEED_Z = 0xDEEFBAAD
rank, world, local = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
if world > 1:
    dist.init_process_group("nccl", device_id=dev)
_st = torch.cuda.Stream(); torch.cuda.set_stream(_st)
ctx = smw.Context(local, stream=_st.cuda_stream)       # torch's work and the library's share one explicit stream
comm = shard.make_comm(ctx)
res = {"world": world, "checks": []}


def ok(name, cond):
    assert cond, "rank %d: %s" % (rank, name)
    res["checks"].append(name)


def dev_t(a):
    return torch.from_numpy(np.ascontiguousarray(a).view(np.int64)).to(dev)


gold_path = os.path.join(ROOT, "tests", "golden", "full_size_golden.json")
gold = json.load(open(gold_path)) if os.path.exists(gold_path) else {}

# 1. range-sharded FRI commit phase, contiguous slices, 2^18 (+ goldens) and a ragged schedule
for name, log_n0, sched in (("fold2_2p18", 18, [2] * 18), ("paper_2p18", 18, [16, 16, 8]), ("mixed_2p15", 15, [4, 2, 8, 16])):
    n0 = 1 << log_n0
    f0 = gen_columns(0x5EED0003, 0, 1, n0)[0]
    prm = smw.FriProverParams(sched, SEED_Z)
    want = smw.fri_build_transcript(ctx, f0, prm).roots()
    k = n0 // world
    c0 = ctx.collective_count()
    got = smw.fri_roots_sharded(ctx, comm, dev_t(f0[rank * k:(rank + 1) * k]).view(1, k, 4), [rank * k], n0, prm)
    ok("fri_sharded_" + name, np.array_equal(got, want))
    ok("fri_sharded_%s_one_collective" % name, ctx.collective_count() - c0 == (1 if world > 1 else 0))
if "shard18" in gold:       # the oracle's roots on the reference's own generator (StdRng seed 99), 2^18
    from oracle import oracle as O
    O.lib()
    f0 = O.StdRng.seed_from_u64(99).fr_rand(1 << 18)
    k = (1 << 18) // world
    for nm, sched in (("fold2", [2] * 18), ("paper", [16, 16, 8])):
        got = smw.fri_roots_sharded(ctx, comm, dev_t(f0[rank * k:(rank + 1) * k]).view(1, k, 4), [rank * k], 1 << 18, smw.FriProverParams(sched, SEED_Z))
        ok("fri_sharded_golden_" + nm, [O.fr_hex(r) for r in got] == gold["shard18"][nm]["roots"])

# 2. one NTT across the ranks (cyclic in, chunks out) + FRI on the chunks, against the single-GPU transform
for log_n in (12, 16, 20):
    N = 1 << log_n
    M = N // world
    x = gen_columns(0x5EED0005, 0, 1, N)[0]
    X = smw.fft(ctx, x)
    chunks = smw.ntt_cyclic_to_chunks(ctx, comm, dev_t(x[rank::world]), log_n)
    ctx.sync()
    got = chunks.cpu().numpy().view(np.uint64)
    for k1 in range(world):
        a = M * k1 + rank * (M // world)
        ok("ntt_2p%d_chunk%d" % (log_n, k1), np.array_equal(got[k1], X[a:a + M // world]))
    firsts = [M * k1 + rank * (M // world) for k1 in range(world)]
    prm = smw.FriProverParams([16, 16, 8] if log_n >= 12 else [2] * log_n, SEED_Z)
    roots = smw.fri_roots_sharded(ctx, comm, chunks, firsts, N, prm)
    ok("fri_on_ntt_chunks_2p%d" % log_n, np.array_equal(roots, smw.fri_build_transcript(ctx, X, prm).roots()))

# 3. Merkle-folded sum-check commit loop, table sharded by contiguous ranges
for k in (4, 9, 13, 17):
    n = 1 << k
    if n < world:
        continue
    table = gen_columns(0x5EED0006, 0, 1, n)[0]
    want = smw.mf_commit_phase(ctx, table, 2025)
    got = smw.mf_commit_phase_sharded(ctx, comm, dev_t(table[rank * (n // world):(rank + 1) * (n // world)]), 2025)
    ok("mf_sharded_k%d" % k, all(np.array_equal(a, b) for a, b in zip(got, want)))

# 4. column roots all-gather through the library
mine = dev_t(np.full((3, 4), rank + 1, dtype=np.uint64))
allr = comm.all_gather(mine).cpu().numpy() if comm is not None else mine.cpu().numpy()
ctx.sync()
ok("all_gather", all(int(allr[3 * r, 0]) == r + 1 for r in range(world)))

if world > 1:
    dist.barrier()
if rank == 0:
    res["ok"] = True
    res["n_checks"] = len(res["checks"])
    print(json.dumps(res))
if comm is not None:
    comm.close()
ctx.close()
if world > 1:
    dist.destroy_process_group()
