"""cfg5-shaped run at N GPUs under torchrun: ONE vector of 2^LOG_N elements (cyclic over the ranks) -> four-step NTT with one NCCL
all-to-all -> range-sharded FRI commit phase on the resulting chunks (NCCL cap all-gathers).  Rank 0 prints one JSON line; at
sizes <= 2^22 the result is checked against the single-GPU path (smw_ntt + smw_fri_build_transcript) on rank 0."""
import json
import os
import sys
import time

import numpy as np
import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from stark_mlwe_b200 import api as smw, shard  # noqa: E402
from bench import gen_columns  # noqa: E402

rank, world, local = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
if world > 1:
    dist.init_process_group("nccl", device_id=dev)
ctx = smw.Context(local)
LOG_N = int(os.environ.get("NTT_LOG_N", 22))
N = 1 << LOG_N
M = N // world
x_all = gen_columns(0x5EED0005, 0, 1, N)[0]
mine = torch.from_numpy(np.ascontiguousarray(x_all[rank::world]).view(np.int64)).to(dev)      # cyclic slice, resident in HBM
nb, fb = shard.GpuNttBackend(ctx), shard.GpuFriBackend(ctx)
gather = lambda a: shard.all_gather_rows(a, device=dev)  # noqa: E731
sched = [16, 16, 8]
res = {}
for it in range(3):
    buf = mine.clone()
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize(); t0 = time.perf_counter()
    chunks = shard.ntt_cyclic_to_chunks(nb, buf, LOG_N, rank, world)
    ctx.sync(); torch.cuda.synchronize(); t1 = time.perf_counter()
    firsts = [M * k1 + rank * (M // world) for k1 in range(world)]
    roots = shard.fri_roots_sharded_chunks(fb, [chunks[c] for c in range(world)], firsts, N, world, sched, 0xDEEFBAAD, gather=gather)
    ctx.sync(); torch.cuda.synchronize(); t2 = time.perf_counter()
    dt = torch.tensor([t1 - t0, t2 - t1], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(dt, op=dist.ReduceOp.MAX)
ok = None
if rank == 0 and LOG_N <= 22:
    X = smw.fft(ctx, x_all)
    want = smw.fri_build_transcript(ctx, X, smw.FriProverParams(sched, 0xDEEFBAAD)).roots()
    got0 = chunks[0].cpu().numpy().view(np.uint64)
    ok = bool(np.array_equal(roots, want)) and bool(np.array_equal(got0, X[rank * (M // world):(rank + 1) * (M // world)]))
    assert ok, "distributed NTT + FRI differs from the single-GPU path"
if rank == 0:
    print(json.dumps({"log_n": LOG_N, "world": world, "schedule": sched, "ntt_ms": float(dt[0]) * 1e3, "fri_ms": float(dt[1]) * 1e3,
                      "elems_per_s_ntt": N / float(dt[0]), "all_to_all_bytes_per_rank": M * 32 * (world - 1) // max(world, 1),
                      "equal_single_gpu": ok}))
if world > 1:
    dist.barrier(); dist.destroy_process_group()
