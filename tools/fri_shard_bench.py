"""cfg3 at N GPUs: range-sharded FRI commit phase of one 2^22 polynomial under torchrun (one process per GPU, NCCL all-gather
of the per-layer caps).  Prints one JSON line from rank 0; every rank checks that all ranks computed the same roots and (at
world 1 or with --check) that they equal the unsharded GPU result."""
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
if world > 1:
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
ctx = smw.Context(local)
LOG_N = int(os.environ.get("FRI_LOG_N", 22))
n0 = 1 << LOG_N
k = n0 // world
f_all = gen_columns(0x5EED0003, 0, 1, n0)[0]
mine = f_all[rank * k:(rank + 1) * k]
be = shard.GpuFriBackend(ctx)
gather = lambda x: shard.all_gather_rows(x, device=torch.device("cuda", local))  # noqa: E731
out = {}
for name, sched in (("fold2", [2] * LOG_N), ("paper_16_16_8", [16, 16, 8])):
    for it in range(3):                                   # 2 warm-ups (parameter derivation, allocator), 1 timed
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize(); t0 = time.perf_counter()
        roots = shard.fri_roots_sharded(be, mine, n0, rank, world, sched, 0xDEEFBAAD, gather=gather)
        ctx.sync(); torch.cuda.synchronize()
        dt = torch.tensor([time.perf_counter() - t0], dtype=torch.float64, device="cuda")
        if world > 1:
            dist.all_reduce(dt, op=dist.ReduceOp.MAX)
    same = shard.all_gather_rows(roots, device=torch.device("cuda", local)).reshape(world, -1, 4)
    assert all(np.array_equal(same[0], same[r]) for r in range(world)), "ranks disagree on the roots"
    ok = None
    if rank == 0 and LOG_N <= 22:
        want = smw.fri_build_transcript(ctx, f_all, smw.FriProverParams(sched, 0xDEEFBAAD)).roots()
        ok = bool(np.array_equal(roots, want))
        assert ok, "sharded roots differ from the unsharded GPU path"
    leaves = sum(n0 // int(np.prod(sched[:l])) for l in range(len(sched) + 1))
    out[name] = {"n0": n0, "world": world, "ms": float(dt.item()) * 1e3, "n0_per_s": n0 / float(dt.item()), "committed_leaves": int(leaves),
                 "committed_leaves_per_s": leaves / float(dt.item()), "local_layers": shard.fri_shard_plan(n0, sched, world)[0],
                 "roots_equal_unsharded": ok}
if rank == 0:
    print(json.dumps(out))
if world > 1:
    dist.barrier(); dist.destroy_process_group()
