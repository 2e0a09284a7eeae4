#!/usr/bin/env python
"""Times the commit loop of the Merkle-folded sum-check (channel/src/lib.rs:600-657: commit, r_i from the root, fold, repeat) on one
B200 for tables of 2^k elements resident in HBM; prints one JSON object.  The loop is strictly sequential in its k rounds."""
import json, os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from stark_mlwe_b200 import api as smw
from bench import gen_columns_torch

_st = torch.cuda.Stream(); torch.cuda.set_stream(_st)
ctx = smw.Context(0, stream=_st.cuda_stream)       # torch's work and the library's share one explicit stream
out = {}
for k in (12, 16, 20, 22):
    table = gen_columns_torch(0x5EED0006, 0, 1, 1 << k, "cuda")[0]
    for _ in range(2):
        smw.mf_commit_phase(ctx, table, 2025)
    ctx.profile(True)
    torch.cuda.synchronize(); t0 = time.perf_counter()
    reps = 3
    for _ in range(reps):
        roots, rs, fin = smw.mf_commit_phase(ctx, table, 2025)
    torch.cuda.synchronize(); dt = (time.perf_counter() - t0) / reps
    prof = ctx.profile_read(); ctx.profile(False)
    committed = 2 * (1 << k) - 1
    out["k%d" % k] = {"ms": dt * 1e3, "committed_leaves": committed, "committed_leaves_per_s": committed / dt, "rounds": k,
                      "kernel_ms": {kk: v[0] / reps for kk, v in prof.items() if v[1]}, "launches": {kk: v[1] // reps for kk, v in prof.items() if v[1]}}
print(json.dumps(out))
