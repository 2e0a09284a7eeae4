#!/usr/bin/env python
"""Per-repetition wall times of the FRI commit phase (same workloads as tools/fri_bench.py), to show the run-to-run spread."""
import json, os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from stark_mlwe_b200 import api as smw
from bench import gen_columns

ctx = smw.Context(0, stream=torch.cuda.current_stream().cuda_stream)
out = {}
for name, k, sched in (("cfg3_fold2_2p22", 22, [2] * 22), ("paper_16_16_8_2p22", 22, [16, 16, 8])):
    f0 = torch.from_numpy(gen_columns(0x5EED0003, 0, 1, 1 << k)[0].view(np.int64)).cuda()
    prm = smw.FriProverParams(sched, 0xDEEFBAAD)
    ts, tf = [], []
    for _ in range(8):
        torch.cuda.synchronize(); t0 = time.perf_counter()
        st = smw.fri_build_transcript(ctx, f0, prm); roots = st.roots()
        torch.cuda.synchronize(); t1 = time.perf_counter()
        del st
        torch.cuda.synchronize(); t2 = time.perf_counter()
        ts.append(round((t1 - t0) * 1e3, 1)); tf.append(round((t2 - t1) * 1e3, 1))
    out[name] = {"build_ms": ts, "free_ms": tf}
print(json.dumps(out))
