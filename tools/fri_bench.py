#!/usr/bin/env python
"""Times the FRI-like commit phase (deep_ali/src/fri.rs:231-312) on one B200: BASELINE cfg3 (2^22 elements, fold factor 2 down
to one element, per-layer Poseidon-Merkle trees) and the reference's published schedule [16,16,8] at the same size.
Prints one JSON object; f0 is resident on the device when the timed region starts, roots come back to the host inside it."""
import json, os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from stark_mlwe_b200 import api as smw
from bench import gen_columns

_st = torch.cuda.Stream(); torch.cuda.set_stream(_st)
ctx = smw.Context(0, stream=_st.cuda_stream)       # torch's work and the library's share one explicit stream
out = {}
for name, k, sched in (("cfg3_fold2_2p22", 22, [2] * 22), ("paper_16_16_8_2p22", 22, [16, 16, 8]), ("paper_16_16_8_2p18", 18, [16, 16, 8])):
    f0 = torch.from_numpy(gen_columns(0x5EED0003, 0, 1, 1 << k)[0].view(np.int64)).cuda()
    prm = smw.FriProverParams(sched, 0xDEEFBAAD)
    for _ in range(2):
        st = smw.fri_build_transcript(ctx, f0, prm); st.roots(); del st
    ctx.profile(True)
    torch.cuda.synchronize(); t0 = time.perf_counter()
    reps = 3
    for _ in range(reps):
        st = smw.fri_build_transcript(ctx, f0, prm); roots = st.roots(); del st
    torch.cuda.synchronize(); dt = (time.perf_counter() - t0) / reps
    prof = ctx.profile_read(); ctx.profile(False)
    leaves = sum(l for l in [(1 << k)] + [0])
    n, committed = 1 << k, 0
    for m in sched + [1]:
        committed += n; n //= m if m > 1 else 1
    out[name] = {"n0": 1 << k, "schedule": sched if len(sched) < 6 else "[2]x%d" % len(sched), "ms": dt * 1e3, "n0_per_s": (1 << k) / dt, "committed_leaves": committed,
                 "committed_leaves_per_s": committed / dt, "class_ms": {kk: v[0] / reps for kk, v in prof.items() if v[1]}}
print(json.dumps(out))
