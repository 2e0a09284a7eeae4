#!/usr/bin/env python
"""Diagnosis: fold-2 FRI commit phase of 2^22 through smw_fri_build_transcript and through smw_fri_roots_sharded (one rank), wall time
per call with the per-class profile off and on.  SMW_TRACE=1 adds the library's host time per phase on stderr."""
import json, os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from stark_mlwe_b200 import api as smw
from bench import gen_columns

_st = torch.cuda.Stream(); torch.cuda.set_stream(_st)
ctx = smw.Context(0, stream=_st.cuda_stream)
k = int(sys.argv[1]) if len(sys.argv) > 1 else 22
n0 = 1 << k
f0 = torch.from_numpy(gen_columns(0x5EED0003, 0, 1, n0)[0].view(np.int64)).cuda()
prm = smw.FriProverParams([2] * k, 0xDEEFBAAD)
out = {}
def tr(): st = smw.fri_build_transcript(ctx, f0, prm); r = st.roots(); del st; return r
def sh(): return smw.fri_roots_sharded(ctx, None, f0.view(1, n0, 4), [0], n0, prm)
for name, fn in (("transcript", tr), ("sharded", sh)):
    for prof in (False, True):
        fn(); fn()
        ctx.profile(prof)
        ts = []
        for _ in range(3):
            torch.cuda.synchronize(); t0 = time.perf_counter(); fn(); torch.cuda.synchronize(); ts.append((time.perf_counter() - t0) * 1e3)
        pr = ctx.profile_read() if prof else None
        ctx.profile(False)
        out["%s_profile_%s" % (name, "on" if prof else "off")] = {"ms": ts, "class_ms": {kk: (v[0] / 3, v[1] // 3) for kk, v in pr.items() if v[1]} if pr else None}
        sys.stderr.write("==== %s profile=%s done\n" % (name, prof))
print(json.dumps(out))
