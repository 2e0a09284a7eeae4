#!/usr/bin/env python
"""End-to-end `deep_fri_prove` (deep_ali/src/fri.rs:601-641; the only workload the reference itself publishes numbers for,
crates/channel/benchmarkdata.csv: schedule [16,16,8], r = 32, seed_z = 0xDEEFBAAD) on one B200 at 2^LOG_N evaluations:
build_f0 (four ALI sponges + DEEP-ALI merge) -> fri_build_transcript -> fri_prove_queries -> proof size.  One JSON line with the
phase times.  At 2^26 (BASELINE configs[4]) the four strictly serial sponges (n0 / 16 dependent permutations each) are the whole
story: ~0.2 ms per permutation on one warp each.

    python tools/deep_fri_e2e.py 22
"""
import json, os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from stark_mlwe_b200 import api as smw
from bench import gen_columns

log_n = int(sys.argv[1]) if len(sys.argv) > 1 else 20
n0 = 1 << log_n
ctx = smw.Context(0)
a, s, e, t = (gen_columns(0x5EED0005, c, 1, n0)[0] for c in range(4))
smw.deep_fri_prove(ctx, a[:1 << 12], s[:1 << 12], e[:1 << 12], t[:1 << 12], [16, 16, 8], 32, 0xDEEFBAAD)      # warm-up: parameter sets, programs
ctx.sync()
T = [time.perf_counter()]
f0, z = smw.deep_ali_build_f0(ctx, a, s, e, t); ctx.sync(); T.append(time.perf_counter())
st = smw.fri_build_transcript(ctx, f0, smw.FriProverParams([16, 16, 8], 0xDEEFBAAD)); ctx.sync(); T.append(time.perf_counter())
qp = smw.fri_prove_queries(ctx, st, 32); ctx.sync(); T.append(time.perf_counter())
total = T[-1] - T[0]
ref_rate = 1163.0      # elems/s of the reference on its own machine (Apple silicon, one thread), benchmarkdata.csv:9 — linear in n0 over k = 11..18
print(json.dumps({"log_n0": log_n, "schedule": [16, 16, 8], "r": 32, "seconds": total, "n0_per_s": n0 / total,
                  "phase_s": {"build_f0 (4 serial ALI sponges + merge)": T[1] - T[0], "fri_build_transcript": T[2] - T[1], "fri_prove_queries": T[3] - T[2]},
                  "dependent_permutations_per_sponge": n0 // 16, "proof_bytes": qp.size_bytes(),
                  "root0": [int(v) for v in st.roots()[0]],
                  "reference_published": {"elems_per_s": ref_rate, "source": "crates/channel/benchmarkdata.csv:2-9 (Apple silicon, 1 thread)", "extrapolated_seconds": n0 / ref_rate},
                  "speedup_vs_reference_published_rate": (n0 / total) / ref_rate}))
