"""Regenerates tests/golden/*.json with the CPU oracle (oracle/oracle.c).

The reference is Rust and cannot be run in this image, so these are NOT outputs of the reference: the oracle is pinned by
the reference's published `proof_bytes` (proof_bytes.json, copied from crates/channel/benchmarkdata.csv:2-9) and these
files freeze what the pinned oracle computes, so that (a) a later change to the oracle cannot drift silently
(tests/test_golden.py, CPU) and (b) the GPU can be checked at sizes where running the oracle in the test would be slow
(tests/test_gpu_golden.py).

    python tests/golden/make_golden.py
"""
import json
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
from oracle import oracle as O          # noqa: E402
from bench import gen_columns, SEED, SHIFT   # noqa: E402  (the bench's synthetic-input generator)


def hx(a):
    return O.fr_hex(a)


def main():
    O.lib()
    O.set_threads(O.max_threads())
    # 1. the reference's published KAT (copied data, not computed)
    json.dump({"source": "crates/channel/benchmarkdata.csv:2-9 (column proof_bytes; schedule [16,16,8], r = 32, seed_z = 0xDEEFBAAD)",
               "proof_bytes": {"11": 39592, "12": 52000, "13": 60968, "14": 72936, "15": 87736, "16": 101976, "17": 119952, "18": 140032}},
              open(os.path.join(HERE, "proof_bytes.json"), "w"), indent=1)
    # 2. LDE + commit roots on the bench generator (cfg1/cfg2 shape at several sizes)
    shift = O.to_mont([SHIFT])[0]
    lde = {}
    for log_n in (4, 8, 12, 16):
        cols = gen_columns(SEED, 0, 2, 1 << log_n)
        lde[str(log_n)] = [hx(O.commitment_commit(O.lde(cols[c], 2, shift), c).root()) for c in range(2)]
    # 3. FRI commit-phase roots: the published schedule at k = 11, 14 and fold factor 2 at k = 12
    fri = {}
    for name, k, sched in (("paper_k11", 11, [16, 16, 8]), ("paper_k14", 14, [16, 16, 8]), ("fold2_k12", 12, [2] * 12), ("uni32x3_k15", 15, [32, 32, 32])):
        f0 = gen_columns(0x5EED0003, 0, 1, 1 << k)[0]
        st = O.FriProverState(f0, sched, 0xDEEFBAAD)
        fri[name] = {"k": k, "schedule": sched, "roots": [hx(r) for r in st.roots()], "z": [hx(st.z(l)) for l in range(len(sched))]}
    # 4. Merkle-folded sum-check commit loop
    table = gen_columns(0x5EED0006, 0, 1, 1 << 10)[0]
    roots, rs, fin = O.mf_commit_phase(table, 2025)
    mf = {"k": 10, "ds_tag": 2025, "roots": [hx(r) for r in roots], "r": [hx(r) for r in rs], "final": hx(fin)}
    # 5. permutation outputs for the five widths on the state (1, ..., t)
    perm = {}
    for t in (9, 17, 33, 65, 129):
        out = O.permute(O.Params.for_width(t), O.to_mont(range(1, t + 1)).reshape(1, t, 4))
        perm[str(t)] = [hx(out[0, 0]), hx(out[0, t - 1])]
    json.dump({"generator": "tests/golden/make_golden.py (CPU oracle; inputs = bench.gen_columns)", "lde_commit_roots_blowup4_shift5": lde,
               "fri": fri, "mf": mf, "permute_1_to_t": perm}, open(os.path.join(HERE, "oracle_golden.json"), "w"), indent=1)
    print("written", os.listdir(HERE))


if __name__ == "__main__":
    main()
