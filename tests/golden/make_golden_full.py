"""Full-size goldens: what the pinned CPU oracle computes at the sizes BASELINE.json's configs name.

One-off (tens of minutes on 8 host threads); writes tests/golden/full_size_golden.json, one section at a time so an
interrupted run keeps what it finished.  These are oracle outputs (the oracle is pinned by the reference's published
`proof_bytes`, tests/test_oracle_anchors.py), not outputs of the Rust reference, which cannot be built in this image.

    python tests/golden/make_golden_full.py [section ...]       # sections: cfg3_fold2 cfg3_paper cfg2 shard18 cfg4_slice cfg4

Inputs are bench.gen_columns and, for `shard18`, the reference's own StdRng + F::rand (seed 99), the
inputs tests/test_gpu_fri_sharded.py uses.
"""
import json
import os
import sys
import time

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
from oracle import oracle as O          # noqa: E402
from bench import gen_columns, SHIFT    # noqa: E402

OUT = os.path.join(HERE, "full_size_golden.json")
SEED_Z = 0xDEEFBAAD


def hx(a):
    return O.fr_hex(a)


def load():
    return json.load(open(OUT)) if os.path.exists(OUT) else {"generator": "tests/golden/make_golden_full.py (CPU oracle)"}


def save(d):
    json.dump(d, open(OUT + ".tmp", "w"), indent=1)
    os.replace(OUT + ".tmp", OUT)


def fri_section(f0, sched):
    st = O.FriProverState(f0, sched, SEED_Z)
    roots = st.roots()
    return {"n0": int(f0.shape[0]), "schedule": sched, "seed_z": SEED_Z, "roots": [hx(r) for r in roots],
            "roots_seed": hx(O.tr_hash_fields_tagged(b"FRI/seed", roots)),
            "final_f": hx(st.f(len(sched))[0]) if st.layer_n(len(sched)) == 1 else None,
            "arity": [st.layer_arity(l) for l in range(len(sched) + 1)]}


def lde_root(seed, col, log_n, log_blowup):
    c = gen_columns(seed, col, 1, 1 << log_n)[0]
    ext = O.lde(c, log_blowup, O.to_mont([SHIFT])[0])
    del c
    return hx(O.commitment_commit(ext, col).root())


SECTIONS = {
    # BASELINE configs[2]: 2^22-element composition polynomial, fold factor 2 to one element (and the published schedule)
    "cfg3_fold2": lambda: fri_section(gen_columns(0x5EED0003, 0, 1, 1 << 22)[0], [2] * 22),
    "cfg3_paper": lambda: fri_section(gen_columns(0x5EED0003, 0, 1, 1 << 22)[0], [16, 16, 8]),
    # BASELINE configs[1]: one full column 2^20 -> 2^22 (blow-up 4, shift 5), commitment root, first and last column
    "cfg2": lambda: {"log_n": 20, "log_blowup": 2, "seed": 0x5EED0002, "roots": {str(c): lde_root(0x5EED0002, c, 20, 2) for c in (0, 63)}},
    # the 2^18 inputs of tests/test_gpu_fri_sharded.py (StdRng seed 99)
    "shard18": lambda: {name: fri_section(O.StdRng.seed_from_u64(99).fr_rand(1 << 18), s) for name, s in (("fold2", [2] * 18), ("paper", [16, 16, 8]))},
    # BASELINE configs[3] shape (blow-up 8): 2^22 -> 2^25 first, then one column at the full 2^24 -> 2^27
    "cfg4_slice": lambda: {"log_n": 22, "log_blowup": 3, "seed": 0x5EED0004, "roots": {"0": lde_root(0x5EED0004, 0, 22, 3)}},
    "cfg4": lambda: {"log_n": 24, "log_blowup": 3, "seed": 0x5EED0004, "roots": {"0": lde_root(0x5EED0004, 0, 24, 3)}},
}


def main():
    O.lib()
    th = int(os.environ.get("GOLDEN_THREADS", O.max_threads()))
    O.set_threads(th)
    names = sys.argv[1:] or list(SECTIONS)
    for name in names:
        d = load()
        if name in d:
            print(name, "already present", flush=True)
            continue
        t0 = time.time()
        d[name] = SECTIONS[name]()
        d[name]["oracle_seconds"] = round(time.time() - t0, 1)
        d[name]["threads"] = th
        save(d)
        print(name, "done in %.0f s" % (time.time() - t0), flush=True)


if __name__ == "__main__":
    main()
