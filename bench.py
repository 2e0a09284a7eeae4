#!/usr/bin/env python
"""bench.py — committed Fr elements per second of the commitment hot path on B200.

Headline step (`value`, `e2e`) = BASELINE.json configs[1]: coset LDE (blow-up 4, shift 5) of `cols` columns of 2^20 rows, then one
arity-16 Poseidon-Merkle tree per extended column (MerkleCommitment::commit semantics, tree_label = global column index), roots
to the host.  With N GPUs every rank takes its own `cols` columns (column sharding, no data-path collective; the roots are
all-gathered over NCCL inside the timed region) — "scaling": weak.

The same JSON line carries the other halves of the metric as sub-records, each timed the same way (CUDA events + wall clock,
max over ranks, inputs resident in HBM):
  "fri"   cfg3: the FRI commit phase (deep_ali/src/fri.rs:231-312) of a 2^22-element polynomial, fold factor 2 to one element and
          the published schedule [16,16,8]; at N > 1 the polynomial is range-sharded over the ranks (one NCCL all-gather per proof)
  "cfg4"  STRONG scaling: 256 columns in total, split over the ranks, blow-up 8 (BASELINE configs[3] shape; rows per column
          reduced so that a step stays below a minute on one GPU — the full 2^24 rows run at N = 8 only)
  "cfg5"  one column of 2^26: NTT (four-step across the ranks with one NCCL all-to-all at N > 1) + range-sharded FRI [16,16,8]

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference]

Keys follow the driver contract: value (inputs resident in HBM), e2e (host buffers through the C ABI, H2D/D2H inside the timed
region), roofline (dominant kernel k_merkle_level52<17>: instructions it issues on the FP64 pipe / that pipe's live-measured rate),
cpu_baseline (the CPU oracle = C port of the reference, on a bounded sample whose roots are also used to check the GPU's),
clocks, gpu_launches.
"""
import argparse
import json
import os
import subprocess
import sys
import tempfile
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "committed Fr elems/s (LDE+Poseidon Merkle+FRI) at 1/2/4/8 B200 vs host CPU"
UNIT = "Fr elems/s"
SEED = 0x5EED0002
SHIFT = 5                    # coset shift g = field generator
SEED_Z = 0xDEEFBAAD
MM_PERM_T17 = 21408          # dense Fr multiplications per t = 17 permutation, as the reference computes
MM_PERM_T9 = 5904
MAC_PER_MM = 136             # 32x32->64 multiply-accumulates per 8-limb CIOS Montgomery multiplication
# dram__bytes_read.sum + dram__bytes_write.sum of ONE k_merkle_level52<17> launch, from the committed `ncu --set full` capture named
# below (level 0 of 16 columns x 2^18 leaves; algorithmic bytes of that launch: 16 * 2^18 * 32 * (1 + 1/16) = 142 606 336).  It is
# a profile constant, not something this run measured; the JSON says so.
TRAFFIC_NCU = {"bytes": 136450048 + 6638592, "source": "profiles/r01ac_ncu_full_merkle_level52_details.csv (level 0, 16 columns x 2^18 leaves)",
               "algorithmic_bytes_of_that_launch": 142606336}
# the same for ONE k_tc_stage<1> launch (one full round of 262 144 permutations = 2 048 byte tiles: 17 words x 4 KB read and written per
# tile = 2 x 142 606 336 B; the capture saw 142.94 MB read and 96.71 MB written, the rest of the output still in L2 when the kernel ended)
TRAFFIC_NCU_TC = {"bytes": 142940416 + 96708864, "source": "profiles/r04a_ncu_tc_stage_summary.txt (k_tc_stage<1>, level 0 of 16 columns x 2^18 leaves)",
                  "algorithmic_bytes_of_that_launch": 2 * 142606336}


def splitmix64(x):
    x = (x + np.uint64(0x9E3779B97F4A7C15))
    z = x
    z = (z ^ (z >> np.uint64(30))) * np.uint64(0xBF58476D1CE4E5B9)
    z = (z ^ (z >> np.uint64(27))) * np.uint64(0x94D049BB133111EB)
    return z ^ (z >> np.uint64(31))


def gen_columns(seed, col0, n_cols, n_rows, out=None):
    """G(seed, c, i): four splitmix64 outputs of counter seed ^ (c << 40) ^ (4 i + k) as LE limbs, top limb masked to 62
    bits (value < 2^254 < p: a valid ark-ff memory image).  -> (n_cols, n_rows, 4) uint64"""
    if out is None:
        out = np.empty((n_cols, n_rows, 4), dtype=np.uint64)
    idx = np.arange(n_rows * 4, dtype=np.uint64)
    with np.errstate(over="ignore"):
        for c in range(n_cols):
            ctr = np.uint64(seed) ^ (np.uint64(col0 + c) << np.uint64(40)) ^ idx
            v = splitmix64(ctr).reshape(n_rows, 4)
            v[:, 3] &= np.uint64((1 << 62) - 1)
            out[c] = v
    return out


def _s64(v):
    v &= (1 << 64) - 1
    return v - (1 << 64) if v >= (1 << 63) else v


def gen_columns_torch(seed, col0, n_cols, n_rows, device):
    """gen_columns on a torch device (bit-identical; int64 arithmetic wraps, logical shifts are masked arithmetic ones): the
    cfg4 / cfg5 inputs are 2-17 GB per rank and are generated where they are used"""
    import torch

    def lsr(z, s):
        return (z >> s) & ((1 << (64 - s)) - 1)
    idx = torch.arange(n_rows * 4, dtype=torch.int64, device=device)
    out = torch.empty((n_cols, n_rows, 4), dtype=torch.int64, device=device)
    for c in range(n_cols):
        x = (idx ^ _s64(seed ^ ((col0 + c) << 40))) + _s64(0x9E3779B97F4A7C15)
        z = (x ^ lsr(x, 30)) * _s64(0xBF58476D1CE4E5B9)
        z = (z ^ lsr(z, 27)) * _s64(0x94D049BB133111EB)
        z = (z ^ lsr(z, 31)).view(n_rows, 4)
        z[:, 3] &= (1 << 62) - 1
        out[c] = z
        del x, z
    return out


class ClockSampler:
    Q = "clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"

    def __init__(self, gpu_index):
        self.f = tempfile.NamedTemporaryFile("w+", suffix=".csv", delete=False)
        self.p = None
        try:
            self.p = subprocess.Popen(["nvidia-smi", "-i", str(gpu_index), "--query-gpu=" + self.Q, "--format=csv,noheader,nounits", "-lms", "200"],
                                      stdout=self.f, stderr=subprocess.DEVNULL)
        except Exception:
            self.p = None

    def stop(self):
        out = {"sm_mhz": None, "sm_max_mhz": None, "reasons": []}
        if self.p is None:
            return out
        self.p.terminate()
        try:
            self.p.wait(timeout=5)
        except Exception:
            self.p.kill()
        self.f.flush()
        self.f.seek(0)
        sm, mx, pw, reasons = [], [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for line in self.f.read().splitlines():
            parts = [x.strip() for x in line.split(",")]
            if len(parts) < 7:
                continue
            try:
                sm.append(float(parts[0])); mx.append(float(parts[1])); pw.append(float(parts[2]))
            except ValueError:
                continue
            for nm, v in zip(names, parts[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(nm)
        try:
            os.unlink(self.f.name)
        except OSError:
            pass
        if sm:
            # "under load" = samples in the upper half of the power range seen
            thr = (max(pw) + min(pw)) / 2 if pw else 0
            load = [s for s, p in zip(sm, pw) if p >= thr] or sm
            out = {"sm_mhz": float(np.median(load)), "sm_max_mhz": max(mx), "power_w_max": max(pw), "samples": len(sm), "reasons": sorted(reasons)}
        return out


def cpu_threads():
    """host threads for the CPU arm: every core the process may run on, whatever OMP_NUM_THREADS says (torchrun sets it to 1)"""
    try:
        return max(1, len(os.sched_getaffinity(0)))
    except AttributeError:
        return max(1, os.cpu_count() or 1)


def cpu_step_fn(O, args):
    """one step of the reference's algorithm (C port, oracle/) on the bounded CPU sample: coset LDE + arity-16 commit per column"""
    ln, nc, lb = args.cpu_log_n, args.cpu_cols, args.log_blowup
    cols = gen_columns(SEED, 0, nc, 1 << ln)
    shift = O.to_mont([SHIFT])[0]

    def step():
        return [O.commitment_commit(O.lde(cols[c], lb, shift), c).root() for c in range(nc)]
    committed = nc * (1 << (ln + lb))
    sample = "%d columns x 2^%d rows, blow-up %d (%d committed elems per step; one column of BASELINE configs[0])" % (nc, ln, 1 << lb, committed)
    return step, cols, committed, sample


def run_reference(args, rank, world):
    """The reference's own algorithm for the path on the host cores: the CPU oracle (a C port of the Rust reference — the
    reference itself cannot be built here: no cargo/rustc), all host threads, on a bounded sample of the workload."""
    if rank != 0:
        return
    from oracle import oracle as O
    O.lib()
    th = cpu_threads()
    O.set_threads(th)
    step, _cols, committed, sample = cpu_step_fn(O, args)
    for _ in range(args.warmup):
        step()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        step()
    dt = (time.perf_counter() - t0) / args.steps
    O.set_threads(1)
    t1 = time.perf_counter(); step(); single = committed / (time.perf_counter() - t1)
    val = committed / dt
    print(json.dumps({
        "impl": "reference", "metric": METRIC, "value": val, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": dt * 1e3, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u64 (4x64-bit Montgomery limbs)",
        "data": "synthetic",
        "config": {"workload": "cfg2: 2^%d-row trace, %d columns, blow-up %d: coset LDE + arity-16 Poseidon-Merkle commit per column" % (args.log_n, args.cols, 1 << args.log_blowup),
                   "sample": sample, "field": "Pallas Fr", "threads": th,
                   "note": "C port of the reference (oracle/), OpenMP over all host cores; the Rust reference is single-threaded: single_thread_value"},
        "cpu_baseline": {"value": val, "unit": UNIT, "cores": th, "kind": "port", "sample": sample, "single_thread_value": single},
        "e2e": {"value": val, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }))


def tree_perms(n_leaves, arity, rate):
    """permutations of one arity-`arity` tree over n_leaves raw leaves (merkle/src/lib.rs:164-181; 4 DS fields + children + 1 per node)"""
    perms, cn = 0, n_leaves
    while cn > 1:
        full, rem = cn // arity, cn % arity
        perms += full * ((4 + arity + 1 + rate - 1) // rate) + (((4 + rem + 1 + rate - 1) // rate) if rem else 0)
        cn = (cn + arity - 1) // arity
    return perms


def sum_levels_bytes(N):
    b, cn = 0, N
    while cn > 1:
        nn = (cn + 15) // 16
        b += nn * 32 + (nn * 32 if nn > 1 else 0)     # written once, read once by the next level
        cn = nn
    return b


def peak_hbm():
    try:
        return float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"])
    except Exception:
        return 6650.0     # B200_PROFILING.md fallback


# ------------------------------------------------------------------------------------------------------------------ the GPU arm
class Timer:
    """CUDA events on the context's stream + wall clock, max over ranks (a call returns only after its results are on the host,
    so device time and wall time agree; the larger is kept)"""

    def __init__(self, torch, dist, stream, world):
        self.torch, self.dist, self.stream, self.world = torch, dist, stream, world

    def barrier(self):
        if self.world > 1:
            self.dist.barrier()
        self.torch.cuda.synchronize()

    def run(self, fn, steps):
        torch = self.torch
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        self.barrier()
        w0 = time.perf_counter()
        e0.record(self.stream)
        r = None
        for _ in range(steps):
            r = fn()
        e1.record(self.stream)
        self.barrier()
        wall = time.perf_counter() - w0
        t = torch.tensor([max(e0.elapsed_time(e1) * 1e-3, wall)], dtype=torch.float64, device="cuda")
        if self.world > 1:
            self.dist.all_reduce(t, op=self.dist.ReduceOp.MAX)
        return float(t.item()), r


def fri_leg(ctx, smw, comm, T, rank, world, torch, steps):
    """cfg3: FRI commit phase of ONE 2^22-element polynomial (deep_ali/src/fri.rs:231-312), fold factor 2 to one element and the
    published [16,16,8]; range-sharded over the ranks at N > 1 (smw_fri_roots_sharded: one all-gather per proof).  The roots are
    checked against the goldens of the pinned oracle (tests/golden/full_size_golden.json) in the same run."""
    log_n0 = 22
    n0 = 1 << log_n0
    k = n0 // world
    f_loc = gen_columns_torch(0x5EED0003, 0, 1, n0, "cuda")[0][rank * k:(rank + 1) * k].contiguous().view(1, k, 4)
    gold = {}
    try:
        gold = json.load(open(os.path.join(ROOT, "tests", "golden", "full_size_golden.json")))
    except Exception:
        pass
    out = {"n0": n0, "sharding": "contiguous ranges, %d per rank; one NCCL all-gather per proof" % k if world > 1 else "one GPU"}
    P = 0x40000000000000000000000000000000224698FC0994A8DD8C46EB2100000001
    rinv = pow(1 << 256, -1, P)
    for name, sched, gsec in (("fold2", [2] * log_n0, "cfg3_fold2"), ("paper_16_16_8", [16, 16, 8], "cfg3_paper")):
        prm = smw.FriProverParams(sched, SEED_Z)
        fn = lambda: smw.fri_roots_sharded(ctx, comm, f_loc, [rank * k], n0, prm)      # noqa: E731
        fn(); fn()
        ctx.profile(True)
        c0 = ctx.collective_count()
        reps = []                                        # every proof timed on its own: the record is the MEDIAN (a proof is 0.1-0.2 s, so one host
        for _ in range(max(steps, 3)):                   # hiccup - a slow cudaFree of the arenas, say - would otherwise decide the mean); all reps are listed
            t1, roots = T.run(fn, 1)
            reps.append(t1)
        n_reps = len(reps)
        secs = sorted(reps)[n_reps // 2]
        prof = ctx.profile_read(); ctx.profile(False)
        prof = {kk: (v[0] / n_reps, v[1] // n_reps) for kk, v in prof.items()}
        committed, n = 0, n0
        for m in sched + [1]:
            committed += n
            n //= m if m > 1 else 1
        rec = {"ms": secs * 1e3, "ms_reps": [round(t * 1e3, 3) for t in reps], "n0_per_s": n0 / secs, "committed_leaves": committed, "committed_leaves_per_s": committed / secs,
               "kernel_ms": {kk: v[0] for kk, v in prof.items() if v[1]}, "launches": {kk: v[1] for kk, v in prof.items() if v[1]},
               "collectives_per_proof": (ctx.collective_count() - c0) // n_reps}
        if gsec in gold:
            hexes = ["0x%064x" % ((int(r[0]) | int(r[1]) << 64 | int(r[2]) << 128 | int(r[3]) << 192) * rinv % P) for r in roots]
            rec["roots_equal_oracle_golden"] = hexes == gold[gsec]["roots"]
        out[name] = rec
    # executed-op roofline of the t = 9 kernels that carry the fold-2 phase (pair leaves + arity-2 levels: one permutation each)
    if rank == 0:
        p9 = smw.poseidon_params_for_width(ctx, 9)
        fp_interp = p9.executed_fp64_ops()
        cnt9 = ctx.tc_node_counts(p9)                    # one-permutation hashes on the tensor-core pipeline (launches >= min_nodes), else the interpreter
        fp = cnt9["fp64_instr"] or fp_interp
        r = out["fold2"]
        perms = (2 * n0 - 1) + (2 * n0 - 2 - log_n0)     # leaves: 2 n0 - 1 over all layers; arity-2 nodes: sum over layers of (n_l - 1)
        kms = sum(v for kk, v in r["kernel_ms"].items() if kk in ("pair_leaves", "merkle_level"))
        out["fold2_t9_kernels"] = {"permutations_per_proof": perms, "fp64_ops_per_permutation": fp, "fp64_ops_per_permutation_interpreter": fp_interp,
                                   "int8_mac_per_permutation": cnt9["int8_mac"], "path": "tensor-core pipeline (k_tc_stage / k_tc_chain / k_tc_head, t = 9, groups of %d partial rounds) for launches of >= %d hashes" % (cnt9["group"], cnt9["min_nodes"]) if cnt9["fp64_instr"] else "thread-per-permutation FP64 interpreter",
                                   "kernel_ms": kms, "share_of_proof_time": kms / r["ms"], "permutations_per_s_all_ranks": perms / (kms * 1e-3) if kms else None}
    return out


def strong_leg(ctx, smw, comm, T, rank, world, torch, shard, log_n, log_blowup, total_cols, warm=True):
    """cfg4 shape, STRONG scaling: `total_cols` columns in all, split over the ranks (column c -> rank c * world / total), blow-up 8,
    extended columns streamed through HBM in chunks (only the roots persist), roots all-gathered through the library's communicator"""
    c0, c1 = shard.column_range(rank, world, total_cols)
    nc = c1 - c0
    n = 1 << log_n
    dev = gen_columns_torch(0x5EED0004, c0, nc, n, "cuda")
    shift = ctx.from_canonical(np.array([[SHIFT, 0, 0, 0]], dtype=np.uint64))[0]

    def step():
        roots = smw.lde_commit(ctx, dev, log_blowup, shift, label0=c0)
        if comm is not None:
            mine = torch.from_numpy(roots.view(np.int64)).cuda()
            return comm.all_gather(mine).cpu().numpy().view(np.uint64)
        return roots
    if warm:
        step()
    secs, roots = T.run(step, 1)
    committed = total_cols * (n << log_blowup)
    rec = {"workload": "cfg4 shape: %d columns in all x 2^%d rows, blow-up %d, column-sharded %d per GPU" % (total_cols, log_n, 1 << log_blowup, nc),
           "scaling": "strong", "ms": secs * 1e3, "committed_elems": committed, "value": committed / secs, "unit": UNIT, "roots_gathered": int(roots.shape[0])}
    del dev
    torch.cuda.empty_cache()
    return rec


def cfg5_leg(ctx, smw, comm, T, rank, world, torch, log_n):
    """cfg5: ONE column of 2^log_n: NTT (cyclic over the ranks -> four-step with one all-to-all) then the FRI commit phase [16,16,8]
    on the resulting chunks (one all-gather)"""
    N = 1 << log_n
    M = N // world
    x_loc = gen_columns_torch(0x5EED0005, 0, 1, N, "cuda")[0][rank::world].contiguous()
    prm = smw.FriProverParams([16, 16, 8], SEED_Z)
    firsts = [M * k1 + rank * (M // world) for k1 in range(world)]
    state = {}

    def ntt():
        state["chunks"] = smw.ntt_cyclic_to_chunks(ctx, comm, x_loc.clone(), log_n)
        ctx.sync()

    def fri():
        return smw.fri_roots_sharded(ctx, comm, state["chunks"], firsts, N, prm)
    ntt(); fri()
    c0 = ctx.collective_count()
    t_ntt, _ = T.run(ntt, 1)
    t_fri, roots = T.run(fri, 1)
    rec = {"workload": "cfg5: one column of 2^%d, NTT + FRI commit phase [16,16,8]" % log_n, "ntt_ms": t_ntt * 1e3, "fri_ms": t_fri * 1e3,
           "ntt_elems_per_s": N / t_ntt, "fri_n0_per_s": N / t_fri, "all_to_all_bytes_per_rank": M * 32 * (world - 1) // world,
           "collectives": ctx.collective_count() - c0, "root0": [int(v) for v in roots[0]]}
    if world > 1:
        # the single-GPU path on the full vector (2 GiB at 2^26: it fits one B200) as the checker of the distributed one: equal layer
        # roots mean an equal transform, every root depends on every transformed element
        del x_loc
        state.clear()
        torch.cuda.empty_cache()
        ok = None
        if rank == 0:
            X = gen_columns_torch(0x5EED0005, 0, 1, N, "cuda")[0].contiguous()
            smw._ntt(ctx, X, 0)
            want = smw.fri_roots_sharded(ctx, None, X.view(1, N, 4), [0], N, prm)
            ok = bool(np.array_equal(want, roots))
            del X
            torch.cuda.empty_cache()
        rec["equal_single_gpu"] = ok              # reported, not asserted: a raise on one rank would leave the others in a collective
    return rec


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--log-n", type=int, default=20)
    ap.add_argument("--cols", type=int, default=64, help="columns per GPU")
    ap.add_argument("--log-blowup", type=int, default=2)
    ap.add_argument("--cpu-log-n", type=int, default=16, help="rows (log2) of the CPU sample: one column of BASELINE configs[0]")
    ap.add_argument("--cpu-cols", type=int, default=2)
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-interp", action="store_true", help="skip the one comparison step on the thread-per-node interpreter")
    ap.add_argument("--no-fri", action="store_true", help="skip the cfg3 / cfg4 / cfg5 sub-records")
    ap.add_argument("--cfg4-log-n", type=int, default=-1, help="rows (log2) of the strong-scaling leg; default 20, and the full 24 at 8 GPUs")
    ap.add_argument("--cfg5-log-n", type=int, default=26)
    args = ap.parse_args()

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if args.impl == "reference":
        run_reference(args, rank, world)
        return

    import torch
    import torch.distributed as dist
    from stark_mlwe_b200 import api as smw, shard
    assert torch.cuda.is_available(), "bench.py needs a CUDA device: the product has no CPU path"
    torch.cuda.set_device(local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        import datetime
        dist.init_process_group("nccl", device_id=torch.device("cuda", local), timeout=datetime.timedelta(seconds=240))
    stream = torch.cuda.Stream()                              # ONE explicit stream for torch's own work (input generation, copies, events) and the
    torch.cuda.set_stream(stream)                             # library's kernels and NCCL calls: everything is ordered, and torch events bracket the kernels
    ctx = smw.Context(local, stream=stream.cuda_stream)
    comm = shard.make_comm(ctx)                               # the library's own NCCL communicator (None at N = 1)
    T = Timer(torch, dist, stream, world)
    n, nc, lb = 1 << args.log_n, args.cols, args.log_blowup
    N = n << lb
    committed_per_rank = nc * N
    col0 = shard.column_range(rank, world, world * nc)[0]

    # ---- inputs: pinned host buffer (e2e leg) and a resident device copy (value leg)
    host = torch.empty((nc, n, 4), dtype=torch.int64).pin_memory()
    gen_columns(SEED, col0, nc, n, out=host.numpy().view(np.uint64))
    dev = host.to("cuda", non_blocking=False)
    shift = ctx.from_canonical(np.array([[SHIFT, 0, 0, 0]], dtype=np.uint64))[0]

    def step(src):
        roots = smw.lde_commit(ctx, src, lb, shift, label0=col0)          # (nc, 4) uint64 on the host; tree_label = global column index
        if comm is not None:                                              # the only collective: 32 B x columns, NCCL inside the library
            return comm.all_gather(torch.from_numpy(roots.view(np.int64)).cuda()).cpu().numpy().view(np.uint64)
        return roots

    for _ in range(args.warmup):
        step(dev)
    # ---- value: inputs resident in HBM (2 GiB of input and 8.6 GB extended per step: far larger than the 126 MB L2)
    sampler = ClockSampler(local) if rank == 0 else None
    l0 = ctx.launch_count()
    ctx.profile(True)
    secs, roots = T.run(lambda: step(dev), args.steps)
    prof = ctx.profile_read()
    ctx.profile(False)
    launches = ctx.launch_count() - l0
    clocks = sampler.stop() if sampler else None
    value = world * committed_per_rank * args.steps / secs

    # ---- the same step on the thread-per-node interpreter (tensor-core path off), once, for the roofline record
    interp = None
    if rank == 0 and world == 1 and not args.no_interp and ctx.tc_node_counts(smw.commitment_default_params(ctx))["fp64_instr"]:
        keep = ctx.tc_node_counts(smw.commitment_default_params(ctx))["min_nodes"]
        ctx.set_tc_min_nodes(0)
        i_secs, roots_i = T.run(lambda: step(dev), 1)
        ctx.set_tc_min_nodes(keep if keep < (1 << 63) else 0)
        assert np.array_equal(roots_i, roots), "tensor-core and interpreter node hashes disagree"
        interp = {"value": world * committed_per_rank / i_secs, "unit": UNIT, "ms_per_step": i_secs * 1e3}

    # ---- e2e: the same call with HOST buffers (pinned input copied H2D, roots D2H) inside the timed region
    e2e = None
    if not args.no_e2e:
        step(host)
        e_secs, roots_e = T.run(lambda: step(host), args.steps)
        assert np.array_equal(roots_e, roots), "host-buffer and device-buffer paths disagree"
        e2e = {"value": world * committed_per_rank * args.steps / e_secs, "unit": UNIT, "h2d_bytes_per_step": int(nc * n * 32 + 32), "d2h_bytes_per_step": int(nc * 32),
               "ms_per_step": e_secs * 1e3 / args.steps}
    del dev, host
    torch.cuda.empty_cache()

    # ---- the other halves of the metric: FRI (cfg3), strong scaling on the cfg4 shape, cfg5
    legs = {}
    if not args.no_fri:
        legs["fri"] = fri_leg(ctx, smw, comm, T, rank, world, torch, 2)
        ln4 = args.cfg4_log_n if args.cfg4_log_n >= 0 else 20
        legs["cfg4"] = strong_leg(ctx, smw, comm, T, rank, world, torch, shard, ln4, 3, 256)
        if world == 8 and args.cfg4_log_n < 0:
            legs["cfg4_full"] = strong_leg(ctx, smw, comm, T, rank, world, torch, shard, 24, 3, 256, warm=False)     # BASELINE configs[3] as stated: 3.4e10 elements, one pass
        legs["cfg5"] = cfg5_leg(ctx, smw, comm, T, rank, world, torch, args.cfg5_log_n)

    if rank == 0:
        out = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": secs * 1e3 / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "Pallas Fr, exact integer arithmetic: u8 limbs x u8 Toeplitz constants -> s32 on the tensor cores for the Poseidon linear layers, 5x52-bit limbs carried in f64 (DFMA.RZ split) for S-boxes and Montgomery reductions, 9x29-bit in the NTT", "data": "synthetic",
            "config": {"workload": "cfg2: 2^%d-row trace, %d columns per GPU, blow-up %d: coset LDE + arity-16 Poseidon-Merkle commit per column" % (args.log_n, nc, 1 << lb),
                       "field": "Pallas Fr (the field the reference uses; SURVEY.md D1)", "columns_total": world * nc, "committed_elems_per_step": world * committed_per_rank,
                       "l2": "inputs (%.1f GB) and extended columns (%.1f GB) exceed the 126 MB L2" % (nc * n * 32 / 1e9, nc * N * 32 / 1e9),
                       "sharding": "columns, %d per GPU; roots all-gathered by the library's NCCL communicator" % nc,
                       "vs_baseline_note": "BASELINE.json publishes no number for this metric (published: {})"},
            "e2e": e2e, "gpu_launches": int(launches), "clocks": clocks, "roofline": roofline_record(ctx, smw, prof, args, nc, N, interp),
        }
        out.update(legs)
        if not args.no_cpu and world == 1:
            out["cpu_baseline"] = cpu_baseline(args, ctx, smw)
        print(json.dumps(out))
    T.barrier()
    if comm is not None:
        comm.close()
    ctx.close()
    if world > 1:
        dist.destroy_process_group()


def peak_int8_tmacs():
    """dense INT8 multiply-accumulates per second of the tensor cores: twice the bf16 rate measured on this pool (MEASURED_PEAKS.json,
    sustained figure: the stages run inside a long step), else twice the fallback of B200_PROFILING.md; 1 MAC = 2 ops"""
    try:
        return float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["bf16_tflops_sustained"]) * 2 / 2, "2 x MEASURED_PEAKS.json bf16_tflops_sustained / 2 ops per MAC"
    except Exception:
        return 1650.0, "2 x the bf16 fallback of B200_PROFILING.md / 2 ops per MAC"


def tc_split(ctx, smw, nc, N):
    """nodes of one step that the tensor-core node hash takes (full arity-16 levels of >= min_nodes nodes per launch) and the rest"""
    cnt = ctx.tc_node_counts(smw.commitment_default_params(ctx))
    tc_nodes, other_perms, cn = 0, 0, N
    while cn > 1:
        nn = (cn + 15) // 16
        if cnt["fp64_instr"] and cn == nn * 16 and nn * nc >= cnt["min_nodes"]:
            tc_nodes += nn * nc
        else:
            full, rem = cn // 16, cn % 16
            other_perms += nc * (full * 2 + (((4 + rem + 1 + 15) // 16) if rem else 0))
        cn = nn
    return cnt, tc_nodes, other_perms


def roofline_record(ctx, smw, prof, args, nc, N, interp=None):
    """Dominant kernels: the staged tensor-core node hash (k_tc_stage<*>, k_tc_chain, k_tc_head; csrc/poseidon_tc.cuh) when the levels are
    large enough, else k_merkle_level52<17>.  Two pipes carry the work, so two fractions are reported, both from counts of what the
    kernels EXECUTE (never the reference's dense count): `frac` = FP64-pipe instructions (S-boxes, Montgomery reductions) / kernel time /
    the DFMA rate measured live in this run; `tensor.frac` = int8 MACs of the linear layers / kernel time / the dense INT8 rate.  The
    epilogues are bound by instruction issue across the ALU and FP64 pipes (profiles/r04a_ncu_tc_stage_summary.txt), so neither reaches 1."""
    ml_ms, ml_launches = prof["merkle_level"]
    perms = nc * tree_perms(N, 16, 16)
    ml_s = ml_ms * 1e-3 / args.steps
    pdef = smw.commitment_default_params(ctx)
    fp_perm, mac_perm = pdef.executed_fp64_ops(), pdef.executed_macs()
    fp, mac = ctx.pipe_rate(1), ctx.pipe_rate(0)
    cnt, tc_nodes, other_perms = tc_split(ctx, smw, nc, N)
    alg_bytes = nc * (N * 32 + sum_levels_bytes(N))
    name = "k_merkle_level52<17> (+ k_nodes_coop52<17> for the tree tops)"
    if tc_nodes:
        name = "tensor-core node hash: k_tc_stage<0..3> + k_tc_chain + k_tc_head (%.1f %% of the nodes; the rest k_merkle_level52<17> / k_nodes_coop52<17>)" % (100.0 * 2 * tc_nodes / perms)
    rec = {"kernel": name, "kernel_ms_per_step": ml_ms / args.steps, "kernel_launches_per_step": ml_launches / args.steps,
           "permutations_per_step": perms, "permutations_per_s": perms / ml_s,
           "hbm": {"achieved": alg_bytes / ml_s / 1e9, "peak": peak_hbm(), "unit": "GB/s", "frac": alg_bytes / ml_s / 1e9 / peak_hbm()},
           "traffic": (TRAFFIC_NCU_TC if tc_nodes else TRAFFIC_NCU)["bytes"], "traffic_source": TRAFFIC_NCU_TC if tc_nodes else TRAFFIC_NCU,
           "step_share_ms": {k: v[0] / args.steps for k, v in prof.items() if v[1]}}
    if fp_perm:
        fp_exec = tc_nodes * cnt["fp64_instr"] + other_perms * fp_perm
        ach = fp_exec / ml_s
        rec.update({"bound": "fp64", "unit": "T fp64 lane-instr/s (DFMA.RZ + DADD)", "achieved": ach / 1e12, "peak": fp["lanes_per_s"] / 1e12, "frac": ach / fp["lanes_per_s"],
                    "fp64_ops_per_permutation": fp_exec / perms, "fp64_ops_per_permutation_interpreter": fp_perm,
                    "peak_measurement": dict(fp, how="16 independent DFMA.RZ chains x 512 threads/SM, ~40 ms per repetition, best of 3 (smw_diag_pipe_rate); "
                                                     "lanes_per_s = warp_instr_per_clk_per_sm x 32 x n_sm x sm_hz_during, sm_hz_during = in-kernel clock64 ticks / event time"),
                    "peak_at_max_clock": fp["warp_instr_per_clk_per_sm"] * 32 * fp["n_sm"] * fp["sm_max_hz"] / 1e12,
                    "frac_at_max_clock": ach / (fp["warp_instr_per_clk_per_sm"] * 32 * fp["n_sm"] * fp["sm_max_hz"])})
        if tc_nodes:
            pk, how = peak_int8_tmacs()
            tm = tc_nodes * cnt["int8_mac"] / ml_s / 1e12
            rec["tensor"] = {"bound": "tensor", "unit": "T int8 MAC/s (tcgen05.mma kind::i8, Toeplitz zeros included)", "achieved": tm, "peak": pk, "frac": tm / pk, "peak_source": how,
                             "int8_mac_per_node": cnt["int8_mac"], "word_products_per_node": cnt["word_products"], "group": cnt["group"], "launches_per_batch": cnt["launches_per_batch"]}
        if interp:
            rec["interpreter_same_run"] = dict(interp, note="the same step with the tensor-core path switched off (smw_ctx_set_tc_min_nodes(0)): thread-per-node FP64 interpreter, "
                                                            "%d FP64 instructions per permutation" % fp_perm)
    else:
        ach = perms * mac_perm / ml_s
        rec.update({"bound": "imad", "unit": "T mac32/s", "achieved": ach / 1e12, "peak": mac["lanes_per_s"] / 1e12, "frac": ach / mac["lanes_per_s"], "mac_per_permutation": mac_perm})
    alg_mac = perms * MM_PERM_T17 * MAC_PER_MM
    rec["algorithmic"] = {"mm_per_permutation_dense": MM_PERM_T17, "mac32_per_mm": MAC_PER_MM, "achieved_T_mac32_per_s": alg_mac / ml_s / 1e12,
                          "imad_wide_peak_T_mac32_per_s": mac["lanes_per_s"] / 1e12, "imad_peak_measurement": mac,
                          "algorithmic_speedup": alg_mac / ml_s / mac["lanes_per_s"],
                          "note": "dense reference count (SURVEY.md 8d) / time / measured IMAD.WIDE rate: how many times faster than a dense 8x32-bit CIOS kernel running at "
                                  "the integer-multiply peak; > 1 because the kernel executes ~5x fewer multiplications (LU mat-vec, sparse partial rounds, lazy reduction) on another pipe"}
    return rec


def cpu_baseline(args, ctx, smw):
    """the oracle (C port of the reference) on the host cores, on a bounded sample of the same workload; the GPU computes the
    same sample and the roots must agree bit for bit"""
    from oracle import oracle as O
    O.lib()
    step, cols, committed, sample = cpu_step_fn(O, args)
    shift = O.to_mont([SHIFT])[0]
    res = {}
    roots = None
    for th in sorted({1, cpu_threads()}):
        O.set_threads(th)
        t0 = time.perf_counter()
        roots = step()
        res[th] = time.perf_counter() - t0
    got = smw.lde_commit(ctx, cols, args.log_blowup, shift, label0=0)
    assert np.array_equal(got, np.stack(roots)), "GPU roots differ from the CPU oracle on the sample"
    th = max(res)
    return {"value": committed / res[th], "unit": UNIT, "cores": th, "kind": "port",
            "sample": sample + "; GPU roots on the same sample: bit-exact", "single_thread_value": committed / res[1]}


if __name__ == "__main__":
    main()
