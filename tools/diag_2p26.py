"""Diagnostic: determinism and correctness of the single-GPU NTT + FRI at 2^25 / 2^26 (the cfg5 sizes)."""
import json, os, sys, time
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from stark_mlwe_b200 import api as smw
from bench import gen_columns_torch
from oracle import oracle as O
O.lib(); O.set_threads(1)
_st = torch.cuda.Stream(); torch.cuda.set_stream(_st)
ctx = smw.Context(0, stream=_st.cuda_stream)       # torch's work and the library's share one explicit stream
out = {}
for log_n in [int(a) for a in sys.argv[1:]] or [24, 25, 26]:
    N = 1 << log_n
    x = gen_columns_torch(0x5EED0005, 0, 1, N, "cuda")[0].contiguous()
    X1 = smw._ntt(ctx, x.clone(), 0); ctx.sync()
    X2 = smw._ntt(ctx, x.clone(), 0); ctx.sync()
    rec = {"ntt_deterministic": bool(torch.equal(X1, X2))}
    xh = x.cpu().numpy().view(np.uint64)
    w = O.root_of_unity(log_n)
    pts = [0, 1, N - 1, 12345677 % N, (N // 2) + 3]
    X1h = X1[pts].cpu().numpy().view(np.uint64)
    rec["ntt_horner_ok"] = [bool(np.array_equal(X1h[k], O.horner(xh, O.fr_pow(w, i)))) for k, i in enumerate(pts)]
    back = smw._ntt(ctx, X1.clone(), 1); ctx.sync()
    rec["inverse_round_trip"] = bool(torch.equal(back, x))
    del back, X2
    prm = smw.FriProverParams([16, 16, 8], 0xDEEFBAAD)
    r1 = smw.fri_roots_sharded(ctx, None, X1.view(1, N, 4), [0], N, prm)
    r2 = smw.fri_roots_sharded(ctx, None, X1.view(1, N, 4), [0], N, prm)
    r3 = smw.fri_roots_sharded(ctx, None, X1.view(4, N // 4, 4), [k * (N // 4) for k in range(4)], N, prm)
    st = smw.fri_build_transcript(ctx, X1, prm)
    r4 = st.roots()
    rec["fri_deterministic"] = bool(np.array_equal(r1, r2))
    rec["fri_4_slices_equal"] = bool(np.array_equal(r1, r3))
    rec["fri_build_transcript_equal"] = bool(np.array_equal(r1, r4))
    rec["root0"] = [int(v) for v in r1[0]]
    # layer 0 spot check: leaf digests of layer 0 against the oracle's hash_leaf_pair on a few indices, through the tree's level 0
    f0 = X1[:64].cpu().numpy().view(np.uint64)
    f1 = st.layers[1].f[:4]
    lv0 = st.layers[0].tree.level(0)[:64]
    s0 = np.repeat(f1, 16, axis=0)
    rec["leaf_digests_ok"] = bool(np.array_equal(lv0, O.hash_leaf_pair(f0, s0)))
    del st, X1, x
    torch.cuda.empty_cache()
    out[str(log_n)] = rec
    print(json.dumps({str(log_n): rec}), flush=True)
