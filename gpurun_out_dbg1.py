import sys; sys.path.insert(0, "/root/repo")
import numpy as np
from oracle import oracle as O
from stark_mlwe_b200 import api as smw
ctx = smw.Context(0)
for t in (9, 17):
    g, o = smw.poseidon_params_for_width(ctx, t), O.Params.for_width(t)
    st = O.StdRng.seed_from_u64(5).fr_rand(3*t).reshape(3, t, 4)
    ref = O.permute(o, st)
    for dense in (True, False):
        got = smw.permute(ctx, st, g, dense=dense)
        print("t", t, "dense" if dense else "sparse", "match:", np.array_equal(got, ref), "lanes ok:", [bool(np.array_equal(got[0,i], ref[0,i])) for i in range(t)])
