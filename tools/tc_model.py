"""Python-integer model of the tensor-core formulation of the Poseidon permutation (csrc/poseidon_tc.cuh), checked against the
oracle's permutation.  Test infrastructure (imports oracle/): run `python tools/tc_model.py`.
 1. round structure: every linear layer is a word-level matrix product; partial rounds in groups of G, only S-box chains sequential;
 2. byte GEMM with REDUCED constants: column c of output word i = sum_j sum_a byte_a(x_j) * byte_c((W_ij * 2^(8a)) mod p), 32 columns per
    word, value < 2^274, congruent to (W x)_i mod p; cheap reduction: y = low254 + p - top * (p - 2^254)  in [0, 2^255)."""
import sys
import os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from oracle import oracle as O

P = 0x40000000000000000000000000000000224698fc0994a8dd8c46eb2100000001
C = P - (1 << 254)


def matvec(M, v): return [sum(a * b for a, b in zip(row, v)) % P for row in M]
def matmul(A, B): return [[sum(A[i][k] * B[k][j] for k in range(len(B))) % P for j in range(len(B[0]))] for i in range(len(A))]
def sbox(x): return pow(x, 5, P)


def byte_gemm(W, x):
    """x: words < 2^256 (any representatives); returns lazily reduced representatives < 2^255 of W x"""
    out = []
    for row in W:
        cols = [0] * 32
        for wij, xj in zip(row, x):
            assert 0 <= xj < (1 << 256)
            for a in range(32):
                xa = (xj >> (8 * a)) & 255
                e = (wij << (8 * a)) % P
                for c in range(32):
                    cols[c] += xa * ((e >> (8 * c)) & 255)
        assert max(cols) < (1 << 26)
        y = sum(v << (8 * c) for c, v in enumerate(cols))
        top, low = y >> 254, y & ((1 << 254) - 1)
        assert top < (1 << 21)
        r = low + P - top * C
        assert 0 <= r < (1 << 255) and (r - sum(w * v for w, v in zip(row, x))) % P == 0
        out.append(r)
    return out


def model(params, state, G, gemm=matvec):
    mds, rcf, rcp = params.export()
    t, rp = params.t, params.rp
    M = [[O.from_mont(mds[i, j])[0] for j in range(t)] for i in range(t)]
    RF = [[O.from_mont(rcf[r, j])[0] for j in range(t)] for r in range(8)]
    RP = [O.from_mont(rcp[k])[0] for k in range(rp)]
    assert rp % G == 0
    Mh = [[0 if j == 0 else M[i][j] for j in range(t)] for i in range(t)]
    m0 = [M[i][0] for i in range(t)]
    pw = [[[int(i == j) for j in range(t)] for i in range(t)]]
    for d in range(G): pw.append(matmul(Mh, pw[-1]))
    rows = [pw[d][0] for d in range(G)]                           # e0^T Mh^d, d < G
    cols = [matvec(pw[d], m0) for d in range(G)]                  # Mh^d m0
    gamma = [cols[d][0] for d in range(G)]                        # e0^T Mh^d m0
    Wv = [pw[G][i] + [cols[G - 1 - m][i] for m in range(G)] for i in range(t)]               # t x (t + G)
    Wd = [[sum(rows[k][i] * Wv[i][c] for i in range(t)) % P for c in range(t + G)] for k in range(G)]
    D0 = [[sum(rows[k][i] * M[i][j] for i in range(t)) % P for j in range(t)] for k in range(G)]
    v = list(state)
    x = [sbox((v[i] + RF[0][i]) % P) for i in range(t)]
    for r in range(3):
        v = [(a + b) % P for a, b in zip(gemm(M, x), RF[r + 1])]
        x = [sbox(a) for a in v]
    vd = gemm(M + D0, x)
    v, dots = vd[:t], vd[t:]
    for g in range(rp // G):
        xs = []
        for k in range(G):
            s = (dots[k] + sum(gamma[k - 1 - m] * xs[m] for m in range(k)) + RP[g * G + k]) % P
            xs.append(sbox(s))
        vd = gemm(Wv + Wd, v + xs)
        v, dots = vd[:t], vd[t:]
    x = [sbox((v[i] + RF[4][i]) % P) for i in range(t)]
    for r in range(4, 7):
        v = [(a + b) % P for a, b in zip(gemm(M, x), RF[r + 1])]
        x = [sbox(a) for a in v]
    return [a % P for a in gemm(M, x)]


if __name__ == "__main__":
    ok = True
    for t, G, gemm in ((17, 8, matvec), (17, 4, matvec), (9, 6, matvec), (17, 8, byte_gemm)):
        prm = O.Params.for_width(t)
        st = O.StdRng.seed_from_u64(5 + t).fr_rand(t)
        want = O.from_mont(O.permute(prm, st.reshape(1, t, 4))[0])
        got = model(prm, O.from_mont(st), G, gemm)
        print(t, prm.rp, G, gemm.__name__, want == got)
        ok = ok and want == got
    sys.exit(0 if ok else 1)
