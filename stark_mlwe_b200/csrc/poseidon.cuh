// poseidon.cuh — device-side Poseidon permutation and sponges over fr29.
//
// Mapping to the SM: ONE THREAD PER PERMUTATION, state resident in shared memory in a
// word-interleaved layout ([element][chunk][thread], conflict-free LDS.128), constants read as
// warp-uniform loads from a global table (L1/L2 resident, 48-byte aligned elements).  The dense
// t x t matrix is applied IN PLACE through its LU factors (same multiply count as the dense product,
// no second state buffer, so 2x the resident threads per SM); the RP partial rounds use the exact
// sparse factorisation (row + rank-1 update per round, one dense (t-1)x(t-1) matrix at the end).
// Every dot product is accumulated lazily (fr29.cuh) and reduced once.
//
// Reference semantics: poseidon/src/lib.rs:219-258 permute_dynamic (== :31-68 permute for t = 17),
// :260-272 dense mat-vec in every round, :276-312 absorb_one / hash_with_ds_dynamic.
// The factorisations are exact identities over the field, so outputs are bit-identical.
#pragma once
#include "fr29.cuh"

namespace smw {

constexpr int EL = 12;   // table element stride in 32-bit words (9 used; 48 B keeps 16-byte alignment)

// Device view of one parameter set (all tables in internal form: x * 2^261 mod p, canonical limbs)
struct PoseidonDev {
    int t, rp;
    const uint32_t* lu_full;   // [t][t]      : U on and above the diagonal, unit-L strictly below   (M = L*U)
    const uint32_t* rc_full;   // [8][t]
    const uint32_t* rc_part;   // [rp]
    const uint32_t* sp_row;    // [rp][t]     : (M00, v^T * Mh^k)            -> new lane 0
    const uint32_t* sp_col;    // [rp][t-1]   : Mh^-(k+1) * w                -> rank-1 update of lanes 1..
    const uint32_t* lu_post;   // [t-1][t-1]  : LU of Mh^rp, applied to lanes 1.. after the partial rounds
};

SMW_DEV F29 tbl_load(const uint32_t* base, int idx) {
    const uint4* q = reinterpret_cast<const uint4*>(base + (size_t)idx * EL);
    uint4 a = __ldg(q), b = __ldg(q + 1), c = __ldg(q + 2);
    F29 r;
    r.v[0] = a.x; r.v[1] = a.y; r.v[2] = a.z; r.v[3] = a.w; r.v[4] = b.x; r.v[5] = b.y; r.v[6] = b.z; r.v[7] = b.w; r.v[8] = c.x;
    return r;
}

// Per-thread state in shared memory: element j = two uint4 (limbs 0-3, 4-7) + one word (limb 8)
struct StateRef {
    uint4* q;       // [T][2][NT]
    uint32_t* w;    // [T][NT]
    int nt, tid;
    SMW_DEV F29 load(int j) const {
        uint4 a = q[(2 * j) * nt + tid], b = q[(2 * j + 1) * nt + tid];
        F29 r;
        r.v[0] = a.x; r.v[1] = a.y; r.v[2] = a.z; r.v[3] = a.w; r.v[4] = b.x; r.v[5] = b.y; r.v[6] = b.z; r.v[7] = b.w; r.v[8] = w[j * nt + tid];
        return r;
    }
    SMW_DEV void store(int j, const F29& x) const {
        q[(2 * j) * nt + tid] = make_uint4(x.v[0], x.v[1], x.v[2], x.v[3]);
        q[(2 * j + 1) * nt + tid] = make_uint4(x.v[4], x.v[5], x.v[6], x.v[7]);
        w[j * nt + tid] = x.v[8];
    }
};
template <int T> SMW_DEV StateRef state_ref(unsigned char* smem, int nt, int tid) {
    StateRef S; S.q = reinterpret_cast<uint4*>(smem); S.w = reinterpret_cast<uint32_t*>(smem + (size_t)T * 2 * nt * 16); S.nt = nt; S.tid = tid; return S;
}
template <int T> __host__ __device__ constexpr size_t state_smem_bytes(int nt) { return (size_t)T * nt * 36; }

// s[off .. off+n) <- (L*U) * s[off .. off+n), in place.  lu is an n x n table.
SMW_DEV void matvec_lu_inplace(const StateRef& S, const uint32_t* lu, int n, int off) {
    // U sweep, ascending rows: row i reads lanes >= i only
#pragma unroll 1
    for (int i = 0; i < n; i++) {
        Acc A; acc_zero(A); int c = 0;
        const uint32_t* row = lu + (size_t)i * n * EL;
#pragma unroll 1
        for (int j = i; j < n; j++) {
            acc_mac(A, S.load(off + j), tbl_load(row, j));
            if (++c == 6) { acc_normalize(A); c = 0; }
        }
        S.store(off + i, acc_reduce(A));
    }
    // L sweep (unit diagonal), descending rows: row i reads lanes <= i only
#pragma unroll 1
    for (int i = n - 1; i >= 1; i--) {
        Acc A; acc_zero(A); int c = 0;
        const uint32_t* row = lu + (size_t)i * n * EL;
#pragma unroll 1
        for (int j = 0; j < i; j++) {
            acc_mac(A, S.load(off + j), tbl_load(row, j));
            if (++c == 6) { acc_normalize(A); c = 0; }
        }
        S.store(off + i, acc_reduce_add(A, S.load(off + i)));
    }
}

template <int T> SMW_DEV void full_round(const StateRef& S, const PoseidonDev& P, int r) {
#pragma unroll 1
    for (int i = 0; i < T; i++) S.store(i, f29_sbox5(f29_add(S.load(i), tbl_load(P.rc_full, r * T + i))));
    matvec_lu_inplace(S, P.lu_full, T, 0);
}

// The permutation.  SPARSE = false keeps the reference's round structure literally (dense matrix in
// every partial round); SPARSE = true is the production path.  Both are tested against the oracle.
template <int T, bool SPARSE> SMW_DEV void permute(const StateRef& S, const PoseidonDev& P) {
#pragma unroll 1
    for (int r = 0; r < 4; r++) full_round<T>(S, P, r);
    if (SPARSE) {
#pragma unroll 1
        for (int k = 0; k < P.rp; k++) {
            F29 x0 = f29_sbox5(f29_add(S.load(0), tbl_load(P.rc_part, k)));
            const uint32_t* row = P.sp_row + (size_t)k * T * EL;
            const uint32_t* col = P.sp_col + (size_t)k * (T - 1) * EL;
            Acc A; acc_zero(A); acc_mac(A, x0, tbl_load(row, 0)); int c = 1;
#pragma unroll 1
            for (int i = 1; i < T; i++) {
                F29 si = S.load(i);
                acc_mac(A, si, tbl_load(row, i));
                if (++c == 6) { acc_normalize(A); c = 0; }
                Acc B; acc_zero(B); acc_mac(B, x0, tbl_load(col, i - 1));
                S.store(i, acc_reduce_add(B, si));
            }
            S.store(0, acc_reduce(A));
        }
        matvec_lu_inplace(S, P.lu_post, T - 1, 1);
    } else {
#pragma unroll 1
        for (int k = 0; k < P.rp; k++) {
            S.store(0, f29_sbox5(f29_add(S.load(0), tbl_load(P.rc_part, k))));
            matvec_lu_inplace(S, P.lu_full, T, 0);
        }
    }
#pragma unroll 1
    for (int r = 4; r < 8; r++) full_round<T>(S, P, r);
}

// DS sponge (poseidon/src/lib.rs:276-312): rate T-1, additive absorb, pad with 1 then zeros.
template <int T, bool SPARSE> struct Sponge {
    StateRef S; const PoseidonDev& P; int cursor;
    SMW_DEV Sponge(const StateRef& s, const PoseidonDev& p) : S(s), P(p), cursor(0) {
#pragma unroll 1
        for (int i = 0; i < T; i++) S.store(i, f29_zero());
    }
    SMW_DEV void absorb(const F29& x) {
        S.store(cursor, f29_add(S.load(cursor), x));
        if (++cursor == T - 1) { cursor = 0; permute<T, SPARSE>(S, P); }
    }
    SMW_DEV F29 finish() {
        absorb(f29_const(C_ONE29));
        if (cursor != 0) { cursor = 0; permute<T, SPARSE>(S, P); }   // absorbing zeros only advances the cursor
        return S.load(0);
    }
};

SMW_DEV void load_ext(const uint32_t* p, uint32_t w[8]) {
    const uint4* q = reinterpret_cast<const uint4*>(p);
    uint4 a = q[0], b = q[1];
    w[0] = a.x; w[1] = a.y; w[2] = a.z; w[3] = a.w; w[4] = b.x; w[5] = b.y; w[6] = b.z; w[7] = b.w;
}
SMW_DEV void store_ext(uint32_t* p, const uint32_t w[8]) {
    uint4* q = reinterpret_cast<uint4*>(p);
    q[0] = make_uint4(w[0], w[1], w[2], w[3]); q[1] = make_uint4(w[4], w[5], w[6], w[7]);
}

} // namespace smw
