// poseidon.cuh — device-side Poseidon permutation and sponges over fr29.
//
// Mapping to the SM: ONE THREAD PER PERMUTATION, state resident in shared memory in a
// word-interleaved layout ([lane][chunk][thread], conflict-free LDS.128), constants read as
// warp-uniform loads from one global table (L1/L2 resident, 48-byte aligned elements).
//
// The permutation is not written out as code: it is a PROGRAM of "row operations" built once per
// parameter set on the host (poseidon_prog.cu) and run by one small interpreter loop (run_ops below).
// A row operation is
//        x = sum_{j < n} A_j * B_j            lazily accumulated (fr29.cuh), ONE Montgomery reduction
//        x += S[dst]   (optional)             unit diagonal of L / rank-1 update of a sparse partial round
//        x += rc       (optional)             the NEXT round's round constant, folded into the producer
//        S[dst] = x
// with A_j a state lane (or an external input element) and B_j a table entry (or a state lane, for the
// S-box multiplications).  Why an interpreter: round 1a inlined the rounds as straight-line code, 360 KB of
// SASS for the Merkle kernel; with ten warps per SM in different phases the instruction cache thrashed and
// "no instruction" was the top stall (profiles/r01a_ncu_full_merkle_level17_details.csv: 6.2 of 10.6 cycles per
// issue).  Here every warp executes the same ~8 KB loop whose body is one 81-multiply block, so the hot
// code stays in the L0/L1 instruction caches whatever the phase of each warp.
//
// The dense t x t matrix is applied IN PLACE through its LU factors (same multiply count as the dense
// product, no second state buffer); the RP partial rounds use the exact sparse factorisation (one row
// dot product + rank-1 update per round, one dense (t-1)x(t-1) matrix at the end).  These are exact
// identities over the field, so outputs are bit-identical to the reference's dense rounds.
//
// Reference semantics: poseidon/src/lib.rs:219-258 permute_dynamic (== :31-68 permute for t = 17),
// :260-272 dense mat-vec in every round, :276-312 absorb_one / hash_with_ds_dynamic.
#pragma once
#include "smw_internal.h"
#include "fr29.cuh"

namespace smw {

constexpr int EL = 12;   // table element stride in 32-bit words (9 used; 48 B keeps 16-byte alignment)

// row-operation flags (bits 24..31 of Op.y)
enum : uint32_t {
    OPF_BSTATE  = 1,    // B operand is state lane Op.x (S-box multiplications) instead of table element Op.x
    OPF_ADDSELF = 2,    // add the old S[dst] to the reduced result
    OPF_ADDRC   = 4,    // add table element Op.z (round constant / padding one)
    OPF_LOAD    = 8,    // no products: x = S[Op.w]
    OPF_SQR     = 16,   // x = S[j0]^2 (45 multiplies instead of 81)
    OPF_AEXT    = 32    // A operand is external input element Op.w (raw 29-bit repack; the B operand carries the
                        // conversion constant), single term
};
// Op = uint4 { x: B offset (table element index, or state lane), y: j0 | n << 8 | dst << 16 | flags << 24,
//              z: table element index of the constant to add, w: source lane (LOAD) / input index (AEXT) / A-lane stride (else) }
__host__ __device__ inline uint4 make_op(uint32_t b_off, int j0, int n, int dst, uint32_t flags, uint32_t rc_off = 0, uint32_t aux = 0) {
    if (!(flags & (OPF_LOAD | OPF_AEXT)) && aux == 0) aux = 1;          // A-lane stride defaults to 1
    uint4 o; o.x = b_off; o.y = (uint32_t)j0 | ((uint32_t)n << 8) | ((uint32_t)dst << 16) | (flags << 24); o.z = rc_off; o.w = aux; return o;
}

// Segs (smw_internal.h): up to 8 (first, count) ranges of the op array, run in order, e.g.
// [absorb block 0][permutation][absorb block 1][permutation]

// Device view of one parameter set
struct PoseidonDev {
    int t, rp;
    const uint32_t* tbl;     // every table in internal form (x * 2^261 mod p, canonical limbs), EL words per element
    const uint4* prog;       // row operations
    int perm_first, perm_count;       // the production permutation (sparse partial rounds)
    int dense_first, dense_count;     // the literal round structure (dense matrix in every partial round): test aid
    uint32_t k_in, k_out, k_rr, k_one;   // table indices of 2^266, 2^256, 2^522, 2^261 (mod p)
};

SMW_DEV F29 tbl_load(const uint32_t* base, uint32_t idx) {
    const uint4* q = reinterpret_cast<const uint4*>(base + (size_t)idx * EL);
    uint4 a = __ldg(q), b = __ldg(q + 1), c = __ldg(q + 2);
    F29 r;
    r.v[0] = a.x; r.v[1] = a.y; r.v[2] = a.z; r.v[3] = a.w; r.v[4] = b.x; r.v[5] = b.y; r.v[6] = b.z; r.v[7] = b.w; r.v[8] = c.x;
    return r;
}

// Per-thread state in shared memory: lane j = two uint4 (limbs 0-3, 4-7) + one word (limb 8); T + 2 lanes (lanes T, T + 1 are scratch)
struct StateRef {
    uint4* q;       // [T+2][2][NT]
    uint32_t* w;    // [T+2][NT]
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
    StateRef S; S.q = reinterpret_cast<uint4*>(smem); S.w = reinterpret_cast<uint32_t*>(smem + (size_t)(T + 2) * 2 * nt * 16); S.nt = nt; S.tid = tid; return S;
}
template <int T> __host__ __device__ constexpr size_t state_smem_bytes(int nt) { return (size_t)(T + 2) * nt * 36; }

SMW_DEV void load_ext(const uint32_t* p, uint32_t w[8]) {
    const uint4* q = reinterpret_cast<const uint4*>(p);
    uint4 a = q[0], b = q[1];
    w[0] = a.x; w[1] = a.y; w[2] = a.z; w[3] = a.w; w[4] = b.x; w[5] = b.y; w[6] = b.z; w[7] = b.w;
}
SMW_DEV void store_ext(uint32_t* p, const uint32_t w[8]) {
    uint4* q = reinterpret_cast<uint4*>(p);
    q[0] = make_uint4(w[0], w[1], w[2], w[3]); q[1] = make_uint4(w[4], w[5], w[6], w[7]);
}

// "no external inputs" functor for programs without OPF_AEXT operations
struct NoInput { SMW_DEV F29 operator()(uint32_t) const { return f29_zero(); } };

// The interpreter.  One copy per kernel (noinline: the sponges call it once per segment list), about a thousand
// instructions; all warps of the SM loop over the same code.  The state lives in the CTA's dynamic shared memory and is
// re-derived from `smem` here (not passed as pointers) so that the compiler keeps LDS/STS addressing.
//   * the term loop is unrolled by two with ping-pong operand registers: the next term's operands are in flight during
//     the current term's 81 multiplies and no register moves are needed;
//   * the next operation's descriptor and its first table operand are fetched one operation ahead (table loads mostly
//     miss the small L1 left beside 207 KB of shared memory, and go to L2).
template <int T, class In>
__device__ __noinline__ void run_ops(int nt, const uint32_t* __restrict__ tbl, const uint4* __restrict__ prog, int first, int count, In in) {
    extern __shared__ __align__(16) unsigned char smem[];
    const StateRef S = state_ref<T>(smem, nt, threadIdx.x);
    uint4 d = __ldg(prog + first);
    F29 bpre = ((d.y >> 24) & (OPF_BSTATE | OPF_LOAD | OPF_SQR)) ? f29_zero() : tbl_load(tbl, d.x);
#pragma unroll 1
    for (int k = 0; k < count; k++) {
        const uint4 dn = __ldg(prog + first + k + 1);        // every program is followed by one padding op
        const uint32_t fl = d.y >> 24;
        const int j0 = d.y & 255, n = (d.y >> 8) & 255, dst = (d.y >> 16) & 255;
        F29 x;
        if (fl & OPF_LOAD) {
            x = S.load((int)d.w);
        } else {
            Acc A; acc_zero(A);
            if (fl & OPF_SQR) {
                acc_sqr(A, S.load(j0));
            } else {
                F29 a0 = (fl & OPF_AEXT) ? in(d.w) : S.load(j0);
                F29 b0 = (fl & OPF_BSTATE) ? S.load((int)d.x) : bpre;
                F29 a1 = a0, b1 = b0;
                const int st = (fl & OPF_AEXT) ? 1 : (int)d.w;              // A-lane stride (1 except the two-S-box lane updates)
                int c = 0, j = 1;
#pragma unroll 1
                while (true) {
                    if (j < n) { a1 = S.load(j0 + j * st); b1 = tbl_load(tbl, d.x + j); }
                    if (c == 6) { acc_normalize(A); c = 0; }
                    acc_mac(A, a0, b0); c++;
                    if (j >= n) break;
                    j++;
                    if (j < n) { a0 = S.load(j0 + j * st); b0 = tbl_load(tbl, d.x + j); }
                    if (c == 6) { acc_normalize(A); c = 0; }
                    acc_mac(A, a1, b1); c++;
                    if (j >= n) break;
                    j++;
                }
            }
            if (!((dn.y >> 24) & (OPF_BSTATE | OPF_LOAD | OPF_SQR))) bpre = tbl_load(tbl, dn.x);   // in flight during the reduction
            F29 s = f29_zero();                                   // one copy of the reduction for both cases
            if (fl & OPF_ADDSELF) s = S.load(dst);
#pragma unroll
            for (int i = 0; i < 8; i++) A.c[9 + i] += s.v[i];
            x = acc_reduce(A);
            x.v[8] += s.v[8];
        }
        if (fl & OPF_ADDRC) x = f29_add(x, tbl_load(tbl, d.z));
        S.store(dst, x);
        d = dn;
    }
}
template <int T, class In>
SMW_DEV void run_segs(int nt, const PoseidonDev& P, const Segs& sg, const In& in) {
#pragma unroll 1
    for (int s = 0; s < sg.n; s++) run_ops<T>(nt, P.tbl, P.prog, sg.first[s], sg.count[s], in);
}
template <int T>
SMW_DEV void run_perm(int nt, const PoseidonDev& P, bool sparse = true) {
    run_ops<T>(nt, P.tbl, P.prog, sparse ? P.perm_first : P.dense_first, sparse ? P.perm_count : P.dense_count, NoInput());
}

} // namespace smw
