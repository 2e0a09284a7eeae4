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
// with A_j state lanes j0, j0+1, .. (or an external input element) and B_j consecutive table entries (or a
// state lane, for the S-box multiplications).
//
// Why an interpreter: round 1a inlined the rounds as straight-line code, 360 KB of SASS for the Merkle kernel; with
// ten warps per SM in different phases the instruction cache thrashed (profiles/r01a_*).  Why a SOFTWARE-PIPELINED
// interpreter (v3): the smem-resident state limits an SM to ten warps (2.5 per scheduler), and in v2 a warp spent more
// than half of its instructions outside the multiply block (Montgomery reduction, carry chains, stores: 58 % of the
// executed SASS, profiles/r01h_ncu_merkle_level17_stalls_pipes_mix.json), during which it issued nothing to the
// multiplier; the fmaheavy pipe idled 34 % of the time with "wait" as the top stall.  Here the reduction + store of
// operation k and the first 81 multiplies of operation k+1 sit in ONE basic block on two accumulator sets (A/B,
// alternating roles in a loop unrolled by two, so no register moves), and ptxas interleaves the dependent ALU chain of
// the reduction with the independent IMAD.WIDE stream.  An operation whose first-term operands are the result of its
// predecessor (OPF_DEP, set by mark_deps() in poseidon_prog.cu) cannot overlap and takes the plain path.
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
    OPF_BSTATE  = 1,    // B operand of the (single) term is state lane Op.x (S-box multiplications) instead of a table element
    OPF_ADDSELF = 2,    // add the old S[dst] to the reduced result
    OPF_ADDRC   = 4,    // add table element Op.z (round constant / padding one)
    OPF_DEP     = 8,    // the first term reads the lane its predecessor writes: no overlap (set by mark_deps, never by builders)
    OPF_SQR     = 16,   // x = S[j0]^2 (always with OPF_BSTATE and Op.x = j0; the plain path spends 45 multiplies instead of 81)
    OPF_AEXT    = 32    // A operand is external input element Op.w (raw 29-bit repack; the B operand carries the
                        // conversion constant), single term
};
// Op = uint4 { x: B offset (table element index, or state lane), y: j0 | n << 8 | dst << 16 | flags << 24,
//              z: table element index of the constant to add, w: input index (AEXT) }
__host__ __device__ inline uint4 make_op(uint32_t b_off, int j0, int n, int dst, uint32_t flags, uint32_t rc_off = 0, uint32_t aux = 0) {
    uint4 o; o.x = b_off; o.y = (uint32_t)j0 | ((uint32_t)n << 8) | ((uint32_t)dst << 16) | (flags << 24); o.z = rc_off; o.w = aux; return o;
}
__host__ __device__ inline uint4 pad_op() { return make_op(0, 0, 1, 0, 0); }     // lane 0 * table element 0: loads only, never stored
__host__ __device__ inline uint4 make_sqr(int lane, int dst) { return make_op((uint32_t)lane, lane, 1, dst, OPF_SQR | OPF_BSTATE); }

// Segs (smw_internal.h): up to 8 (first, count) ranges of the op array, run in order, e.g.
// [absorb block 0][permutation][absorb block 1][permutation]

// Device view of one parameter set
struct PoseidonDev {
    int t, rp;
    const uint32_t* tbl;     // every table in internal form (x * 2^261 mod p, canonical limbs), EL words per element
    const uint32_t* tbl52;   // the same elements as 5 doubles of x * 2^260 mod p (fr52.cuh), same EL-word slots; nullptr when not built
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

// The interpreter.  One copy per kernel (noinline: the sponges call it once per segment list); all warps of the SM loop
// over the same code.  The state lives in the CTA's dynamic shared memory and is re-derived from `smem` here (not passed
// as pointers) so that the compiler keeps LDS/STS addressing.  Runs operations [first, first + count); the operation
// at first + count must exist (every program is followed by one padding operation): it is started and never finished.
// Each iteration finishes the pending operation `pd` (accumulators P: reduce, add, store) while starting operation `nd`
// (accumulators Q).  The initial pending operation is a zero accumulator stored to scratch lane T + 1, which is dead at
// every segment start.  Table operands mostly miss the small L1 left beside 214 KB of shared memory and come from L2,
// so they are always requested an operation (first term) or a term (the others) ahead.
template <int T, class In>
__device__ __noinline__ void run_ops(int nt, const uint32_t* __restrict__ tbl, const uint4* __restrict__ prog, int first, int count, In in) {
    extern __shared__ __align__(16) unsigned char smem[];
    const StateRef S = state_ref<T>(smem, nt, threadIdx.x);
    Acc P; acc_zero(P);
    uint4 pd = make_op(0, 0, 1, T + 1, 0);
    uint4 nd = __ldg(prog + first);
    F29 nb = f29_zero(), prc = f29_zero();                    // nd's first table operand; pd's constant (both fetched one operation ahead)
    if (!((nd.y >> 24) & OPF_BSTATE)) nb = tbl_load(tbl, nd.x);
#pragma unroll 1
    for (int k = 0; k <= count; k++) {
        uint4 nn = nd; if (k < count) nn = __ldg(prog + first + k + 1);
        const uint32_t pfl = pd.y >> 24; const int pdst = (pd.y >> 16) & 255;
        const uint32_t fl = nd.y >> 24; const int j0 = nd.y & 255, n = (nd.y >> 8) & 255;
        F29 sr = f29_zero();                                   // what is added on top of the reduced value: old lane + constant
        if (pfl & OPF_ADDSELF) sr = S.load(pdst);
        const F29 rc = prc;
        prc = f29_zero();
        if (fl & OPF_ADDRC) prc = tbl_load(tbl, nd.z);          // lands during this operation's multiplies, consumed by the next iteration
        F29 b1 = f29_zero();                                   // table operand of term 1 (tables are immutable: no hazard with the pending store)
        if (n > 1) b1 = tbl_load(tbl, nd.x + 1);
        Acc Q;
        if (!(fl & OPF_DEP)) {
            F29 a0, b0 = nb;
            if (fl & OPF_AEXT) a0 = in(nd.w); else a0 = S.load(j0);
            if (fl & OPF_BSTATE) b0 = S.load((int)nd.x);
            // ---- the fused block: 81 independent multiplies of the new operation + the dependent chain of the old one
            acc_mul(Q, a0, b0);
            S.store(pdst, acc_reduce_add2(P, sr, rc));
        } else {
            S.store(pdst, acc_reduce_add2(P, sr, rc));
            F29 a0;
            if (fl & OPF_AEXT) a0 = in(nd.w); else a0 = S.load(j0);
            if (fl & OPF_SQR) { acc_zero(Q); acc_sqr(Q, a0); }
            else { F29 b0 = nb; if (fl & OPF_BSTATE) b0 = S.load((int)nd.x); acc_mul(Q, a0, b0); }
        }
        if (k < count && !((nn.y >> 24) & OPF_BSTATE)) nb = tbl_load(tbl, nn.x);
        // remaining terms, two per iteration with alternating operand registers (the next term's are in flight during the multiplies)
        {
            int c = 1;
#pragma unroll 1
            for (int j = 1; j < n; j++) {
                F29 a1 = S.load(j0 + j), bb = b1;
                if (j + 1 < n) b1 = tbl_load(tbl, nd.x + j + 1);
                if (c == 6) { acc_normalize(Q); c = 0; }
                acc_mac(Q, a1, bb); c++;
            }
        }
        P = Q; pd = nd; nd = nn;
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
