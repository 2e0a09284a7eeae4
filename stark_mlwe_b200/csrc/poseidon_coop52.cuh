// poseidon_coop52.cuh — the LATENCY formulation of the Poseidon permutation: one permutation per group of lanes
// (t = 17: a whole warp, t = 9: a half warp), one state lane per thread, fr52 arithmetic (fr52.cuh).
//
// Why a second formulation: the batch kernels (poseidon52.cuh) run one permutation per THREAD — full lane utilisation,
// but a lone warp then needs 1.6 ms for a t = 17 permutation (1.27 M dependent-issue instructions), and everything that
// is not a multi-wave launch pays exactly that: the top levels of every tree (3.2 ms per level at t = 17,
// profiles/r01af_ncu_launches_2p16x16.csv), the strictly serial transcript sponges (tr_hash_fields_tagged over n0 elements,
// deep_ali/src/fri.rs:28-36, 541-548), the per-round challenges of the Merkle-folded sum-check.  Here the 17 (9) state
// lanes of ONE permutation sit in 17 (9) threads:
//   * full rounds (poseidon/src/lib.rs:225-235, 248-257): S-boxes lane-parallel; the dense mat-vec as one lazily
//     accumulated row product per lane (operands broadcast through shared memory), one Montgomery reduction per lane;
//   * partial rounds (:237-246) in the exact sparse form the batch kernels use (row product + rank-1 update, tables of
//     poseidon_params.cu): the S-box chain x -> x^2 -> x^4 -> x^5 of lane 0 is the only serial part; the row product
//     is ONE multiplication per lane followed by a cross-lane sum (redux.sync on 26-bit halves), and the rank-1 updates
//     of two rounds ride in the lanes that idle during the next S-box (two-round blocks, so a lane grows by < 1.02 p
//     per block and stays below 2^260 over rp/2 <= 48 blocks);
//   * the deferred (t-1) x (t-1) matrix as a dense row product per lane.
// A permutation is ~150 k warp instructions instead of 1.27 M per thread: a lone warp finishes in ~0.2 ms.
// Exact field arithmetic, same tables: outputs are bit-identical to the batch kernels and to the reference.
#pragma once
#include "poseidon52.cuh"

namespace smw {

// table element indices used by the cooperative kernels (TableLayout, poseidon_prog.cu)
struct CoopTab { uint32_t mds, rc_full, rc_part, sp_row, sp_col, post, k_in, k_out, k_rr, k_one; int rp; };
static inline CoopTab coop_tab(const smw_params* p) {
    const TableLayout& L = p->layout;
    return CoopTab{L.mds, L.rc_full, L.rc_part, L.sp_row, L.sp_col, L.post, L.k_in, L.k_out, L.k_rr, L.k_one, p->rp};
}

// The multiplication slots are inlined at their ~15 call sites (the whole permutation is ~90 KB of code): measured 11 % faster than
// one shared noinline body (proof_bytes KAT k = 18: 3.52 -> 3.14 s, profiles/r02h_kat*.log) — the call overhead (30 operand
// registers in, 10 out) outweighs the instruction-cache footprint for a lone warp.  SMW_COOP_NOINLINE selects the shared body.
// SMW_COOP_REDUCE_LAT (tuning builds): the short-dependency-chain reduction of fr52.cuh (integer quotient digit and low halves).
// Measured equal to the all-FP64 one (deep_fri_prove 2^18: 3.05 s both, profiles/r02j_ab_coop.txt): a lone warp here is bound by
// its issue rate (~2.65 cycles per instruction), not by the five-step chain, so the shorter chain buys what its extra
// instructions cost.  The default is the one reduction the batch kernels use as well.
#ifdef SMW_COOP_REDUCE_LAT
#define COOP_REDUCE acc52_reduce_add2_lat
#else
#define COOP_REDUCE acc52_reduce_add2
#endif
#ifdef SMW_COOP_NOINLINE
#define SMW_COOP_FN static __device__ __noinline__
#else
#define SMW_COOP_FN static __device__ __forceinline__
#endif
// r = (a b + M p) / 2^260 + s : one multiplication slot (~350 instructions)
SMW_COOP_FN F52 coop_mr(F52 a, F52 b, F52 s) {
    Acc52 A; acc52_mul(A, a, b);
    return COOP_REDUCE(A, 1, s, f52_zero());
}
// r = (a b + M p) / 2^260 + s + t
SMW_COOP_FN F52 coop_mr_t(F52 a, F52 b, F52 s, F52 t) {
    Acc52 A; acc52_mul(A, a, b);
    return COOP_REDUCE(A, 1, s, t);
}
// r = (a^2 + M p) / 2^260 (15 limb products): the two squarings of every S-box
SMW_COOP_FN F52 coop_sq(F52 a) {
    Acc52 A; acc52_sqr(A, a);
    return COOP_REDUCE(A, 1, f52_zero(), f52_zero());
}
// r = (a1 b1 + a2 b2 + M p) / 2^260 + s
SMW_COOP_FN F52 coop_mr2(F52 a1, F52 b1, F52 a2, F52 b2, F52 s) {
    Acc52 A; acc52_mul(A, a1, b1); acc52_mac(A, a2, b2);
    return COOP_REDUCE(A, 2, s, f52_zero());
}
SMW_DEV F52 sel52(bool c, const F52& a, const F52& b) { F52 r;
#pragma unroll
    for (int i = 0; i < 5; i++) r.v[i] = c ? a.v[i] : b.v[i];
    return r; }

// Lanes of a warp are split into 32 / G groups of G lanes; `lane` is the index inside the group.  sh points at the
// group's broadcast buffer: T slots of 6 doubles.
template <int T, int G> struct CoopCx {
    const uint32_t* tbl; CoopTab L; double* sh; int lane; int group;
    static constexpr unsigned FULL = 0xffffffffu;
    SMW_DEV F52 tab(uint32_t idx) const { return tbl52_load(tbl, idx); }
    SMW_DEV F52 bcast0(const F52& x) const { F52 r;                       // value of the group's lane 0, to every lane of the group
#pragma unroll
        for (int i = 0; i < 5; i++) r.v[i] = __shfl_sync(FULL, x.v[i], 0, G);
        return r; }
    SMW_DEV F52 bcast(const F52& x, int src) const { F52 r;             // value of the group's lane `src`, to every lane of the group
#pragma unroll
        for (int i = 0; i < 5; i++) r.v[i] = __shfl_sync(FULL, x.v[i], src, G);
        return r; }
    SMW_DEV void publish(int slot, const F52& x) const {
        double2* q = reinterpret_cast<double2*>(sh + slot * 6);
        q[0] = make_double2(x.v[0], x.v[1]); q[1] = make_double2(x.v[2], x.v[3]); sh[slot * 6 + 4] = x.v[4];
    }
    SMW_DEV F52 fetch(int slot) const {
        const double2* q = reinterpret_cast<const double2*>(sh + slot * 6);
        const double2 a = q[0], b = q[1]; F52 r; r.v[0] = a.x; r.v[1] = a.y; r.v[2] = b.x; r.v[3] = b.y; r.v[4] = sh[slot * 6 + 4];
        return r;
    }
    // sum over the lanes of the group of values < 2^260 with limbs < 2^52 (at most 32 of them: the sum stays far below 2^64
    // per limb); every lane receives the sum with limbs carried back below 2^52.  redux.sync adds 32-bit values, so a
    // limb travels as two 26-bit halves (32 x 2^26 = 2^31).
    SMW_DEV F52 group_sum(const F52& x) const {
        uint64_t sum[5];
#pragma unroll
        for (int k = 0; k < 5; k++) {
            const uint64_t l = double_to_u52(x.v[k]);
            const uint32_t lo = (uint32_t)l & 0x3ffffffu, hi = (uint32_t)(l >> 26);
            uint32_t slo = 0, shi = 0;
#pragma unroll
            for (int g = 0; g < 32 / G; g++) {
                const uint32_t a = __reduce_add_sync(FULL, group == g ? lo : 0u), b = __reduce_add_sync(FULL, group == g ? hi : 0u);
                if (group == g) { slo = a; shi = b; }
            }
            sum[k] = (uint64_t)slo + ((uint64_t)shi << 26);
        }
        F52 r; uint64_t carry = 0;
#pragma unroll
        for (int k = 0; k < 5; k++) { const uint64_t v = sum[k] + carry; r.v[k] = u52_to_double(k < 4 ? (v & M52) : v); carry = v >> 52; }
        return r;
    }
};

// y[row] = sum_{j < n} mat[row][j] * v[j] + rc, for the lanes `first_lane + row`, row < n (v[j] = state of lane first_lane + j).
// Other lanes keep their value.  One body for the full t x t matrix and the deferred (t-1) x (t-1) one.
template <int T, int G>
__device__ __noinline__ F52 coop_matvec(const CoopCx<T, G> c, F52 s, uint32_t mat, int n, int first_lane, F52 rc /* this lane's constant, added on top */) {
    const int row = c.lane - first_lane; const bool mine = row >= 0 && row < n;
    if (c.lane < T) c.publish(c.lane, s);
    __syncwarp();
    const uint32_t base = mat + (uint32_t)(mine ? row : 0) * (uint32_t)n;
    Acc52 A; acc52_mul(A, c.fetch(first_lane), c.tab(base));
    F52 b = c.tab(base + (n > 1 ? 1 : 0));
#pragma unroll 1
    for (int j = 1; j < n; j++) {
        const F52 bb = b;
        if (j + 1 < n) b = c.tab(base + j + 1);
        acc52_mac(A, c.fetch(first_lane + j), bb);
    }
    const F52 y = COOP_REDUCE(A, (uint32_t)n, rc, f52_zero());
    __syncwarp();
    return sel52(mine, y, s);
}

// the permutation on the group's lanes: lane i < T holds state[i] (value < 2^258, limbs < 2^52), other lanes hold zero
template <int T, int G>
__device__ __noinline__ F52 coop_perm(const CoopCx<T, G> c, F52 s) {
    constexpr int n = T - 1;
    const CoopTab& L = c.L;
    const int lane = c.lane; const bool in = lane < T, l0 = lane == 0, lu = lane >= 1 && lane < T;
    const int lc = in ? lane : 0;                         // clamped lane for table addressing
    const F52 zero = f52_zero();
    // ARK of full round 0: s * 1 + rc (the product with the Montgomery one also carries lazily added inputs back below 2^52 per limb)
    s = coop_mr(s, c.tab(L.k_one), in ? c.tab(L.rc_full + lc) : zero);
    // full rounds 0..3
#pragma unroll 1
    for (int r = 0; r < 4; r++) {
        const F52 x2 = coop_sq(s), x4 = coop_sq(x2); s = coop_mr(x4, s, zero);
        F52 rc = zero;                                    // the next round's constant: full round r + 1, or partial round 0 (lane 0 only)
        if (r < 3) { if (in) rc = c.tab(L.rc_full + (uint32_t)(r + 1) * T + lc); } else if (l0) rc = c.tab(L.rc_part);
        s = coop_matvec<T, G>(c, s, L.mds, T, 0, rc);
    }
#ifndef SMW_COOP_4SLOT
    // Partial rounds, two per block, THREE dependent multiplication slots per round instead of four.  The serial chain of a round is
    // l0 -> x^2 -> x^4 -> x^5 -> M00 x^5 + row . u + c; since x^5 = x^4 l0, the last two links collapse into x^4 (M00 l0) + (row . u) + c
    // with y = M00 l0 computed by a spare lane BESIDE the first squaring, and the row products (which do not depend on this round's
    // S-box) computed by the u lanes beside the second.  x^5 itself (needed for the rank-1 updates and the G term) is computed by
    // another spare lane beside the last slot.  Roles: lane 0 the S-box lane, lanes 1..n hold u, lane T computes y, lane T + 1
    // computes x^5, lane T + 2 the G term of the block's second round (T + 2 < G for both widths).
    const bool ly = lane == T, lx = lane == T + 1, lg = lane == T + 2;
    F52 pxa = zero, pxb = zero;                           // the previous block's S-box outputs (pending rank-1 update)
#pragma unroll 1
    for (int k = 0; k < L.rp; k += 2) {
        const uint32_t rowA = L.sp_row + (uint32_t)k * (T + 1), rowB = rowA + (T + 1);
        F52 l0b = c.bcast(s, 0);                          // the S-box input, known to the spare lanes
        // slot A1: lane 0 squares; lane T: y = M00 l0; lanes 1..n apply the pending update u += colA x_a + colB x_b of the previous block
        F52 ca = zero, cb = zero;
        if (lu && k > 0) { const uint32_t q = L.sp_col + (uint32_t)((k >> 1) - 1) * 2 * n + 2 * (uint32_t)(lane - 1); ca = c.tab(q); cb = c.tab(q + 1); }
        F52 r = coop_mr2(l0 ? s : (lu ? ca : (ly ? c.tab(rowA + (uint32_t)n) : zero)), l0 ? s : (lu ? pxa : (ly ? l0b : zero)),
                         lu ? cb : zero, lu ? pxb : zero, lu ? s : zero);
        F52 x2 = r, y = r; if (lu) s = r;
        // slot B1: lane 0 squares again; lanes 1..n: row_k[j] u_j (u is not updated inside the block: the second round carries the G term)
        r = coop_mr(l0 ? x2 : (lu ? s : zero), l0 ? x2 : (lu ? c.tab(rowA + (uint32_t)(lane - 1)) : zero), zero);
        F52 sum = c.group_sum(sel52(lu, r, zero));
        F52 x4b = c.bcast(r, 0), yb = c.bcast(y, T);
        // slot C1: lane 0: l0' = x^4 y + sum + c_{k+1}; lane T + 1: x_a = x^4 l0
        r = coop_mr_t((l0 || lx) ? x4b : zero, l0 ? yb : (lx ? l0b : zero), l0 ? sum : zero, l0 ? c.tab(L.rc_part + (uint32_t)k + 1) : zero);
        if (l0) s = r;
        const F52 xa = c.bcast(r, T + 1);
        l0b = c.bcast(r, 0);
        // second round.  slot A2: lane 0 squares; lane T: y = M00 l0'; lane T + 2: G x_a + the next constant
        const uint32_t next_rc = (k + 2 < L.rp) ? L.rc_part + (uint32_t)k + 2 : L.rc_full + 4u * T;
        r = coop_mr(l0 ? s : (ly ? c.tab(rowB + (uint32_t)n + 1) : (lg ? c.tab(rowB + (uint32_t)n) : zero)), l0 ? s : (ly ? l0b : (lg ? xa : zero)),
                    lg ? c.tab(next_rc) : zero);
        x2 = r; y = r; const F52 g = r;
        r = coop_mr(l0 ? x2 : (lu ? s : zero), l0 ? x2 : (lu ? c.tab(rowB + (uint32_t)(lane - 1)) : zero), zero);
        sum = c.group_sum(sel52(lu, r, zero));
        x4b = c.bcast(r, 0); yb = c.bcast(y, T);
        const F52 gb = c.bcast(g, T + 2);
        r = coop_mr_t((l0 || lx) ? x4b : zero, l0 ? yb : (lx ? l0b : zero), l0 ? sum : zero, l0 ? gb : zero);
        if (l0) s = r;
        pxa = xa; pxb = c.bcast(r, T + 1);
    }
#else
    // partial rounds, two per block.  Lane 0: the S-box lane; lanes 1..n: u; lane T: the G term of the second round.
    F52 pxa = zero, pxb = zero;                           // the previous block's S-box outputs (pending rank-1 update)
#pragma unroll 1
    for (int k = 0; k < L.rp; k += 2) {
        const uint32_t rowA = L.sp_row + (uint32_t)k * (T + 1), rowB = rowA + (T + 1);
        // slot A1: lane 0 squares; lanes 1..n apply the pending update u += colA x_a + colB x_b of the previous block
        F52 ca = zero, cb = zero;
        if (lu && k > 0) { const uint32_t q = L.sp_col + (uint32_t)((k >> 1) - 1) * 2 * n + 2 * (uint32_t)(lane - 1); ca = c.tab(q); cb = c.tab(q + 1); }
        F52 r = coop_mr2(l0 ? s : ca, l0 ? s : pxa, l0 ? zero : cb, l0 ? zero : pxb, l0 ? zero : s);
        F52 x2 = r; if (!l0) s = r;
        F52 x4 = coop_sq(x2);
        F52 xa = c.bcast0(coop_mr(x4, s, zero));          // x_a = S-box output of round k (lane 0's s is the S-box input)
        // slot D1: lane 0: M00 x_a + rc_{k+1}; lanes 1..n: row_k[j] u_j; summed over the group -> the new lane 0
        F52 a = l0 ? xa : (lu ? s : zero), b = c.tab(rowA + (l0 ? (uint32_t)n : (uint32_t)(lu ? lane - 1 : 0)));
        r = coop_mr(a, b, l0 ? c.tab(L.rc_part + (uint32_t)k + 1) : zero);
        F52 t0 = c.group_sum(sel52(in, r, zero));
        if (l0) s = t0;
        // second round of the block
        x2 = coop_sq(s); x4 = coop_sq(x2);
        F52 xb = c.bcast0(coop_mr(x4, s, zero));
        const bool lg = lane == T;                        // the lane that carries G x_a
        a = l0 ? xb : (lu ? s : (lg ? xa : zero));
        b = c.tab(rowB + (l0 ? (uint32_t)n + 1 : (lu ? (uint32_t)(lane - 1) : (uint32_t)n)));
        const uint32_t next_rc = (k + 2 < L.rp) ? L.rc_part + (uint32_t)k + 2 : L.rc_full + 4u * T;
        r = coop_mr(a, b, l0 ? c.tab(next_rc) : zero);
        t0 = c.group_sum(sel52(in || lg, r, zero));
        if (l0) s = t0;
        pxa = xa; pxb = xb;
    }
#endif
    {   // the last block's pending update
        F52 ca = zero, cb = zero;
        if (lu) { const uint32_t q = L.sp_col + (uint32_t)((L.rp >> 1) - 1) * 2 * n + 2 * (uint32_t)(lane - 1); ca = c.tab(q); cb = c.tab(q + 1); }
        const F52 r = coop_mr2(ca, pxa, cb, pxb, lu ? s : zero);
        if (lu) s = r;
    }
    // the deferred dense matrix Mh^rp on lanes 1..n, with the ARK of full round 4 (lane 0 already carries its constant)
    s = coop_matvec<T, G>(c, s, L.post, n, 1, lu ? c.tab(L.rc_full + 4u * T + lc) : zero);
    // full rounds 4..7
#pragma unroll 1
    for (int r = 4; r < 8; r++) {
        const F52 x2 = coop_sq(s), x4 = coop_sq(x2); s = coop_mr(x4, s, zero);
        s = coop_matvec<T, G>(c, s, L.mds, T, 0, (r < 7 && in) ? c.tab(L.rc_full + (uint32_t)(r + 1) * T + lc) : zero);
    }
    return sel52(in, s, zero);
}

} // namespace smw
