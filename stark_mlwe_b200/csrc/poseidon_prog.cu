// poseidon_prog.cu — host-side builders of the row-operation programs that poseidon.cuh's interpreter runs.
// Pure index bookkeeping: no field arithmetic happens here (the tables the indices point into are derived on
// the device by poseidon_params.cu).
//
// Reference round structure (poseidon/src/lib.rs:219-258): 4 full rounds, RP partial rounds, 4 full rounds; each
// round = add round constants (all lanes / lane 0), x^5 (all lanes / lane 0), dense matrix.  In program form the
// round-constant addition of round r+1 is folded into the operation that produces each lane in round r.
#include "poseidon_kernels.cuh"

namespace smw {

TableLayout table_layout(int t, int rp) {
    TableLayout L; uint32_t q = 0;
    L.lu_full = q; q += (uint32_t)t * t;
    L.rc_full = q; q += 8u * t;
    L.rc_part = q; q += (uint32_t)rp;
    L.sp_row = q; q += (uint32_t)rp * (t + 1);
    L.sp_col = q; q += (uint32_t)rp * (t - 1);
    L.lu_post = q; q += (uint32_t)(t - 1) * (t - 1);
    L.mds = q; q += (uint32_t)t * t;
    L.k_in = q++; L.k_out = q++; L.k_rr = q++; L.k_one = q++;
    L.post = q; q += (uint32_t)(t - 1) * (t - 1);      // the deferred matrix Mh^rp itself (dense; the cooperative kernels of poseidon_coop52.cuh apply it row per lane)
    L.total = q + 1;          // one padding element: the interpreter never reads past it, but keep the tail defined
    return L;
}

namespace {
struct Builder {
    std::vector<uint4>& ops; int t, rp; TableLayout L; int X, Y;     // X = lane t, Y = lane t + 1: the two scratch lanes
    void mul_lanes(int a, int b, int dst) { ops.push_back(make_op((uint32_t)b, a, 1, dst, OPF_BSTATE)); }
    // x^5 in place (poseidon/src/lib.rs:24-29: x2 = x*x, x4 = x2*x2, x*x4), scratch lane `tmp`
    void sbox(int lane, int tmp) { ops.push_back(make_sqr(lane, tmp)); ops.push_back(make_sqr(tmp, tmp)); mul_lanes(lane, tmp, lane); }
    // two S-boxes interleaved: no operation reads the lane its predecessor writes, so every one of them overlaps (poseidon.cuh)
    void sbox_pair(int la, int lb) {
        ops.push_back(make_sqr(la, X)); ops.push_back(make_sqr(lb, Y));
        ops.push_back(make_sqr(X, X)); ops.push_back(make_sqr(Y, Y));
        mul_lanes(la, X, la); mul_lanes(lb, Y, lb);
    }
    void sbox_all() { int i = 0; for (; i + 1 < t; i += 2) sbox_pair(i, i + 1); if (i < t) sbox(i, X); }
    // y = x^5 of lane `src`, written to lane `dst` != src (partial rounds: lane 0 keeps the S-box input until the row product overwrites it)
    void sbox_to(int src, int dst) { ops.push_back(make_sqr(src, dst)); ops.push_back(make_sqr(dst, dst)); mul_lanes(src, dst, dst); }
    // lanes [off, off + n) <- (L*U) * lanes, in place, lu = n x n table at element index `lu`.
    // rc_of(lane) >= 0: table index of the constant added to that lane's final value (next round's ARK); lane `off`
    // gets its final value in the U sweep but is still read by the L sweep, so its constant is added by add_const at the end.
    template <class RcOf> void matvec_lu(uint32_t lu, int n, int off, RcOf rc_of) {
        for (int i = 0; i < n; i++)                 // U sweep, ascending rows: row i reads lanes >= i only
            ops.push_back(make_op(lu + (uint32_t)i * n + i, off + i, n - i, off + i, 0));
        for (int i = n - 1; i >= 1; i--) {          // L sweep (unit diagonal), descending rows: row i reads lanes < i, adds lane i
            long rc = rc_of(off + i);
            ops.push_back(make_op(lu + (uint32_t)i * n, off, i, off + i, OPF_ADDSELF | (rc >= 0 ? OPF_ADDRC : 0u), rc >= 0 ? (uint32_t)rc : 0u));
        }
    }
    // S[dst] = S[src] (+ constant): a one-term product with the Montgomery one, so that the interpreter has a single operation shape
    void add_const(int src, int dst, long rc) { ops.push_back(make_op(L.k_one, src, 1, dst, rc >= 0 ? OPF_ADDRC : 0u, rc >= 0 ? (uint32_t)rc : 0u)); }
};
}

// Sets OPF_DEP on every operation whose first term reads the lane written by the operation before it in the array (the
// interpreter loads those operands before the predecessor's store).  Later terms, the added old lane and constants are
// loaded after that store and need no flag.  Idempotent; depends only on an operation and its array predecessor.
void mark_deps(std::vector<uint4>& ops, size_t from) {
    for (size_t k = from > 0 ? from : 1; k < ops.size(); k++) {
        const uint32_t fl = (ops[k].y >> 24) & ~(uint32_t)OPF_DEP;
        const int prev_dst = (int)((ops[k - 1].y >> 16) & 255);
        bool dep = false;
        if (!(fl & OPF_AEXT) && (int)(ops[k].y & 255) == prev_dst) dep = true;
        if ((fl & OPF_BSTATE) && (int)ops[k].x == prev_dst) dep = true;
        ops[k].y = (ops[k].y & 0x00ffffffu) | ((fl | (dep ? (uint32_t)OPF_DEP : 0u)) << 24);
    }
}
// multiply-accumulates (32x32+64) one run of operations executes: 81 per term (45 for a square on the plain path) + 45 per reduction
uint64_t count_macs(const std::vector<uint4>& ops, int first, int count) {
    uint64_t m = 0;
    for (int k = first; k < first + count; k++) {
        const uint32_t fl = ops[k].y >> 24; const uint64_t n = (ops[k].y >> 8) & 255;
        m += (n ? n - 1 : 0) * 81 + (((fl & OPF_DEP) && (fl & OPF_SQR)) ? 45 : 81) + 45;
    }
    return m;
}

// FP64-pipe instructions (DFMA.RZ + DADD) one run of operations executes on the fr52 path: 75 per term (25 limb products x
// (high part, subtraction, low part)) + 85 per reduction (5 steps x 14, 10 addend conversions, 5 limb conversions)
uint64_t count_fp64_ops(const std::vector<uint4>& ops, int first, int count) {
    uint64_t m = 0;
    for (int k = first; k < first + count; k++) m += (uint64_t)((ops[k].y >> 8) & 255) * 75 + 85;
    return m;
}

// Appends the permutation program; lanes 0..t-1 in, lanes 0..t-1 out, lanes t, t+1 scratch.
void build_perm_program(std::vector<uint4>& ops, int t, int rp, bool sparse) {
    Builder B{ops, t, rp, table_layout(t, rp), t, t + 1};
    const TableLayout& L = B.L;
    auto rcf = [&](int r, int lane) { return (long)(L.rc_full + (uint32_t)r * t + lane); };
    auto rcp = [&](int k) { return (long)(L.rc_part + (uint32_t)k); };
    // ARK of full round 0
    for (int i = 0; i < t; i++) B.add_const(i, i, rcf(0, i));
    // stage list: F0..F3, P0..P(rp-1), F4..F7
    for (int r = 0; r < 4; r++) {
        B.sbox_all();
        if (r < 3) {
            B.matvec_lu(L.lu_full, t, 0, [&](int lane) { return rcf(r + 1, lane); });
            B.add_const(0, 0, rcf(r + 1, 0));
        } else {                    // next: partial round 0, whose S-box input is lane 0
            B.matvec_lu(L.lu_full, t, 0, [&](int) { return -1L; });
            B.add_const(0, 0, rcp(0));
        }
    }
    if (sparse) {
        // Two partial rounds per block (rp is even for every supported width).  Lane 0 holds the S-box input, lanes
        // 1..t-1 the vector u; x = S-box output of round k goes to lane X = t and y = that of round k + 1 to lane
        // Y = t + 1, so that both row products run over CONSECUTIVE lanes starting at u_1 (their first term never
        // depends on the S-box that precedes them) and u is only updated once per block:
        //   lane0 = (row_k[1..], M00) . (u_1 .., x)                    + c_{k+1}
        //   lane0 = (row_{k+1}[1..], G, M00) . (u_1 .., x, y)          + c_{k+2}      G = sum_i row_{k+1}[i] col_k[i-1]
        //   u_i  += col_k[i-1] x + col_{k+1}[i-1] y                    (one 2-term operation per lane)
        for (int k = 0; k < rp; k += 2) {
            B.sbox_to(0, B.X);
            ops.push_back(make_op(L.sp_row + (uint32_t)k * (t + 1), 1, t, 0, OPF_ADDRC, (uint32_t)rcp(k + 1)));
            B.sbox_to(0, B.Y);
            long next = (k + 2 < rp) ? rcp(k + 2) : rcf(4, 0);
            ops.push_back(make_op(L.sp_row + (uint32_t)(k + 1) * (t + 1), 1, t + 1, 0, OPF_ADDRC, (uint32_t)next));
            for (int i = 1; i < t; i++)
                ops.push_back(make_op(L.sp_col + (uint32_t)(k / 2) * 2 * (t - 1) + 2 * (i - 1), B.X, 2, i, OPF_ADDSELF));
        }
        // the deferred dense (t-1)x(t-1) matrix on lanes 1.., then ARK of full round 4 (lane 0 already carries its constant)
        B.matvec_lu(L.lu_post, t - 1, 1, [&](int lane) { return rcf(4, lane); });
        B.add_const(1, 1, rcf(4, 1));
    } else {
        for (int k = 0; k < rp; k++) {
            B.sbox(0, B.X);
            bool last = (k + 1 == rp);
            B.matvec_lu(L.lu_full, t, 0, [&](int lane) { return last ? rcf(4, lane) : -1L; });
            B.add_const(0, 0, last ? rcf(4, 0) : rcp(k + 1));
        }
    }
    for (int r = 4; r < 8; r++) {
        B.sbox_all();
        if (r < 7) {
            B.matvec_lu(L.lu_full, t, 0, [&](int lane) { return rcf(r + 1, lane); });
            B.add_const(0, 0, rcf(r + 1, 0));
        } else {
            B.matvec_lu(L.lu_full, t, 0, [&](int) { return -1L; });
        }
    }
}

// Absorb program of the DS sponge (poseidon/src/lib.rs:276-312) for `n_ds` small-integer DS fields followed by up to
// `max_inputs` field elements and the padding one: input index idx < count is the element, idx == count the padding
// one (external form of 1), idx > count zero — so one program serves ragged nodes.  The kernel writes the raw DS
// integers into lanes 0..n_ds-1 beforehand.  Returns the segment list [absorb block 0][perm][absorb block 1][perm]...
Segs build_sponge_program(std::vector<uint4>& ops, int t, int rp, int n_ds, int max_inputs, int perm_first, int perm_count) {
    TableLayout L = table_layout(t, rp);
    const int rate = t - 1, total = n_ds + max_inputs + 1;
    Segs sg; sg.n = 0;
    int s = 0;
    while (s < total) {
        int first = (int)ops.size();
        int blk_end = std::min(total, (s / rate + 1) * rate);
        for (; s < blk_end; s++) {
            int lane = s % rate; bool first_block = s < rate;
            if (s < n_ds) ops.push_back(make_op(L.k_rr, lane, 1, lane, 0));           // raw integer * 2^522 / 2^261
            else ops.push_back(make_op(L.k_in, 0, 1, lane, OPF_AEXT | (first_block ? 0u : OPF_ADDSELF), 0, (uint32_t)(s - n_ds)));
        }
        int cnt = (int)ops.size() - first;
        ops.push_back(pad_op());                                                      // padding op (started, never finished)
        sg.first[sg.n] = first; sg.count[sg.n] = cnt; sg.n++;
        sg.first[sg.n] = perm_first; sg.count[sg.n] = perm_count; sg.n++;
    }
    return sg;
}

} // namespace smw
