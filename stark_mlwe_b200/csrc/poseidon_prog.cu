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
    L.total = q + 1;          // one padding element: the interpreter never reads past it, but keep the tail defined
    return L;
}

namespace {
struct Builder {
    std::vector<uint4>& ops; int t, rp; TableLayout L; int TMP;
    // x^5 of lane `lane` in place, lane TMP as scratch (poseidon/src/lib.rs:24-29: x2 = x*x, x4 = x2*x2, x*x4)
    void sbox(int lane) {
        ops.push_back(make_op(0, lane, 1, TMP, OPF_SQR));
        ops.push_back(make_op(0, TMP, 1, TMP, OPF_SQR));
        ops.push_back(make_op((uint32_t)TMP, lane, 1, lane, OPF_BSTATE));
    }
    // x^5 of the value held in TMP, result in lane `out` (which doubles as scratch): the sparse partial rounds
    void sbox_tmp_to(int out) {
        ops.push_back(make_op(0, TMP, 1, out, OPF_SQR));
        ops.push_back(make_op(0, out, 1, out, OPF_SQR));
        ops.push_back(make_op((uint32_t)out, TMP, 1, out, OPF_BSTATE));
    }
    // lanes [off, off + n) <- (L*U) * lanes, in place, lu = n x n table at element index `lu`.
    // rc_of(lane) >= 0: table index of the constant added to that lane's final value (next round's ARK); lane `off`
    // gets its final value in the U sweep but is still read by the L sweep, so its constant is added by a LOAD op at the end.
    template <class RcOf> void matvec_lu(uint32_t lu, int n, int off, RcOf rc_of) {
        for (int i = 0; i < n; i++)                 // U sweep, ascending rows: row i reads lanes >= i only
            ops.push_back(make_op(lu + (uint32_t)i * n + i, off + i, n - i, off + i, 0));
        for (int i = n - 1; i >= 1; i--) {          // L sweep (unit diagonal), descending rows: row i reads lanes < i, adds lane i
            long rc = rc_of(off + i);
            ops.push_back(make_op(lu + (uint32_t)i * n, off, i, off + i, OPF_ADDSELF | (rc >= 0 ? OPF_ADDRC : 0u), rc >= 0 ? (uint32_t)rc : 0u));
        }
    }
    void add_const(int src, int dst, long rc) { ops.push_back(make_op(0, 0, 0, dst, OPF_LOAD | (rc >= 0 ? OPF_ADDRC : 0u), rc >= 0 ? (uint32_t)rc : 0u, (uint32_t)src)); }
};
}

// Appends the permutation program; lanes 0..t-1 in, lanes 0..t-1 out, lane t scratch.
void build_perm_program(std::vector<uint4>& ops, int t, int rp, bool sparse) {
    Builder B{ops, t, rp, table_layout(t, rp), t + 1};      // lane t = X (second S-box output of a partial-round block), lane t + 1 = TMP
    const TableLayout& L = B.L;
    auto rcf = [&](int r, int lane) { return (long)(L.rc_full + (uint32_t)r * t + lane); };
    auto rcp = [&](int k) { return (long)(L.rc_part + (uint32_t)k); };
    // ARK of full round 0
    for (int i = 0; i < t; i++) B.add_const(i, i, rcf(0, i));
    // stage list: F0..F3, P0..P(rp-1), F4..F7
    for (int r = 0; r < 4; r++) {
        for (int i = 0; i < t; i++) B.sbox(i);
        if (r < 3) {
            B.matvec_lu(L.lu_full, t, 0, [&](int lane) { return rcf(r + 1, lane); });
            B.add_const(0, 0, rcf(r + 1, 0));
        } else if (sparse) {        // next: partial round 0, whose S-box input lives in TMP
            B.matvec_lu(L.lu_full, t, 0, [&](int) { return -1L; });
            B.add_const(0, B.TMP, rcp(0));
        } else {
            B.matvec_lu(L.lu_full, t, 0, [&](int) { return -1L; });
            B.add_const(0, 0, rcp(0));
        }
    }
    if (sparse) {
        // Two partial rounds per block (rp is even for every supported width).  With x = S-box output of round k and
        // y = that of round k + 1, the lanes 1.. are only updated once per block:
        //   TMP  = row_k . (x, u_1 ..)                 + c_{k+1}
        //   TMP' = (G, row_{k+1}[1..], M00) . (x, u_1 .., y) + c_{k+2}      G = sum_i row_{k+1}[i] col_k[i-1]
        //   u_i += col_k[i-1] x + col_{k+1}[i-1] y     (one 2-term operation per lane instead of two 1-term ones)
        const int X = t;
        for (int k = 0; k < rp; k += 2) {
            B.sbox_tmp_to(0);                                                          // lane 0 = x
            ops.push_back(make_op(L.sp_row + (uint32_t)k * (t + 1), 0, t, B.TMP, OPF_ADDRC, (uint32_t)rcp(k + 1), 1));
            B.sbox_tmp_to(X);                                                          // lane X = y
            long next = (k + 2 < rp) ? rcp(k + 2) : rcf(4, 0);
            ops.push_back(make_op(L.sp_row + (uint32_t)(k + 1) * (t + 1), 0, t + 1, B.TMP, OPF_ADDRC, (uint32_t)next, 1));
            for (int i = 1; i < t; i++)                                               // A lanes 0 and X: stride t
                ops.push_back(make_op(L.sp_col + (uint32_t)(k / 2) * 2 * (t - 1) + 2 * (i - 1), 0, 2, i, OPF_ADDSELF, 0, (uint32_t)X));
        }
        // the deferred dense (t-1)x(t-1) matrix on lanes 1.., then ARK of full round 4
        B.matvec_lu(L.lu_post, t - 1, 1, [&](int lane) { return rcf(4, lane); });
        B.add_const(1, 1, rcf(4, 1));
        B.add_const(B.TMP, 0, -1);
    } else {
        for (int k = 0; k < rp; k++) {
            B.sbox(0);
            bool last = (k + 1 == rp);
            B.matvec_lu(L.lu_full, t, 0, [&](int lane) { return last ? rcf(4, lane) : -1L; });
            B.add_const(0, 0, last ? rcf(4, 0) : rcp(k + 1));
        }
    }
    for (int r = 4; r < 8; r++) {
        for (int i = 0; i < t; i++) B.sbox(i);
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
        ops.push_back(make_op(0, 0, 0, t, OPF_LOAD, 0, (uint32_t)t));                 // padding op (prefetch target)
        sg.first[sg.n] = first; sg.count[sg.n] = cnt; sg.n++;
        sg.first[sg.n] = perm_first; sg.count[sg.n] = perm_count; sg.n++;
    }
    return sg;
}

} // namespace smw
