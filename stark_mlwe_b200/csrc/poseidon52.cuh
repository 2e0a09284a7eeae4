// poseidon52.cuh — the row-operation interpreter of poseidon.cuh over fr52.cuh (5 x 52-bit limbs on the FP64 pipe), for the
// throughput-bound batch kernels of widths 9 and 17 (Merkle levels, pair leaves, FRI hashed leaves, batched permutation).
// Same programs (poseidon_prog.cu), same table layout (element index -> 48-byte slot, here holding 5 doubles of x * 2^260 mod p
// instead of nine 29-bit limbs of x * 2^261), same software pipelining; the reduction needs the pending operation's term
// count to strip the exponent bits the DFMA split leaves in the accumulators.  Values stay below 2^260: an S-box output is
// < 2^257, a row product of <= 18 terms + old lane + constant < 2^258, and a lane of the sparse partial rounds grows by
// < 2^254.4 per two-round block (32 blocks at t = 17, 30 at t = 9) — the wider parameter sets keep the 29-bit path.
#pragma once
#include "poseidon.cuh"
#include "fr52.cuh"

namespace smw {

SMW_DEV F52 tbl52_load(const uint32_t* base, uint32_t idx) {
    const double2* q = reinterpret_cast<const double2*>(base + (size_t)idx * EL);
    double2 a = __ldg(q), b = __ldg(q + 1); double c = __ldg(reinterpret_cast<const double*>(q + 2));
    F52 r; r.v[0] = a.x; r.v[1] = a.y; r.v[2] = b.x; r.v[3] = b.y; r.v[4] = c;
    return r;
}
// Per-thread state in shared memory: lane j = two double2 (limbs 0-1, 2-3) + one double (limb 4), word-interleaved over the threads
struct StateRef52 {
    double2* q;     // [T+2][2][NT]
    double* w;      // [T+2][NT]
    int nt, tid;
    SMW_DEV F52 load(int j) const {
        double2 a = q[(2 * j) * nt + tid], b = q[(2 * j + 1) * nt + tid];
        F52 r; r.v[0] = a.x; r.v[1] = a.y; r.v[2] = b.x; r.v[3] = b.y; r.v[4] = w[j * nt + tid];
        return r;
    }
    SMW_DEV void store(int j, const F52& x) const {
        q[(2 * j) * nt + tid] = make_double2(x.v[0], x.v[1]);
        q[(2 * j + 1) * nt + tid] = make_double2(x.v[2], x.v[3]);
        w[j * nt + tid] = x.v[4];
    }
};
template <int T> SMW_DEV StateRef52 state_ref52(unsigned char* smem, int nt, int tid) {
    StateRef52 S; S.q = reinterpret_cast<double2*>(smem); S.w = reinterpret_cast<double*>(smem + (size_t)(T + 2) * 2 * nt * 16); S.nt = nt; S.tid = tid; return S;
}
template <int T> __host__ __device__ constexpr size_t state_smem_bytes52(int nt) { return (size_t)(T + 2) * nt * 40; }

struct NoInput52 { SMW_DEV F52 f52(uint32_t) const { return f52_zero(); } };

// See run_ops (poseidon.cuh) for the structure; differences: accumulators are 10 raw-bit columns, squares take the general
// product (a doubled operand would leave the 2^104 window of the split), and the reduction is told the term count.
template <int T, class In>
__device__ __noinline__ void run_ops52(int nt, const uint32_t* __restrict__ tbl, const uint4* __restrict__ prog, int first, int count, In in) {
    extern __shared__ __align__(16) unsigned char smem[];
    const StateRef52 S = state_ref52<T>(smem, nt, threadIdx.x);
    Acc52 P; acc52_mul(P, f52_zero(), f52_zero());            // the initial pending operation: one term of zeros (carries exactly one term's exponent bits)
    uint4 pd = make_op(0, 0, 1, T + 1, 0);
    uint4 nd = __ldg(prog + first);
    F52 nb = f52_zero(), prc = f52_zero();
    if (!((nd.y >> 24) & OPF_BSTATE)) nb = tbl52_load(tbl, nd.x);
#pragma unroll 1
    for (int k = 0; k <= count; k++) {
        uint4 nn = nd; if (k < count) nn = __ldg(prog + first + k + 1);
        const uint32_t pfl = pd.y >> 24; const int pdst = (pd.y >> 16) & 255; const uint32_t pn = (pd.y >> 8) & 255;
        const uint32_t fl = nd.y >> 24; const int j0 = nd.y & 255, n = (nd.y >> 8) & 255;
        F52 sr = f52_zero();
        if (pfl & OPF_ADDSELF) sr = S.load(pdst);
        const F52 rc = prc;
        prc = f52_zero();
        if (fl & OPF_ADDRC) prc = tbl52_load(tbl, nd.z);
        F52 b1 = f52_zero();
        if (n > 1) b1 = tbl52_load(tbl, nd.x + 1);
        Acc52 Q;
        if (!(fl & (OPF_DEP | OPF_AEXT))) {
            // operands without a branch (a state-state product reads lane nd.x, a table product re-reads lane j0 and drops it):
            // one basic block, so the 25 independent limb products can be scheduled into the whole dependent reduction chain
            const bool bs = (fl & OPF_BSTATE) != 0;
            const F52 a0 = S.load(j0), bv = S.load(bs ? (int)nd.x : j0);
            F52 b0;
#pragma unroll
            for (int i = 0; i < 5; i++) b0.v[i] = bs ? bv.v[i] : nb.v[i];
            acc52_mul(Q, a0, b0);
            S.store(pdst, acc52_reduce_add2(P, pn, sr, rc));
        } else {
            S.store(pdst, acc52_reduce_add2(P, pn, sr, rc));
            F52 a0, b0 = nb;
            if (fl & OPF_AEXT) a0 = in.f52(nd.w); else a0 = S.load(j0);
            if (fl & OPF_BSTATE) b0 = S.load((int)nd.x);
            acc52_mul(Q, a0, b0);
        }
        if (k < count && !((nn.y >> 24) & OPF_BSTATE)) nb = tbl52_load(tbl, nn.x);
        // remaining terms
        // two terms per basic block (the tail of one product's additions overlaps the head of the next), one left-over term
        if (n > 1) {
            F52 b2;
            int j = 1;
#pragma unroll 1
            while (j + 1 < n) {
                const F52 a1 = S.load(j0 + j), a2 = S.load(j0 + j + 1);
                b2 = tbl52_load(tbl, nd.x + j + 1);
                acc52_mac(Q, a1, b1);
                if (j + 2 < n) b1 = tbl52_load(tbl, nd.x + j + 2);
                acc52_mac(Q, a2, b2);
                j += 2;
            }
            if (j < n) { const F52 a1 = S.load(j0 + j); acc52_mac(Q, a1, b1); }
        }
        P = Q; pd = nd; nd = nn;
    }
}
template <int T, class In>
SMW_DEV void run_segs52(int nt, const PoseidonDev& P, const Segs& sg, const In& in) {
#pragma unroll 1
    for (int s = 0; s < sg.n; s++) run_ops52<T>(nt, P.tbl52, P.prog, sg.first[s], sg.count[s], in);
}

} // namespace smw
