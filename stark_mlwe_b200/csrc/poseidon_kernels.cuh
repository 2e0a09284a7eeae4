// poseidon_kernels.cuh — kernel templates shared by the per-width translation units.
// Every kernel is: a short prologue (zero the lanes, write the raw DS integers), ONE interpreter (run_ops) fed with
// segment lists built on the host, and a short epilogue (canonicalise + repack the digest).
#pragma once
#include "smw_internal.h"
#include "poseidon.cuh"
#include "poseidon52.cuh"
#include <vector>
#include <algorithm>

namespace smw {

// launch shape per width: threads per CTA and CTAs per SM so that the per-thread state ((T + 1) * 36 B)
// fills shared memory: t=9 -> 2 x 256, t=17 -> 2 x 160 (207 KB), larger widths 1 CTA.
template <int T> struct Cfg;
// NT / MINB: the __launch_bounds__ (register budget for that many resident threads); RUN: threads per CTA actually launched
template <> struct Cfg<9>   { static constexpr int NT = 256, MINB = 2, RUN = 64; };
template <> struct Cfg<17>  { static constexpr int NT = 160, MINB = 2, RUN = 64; };
template <> struct Cfg<33>  { static constexpr int NT = 160, MINB = 1, RUN = 160; };
template <> struct Cfg<65>  { static constexpr int NT = 96,  MINB = 1, RUN = 96; };
template <> struct Cfg<129> { static constexpr int NT = 32,  MINB = 1, RUN = 32; };

static PoseidonDev dev_view(const smw_params* p) {
    PoseidonDev d; d.t = p->t; d.rp = p->rp; d.tbl = p->d_tables; d.prog = p->d_prog;
    d.perm_first = p->perm_first; d.perm_count = p->perm_count; d.dense_first = p->dense_first; d.dense_count = p->dense_count;
    d.tbl52 = p->d_tables52;
    d.k_in = p->layout.k_in; d.k_out = p->layout.k_out; d.k_rr = p->layout.k_rr; d.k_one = p->layout.k_one;
    return d;
}

// ------------------------------------------------------------------ plain (non-__ldg) table access
SMW_DEV F29 gl(const uint32_t* base, size_t idx) { F29 r; const uint32_t* q = base + idx * EL;
#pragma unroll
    for (int i = 0; i < 9; i++) r.v[i] = q[i];
    return r; }
SMW_DEV void gs(uint32_t* base, size_t idx, const F29& x) { uint32_t* q = base + idx * EL;
#pragma unroll
    for (int i = 0; i < 9; i++) q[i] = x.v[i];
    q[9] = 0; q[10] = 0; q[11] = 0; }
SMW_DEV F29 mulc(const F29& a, const F29& b) { return f29_canonicalize(f29_mul(a, b)); }   // canonical in -> canonical out

// 256-bit little-endian integer (any value < 2^256) -> internal form, canonical
SMW_DEV F29 f29_from_le256(const uint32_t w[8]) { return f29_canonicalize(f29_mul(f29_repack_from32(w), f29_const(C_RR29))); }

SMW_DEV F29 raw_ext(const uint32_t* p) { uint32_t w[8]; load_ext(p, w); return f29_repack_from32(w); }
SMW_DEV F29 raw_one_ext() { uint32_t w[8];           // external form of 1 (= 2^256 mod p), repacked; constant-folded
#pragma unroll
    for (int i = 0; i < 8; i++) w[i] = C_R32W[i];
    return f29_repack_from32(w); }
SMW_DEV F29 raw_u64(uint64_t x) { F29 a = f29_zero(); a.v[0] = (uint32_t)x & M29; a.v[1] = (uint32_t)(x >> 29) & M29; a.v[2] = (uint32_t)(x >> 58); return a; }
// lane value after the k_out multiplication (< 2p) -> canonical external words
SMW_DEV void emit_ext(const F29& x, uint32_t* dst) { uint32_t w[8]; f29_repack_to32(f29_canonicalize(x), w); store_ext(dst, w); }
SMW_DEV void zero_lanes(const StateRef& S, int n) {
#pragma unroll 1
    for (int i = 0; i < n; i++) S.store(i, f29_zero());
}

// ------------------------------------------------------------------ batched permutation
SMW_DEV F52 raw_ext52(const uint32_t* p) { uint32_t w[8]; load_ext(p, w); return f52_repack_from32(w); }
SMW_DEV F52 raw_one_ext52() { uint32_t w[8];
#pragma unroll
    for (int i = 0; i < 8; i++) w[i] = C_R32W[i];
    return f52_repack_from32(w); }
SMW_DEV void emit_ext52(const F52& x, uint32_t* dst) { uint64_t l[5]; f52_canonical_limbs(x, l); uint32_t w[8]; f52_limbs_to32(l, w); store_ext(dst, w); }
struct InStates { const uint32_t* st; SMW_DEV F29 operator()(uint32_t idx) const { return raw_ext(st + (size_t)idx * 8); }
                  SMW_DEV F52 f52(uint32_t idx) const { return raw_ext52(st + (size_t)idx * 8); } };
template <int T> __global__ void __launch_bounds__(Cfg<T>::NT, Cfg<T>::MINB)
k_permute_batch(uint32_t* states, size_t batch, Segs sg, PoseidonDev P) {
    extern __shared__ __align__(16) unsigned char smem[];
    size_t b = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (b >= batch) return;
    StateRef S = state_ref<T>(smem, blockDim.x, threadIdx.x);
    uint32_t* st = states + b * T * 8;
    InStates in{st};
    run_segs<T>(blockDim.x, P, sg, in);         // [load lanes][permutation][lanes * 2^256 / 2^261]
#pragma unroll 1
    for (int i = 0; i < T; i++) emit_ext(S.load(i), st + i * 8);
}

// ------------------------------------------------------------------ Merkle level (merkle/src/lib.rs:164-179)
// thread q of column c: parents[c][q] = hash_with_ds_dynamic([arity, level, q + pos_offset, label_c], children[c][q*arity ..])
struct InChildren {
    const uint32_t* src; int cnt;
    SMW_DEV F29 operator()(uint32_t idx) const {
        if ((int)idx < cnt) return raw_ext(src + (size_t)idx * 8);
        return (int)idx == cnt ? raw_one_ext() : f29_zero();
    }
    SMW_DEV F52 f52(uint32_t idx) const {
        if ((int)idx < cnt) return raw_ext52(src + (size_t)idx * 8);
        return (int)idx == cnt ? raw_one_ext52() : f52_zero();
    }
};
template <int T> __global__ void __launch_bounds__(Cfg<T>::NT, Cfg<T>::MINB)
k_merkle_level(const uint32_t* __restrict__ children, size_t n_children, uint32_t* __restrict__ parents, size_t n_parents,
               int arity, uint32_t level, uint64_t label0, uint64_t label_stride, uint64_t pos_offset,
               size_t child_col_stride, size_t parent_col_stride, Segs sg, int out_first, PoseidonDev P) {
    extern __shared__ __align__(16) unsigned char smem[];
    size_t q = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (q >= n_parents) return;
    const int col = blockIdx.y;
    StateRef S = state_ref<T>(smem, blockDim.x, threadIdx.x);
    zero_lanes(S, T);
    S.store(0, raw_u64((uint64_t)arity)); S.store(1, raw_u64(level)); S.store(2, raw_u64(q + pos_offset)); S.store(3, raw_u64(label0 + label_stride * col));
    size_t base = q * (size_t)arity;
    InChildren in;
    in.cnt = (int)((n_children - base < (size_t)arity) ? (n_children - base) : (size_t)arity);
    in.src = children + ((size_t)col * child_col_stride + base) * 8;
    const int my_seg = 2 * ((4 + in.cnt + 1 + (T - 2)) / (T - 1)) - 1;     // the permutation after which this node's digest is state[0]
    F29 out = f29_zero();
#pragma unroll 1
    for (int s = 0; s < sg.n; s++) {
        run_ops<T>(blockDim.x, P.tbl, P.prog, sg.first[s], sg.count[s], in);
        if (s == my_seg) out = S.load(0);
    }
    S.store(0, out);
    run_ops<T>(blockDim.x, P.tbl, P.prog, out_first, 1, in);
    emit_ext(S.load(0), parents + ((size_t)col * parent_col_stride + q) * 8);
}
// ------------------------------------------------------------------ explicit groups (verification, merkle/src/lib.rs:632-690, 748-765)
// thread g: out[g] = hash_with_ds_dynamic([arity, level, pos[g], label], pool[child_idx[g*row + c]], c < cnt[g] <= row).  The children
// of a recomputed parent are scattered (opened values, siblings, parents of the level below), so they are addressed through
// an index array into one pool; cnt = 2 with level = u32::MAX is the pair-leaf digest.
struct InGathered {
    const uint32_t* pool; const uint64_t* idx; int cnt;
    SMW_DEV F29 operator()(uint32_t i) const {
        if ((int)i < cnt) return raw_ext(pool + (size_t)idx[i] * 8);
        return (int)i == cnt ? raw_one_ext() : f29_zero();
    }
};
template <int T> __global__ void __launch_bounds__(Cfg<T>::NT, Cfg<T>::MINB)
k_hash_groups(const uint32_t* __restrict__ pool, const uint64_t* __restrict__ child_idx, const uint8_t* __restrict__ cnt, const uint64_t* __restrict__ pos,
              size_t n_groups, int arity, int row, uint32_t level, uint64_t label, uint32_t* __restrict__ out, Segs sg, int out_first, PoseidonDev P) {
    extern __shared__ __align__(16) unsigned char smem[];
    size_t g = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (g >= n_groups) return;
    StateRef S = state_ref<T>(smem, blockDim.x, threadIdx.x);
    zero_lanes(S, T);
    S.store(0, raw_u64((uint64_t)arity)); S.store(1, raw_u64(level)); S.store(2, raw_u64(pos[g])); S.store(3, raw_u64(label));
    InGathered in{pool, child_idx + g * (size_t)row, (int)cnt[g]};
    const int my_seg = 2 * ((4 + in.cnt + 1 + (T - 2)) / (T - 1)) - 1;
    F29 o = f29_zero();
#pragma unroll 1
    for (int s = 0; s < sg.n; s++) {
        run_ops<T>(blockDim.x, P.tbl, P.prog, sg.first[s], sg.count[s], in);
        if (s == my_seg) o = S.load(0);
    }
    S.store(0, o);
    run_ops<T>(blockDim.x, P.tbl, P.prog, out_first, 1, in);
    emit_ext(S.load(0), out + g * 8);
}
// pair leaves (merkle/src/lib.rs:382-400): leaf i = H(DS(arity, u32::MAX, i, label), [f_i, cp_{i / cp_div}])
struct InPair {
    const uint32_t* f; const uint32_t* cp;
    SMW_DEV F29 operator()(uint32_t idx) const {
        if (idx == 0) return raw_ext(f);
        if (idx == 1) return cp ? raw_ext(cp) : f29_zero();
        return idx == 2 ? raw_one_ext() : f29_zero();
    }
    SMW_DEV F52 f52(uint32_t idx) const {
        if (idx == 0) return raw_ext52(f);
        if (idx == 1) return cp ? raw_ext52(cp) : f52_zero();
        return idx == 2 ? raw_one_ext52() : f52_zero();
    }
};
template <int T> __global__ void __launch_bounds__(Cfg<T>::NT, Cfg<T>::MINB)
k_pair_leaves(const uint32_t* __restrict__ f, const uint32_t* __restrict__ cp, size_t n, int arity, uint64_t label, uint64_t pos_offset,
              uint32_t* __restrict__ out, size_t cp_div, Segs sg, int out_first, PoseidonDev P) {
    extern __shared__ __align__(16) unsigned char smem[];
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    StateRef S = state_ref<T>(smem, blockDim.x, threadIdx.x);
    zero_lanes(S, T);
    S.store(0, raw_u64((uint64_t)arity)); S.store(1, raw_u64(0xffffffffull)); S.store(2, raw_u64(i + pos_offset)); S.store(3, raw_u64(label));
    InPair in{f + i * 8, cp ? cp + (i / cp_div) * 8 : nullptr};
    run_segs<T>(blockDim.x, P, sg, in);          // 4 + 2 + 1 absorbed lanes <= rate for every width: one permutation
    run_ops<T>(blockDim.x, P.tbl, P.prog, out_first, 1, in);
    emit_ext(S.load(0), out + i * 8);
}
// one hash with arbitrary input length (poseidon/src/lib.rs:288-312), single thread: latency only
template <int T> __global__ void k_hash_ds_single(const uint32_t* ds, int nds, const uint32_t* in, size_t n, uint32_t* out, PoseidonDev P) {
    extern __shared__ __align__(16) unsigned char smem[];
    if (threadIdx.x != 0 || blockIdx.x != 0) return;
    StateRef S = state_ref<T>(smem, 1, 0);
    zero_lanes(S, T);
    int cursor = 0;
    const size_t total = (size_t)nds + n + 1;
#pragma unroll 1
    for (size_t k = 0; k < total; k++) {
        uint32_t w[8];
        if (k < (size_t)nds) load_ext(ds + k * 8, w);
        else if (k < (size_t)nds + n) load_ext(in + (k - nds) * 8, w);
        else {
#pragma unroll
            for (int i = 0; i < 8; i++) w[i] = C_R32W[i];
        }
        S.store(cursor, f29_add(S.load(cursor), f29_from_ext(w)));
        if (++cursor == T - 1) { cursor = 0; run_perm<T>(1, P); }
    }
    if (cursor != 0) run_perm<T>(1, P);              // absorbing zeros only advances the cursor
    uint32_t w[8]; f29_to_ext(S.load(0), w); store_ext(out, w);
}

// ------------------------------------------------------------------ the same kernels on the FP64-pipe multiplier (poseidon52.cuh)
template <int T> struct Cfg52;
#ifdef SMW_T9_NT
template <> struct Cfg52<9>  { static constexpr int NT = SMW_T9_NT, MINB = 2, RUN = 32; };   // tuning builds
#else
template <> struct Cfg52<9>  { static constexpr int NT = 160, MINB = 2, RUN = 32; };     // 204 registers, 10 one-warp CTAs per SM: fold-2 cfg3 483-494 ms against 531-554 ms at 128 registers / 16 warps (profiles/r01z_t9_register_budget.txt)
#endif
#ifndef SMW_T17_NT
#define SMW_T17_NT 288
#endif
template <> struct Cfg52<17> { static constexpr int NT = SMW_T17_NT, MINB = 1, RUN = 32; };     // 19 lanes x 40 B: 9 one-warp CTAs per SM (760 B per thread)
SMW_DEV void zero_lanes52(const StateRef52& S, int n) {
#pragma unroll 1
    for (int i = 0; i < n; i++) S.store(i, f52_zero());
}
template <int T> __global__ void __launch_bounds__(Cfg52<T>::NT, Cfg52<T>::MINB)
k_permute_batch52(uint32_t* states, size_t batch, Segs sg, PoseidonDev P) {
    extern __shared__ __align__(16) unsigned char smem[];
    size_t b = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (b >= batch) return;
    StateRef52 S = state_ref52<T>(smem, blockDim.x, threadIdx.x);
    uint32_t* st = states + b * T * 8;
    InStates in{st};
    run_segs52<T>(blockDim.x, P, sg, in);
#pragma unroll 1
    for (int i = 0; i < T; i++) emit_ext52(S.load(i), st + i * 8);
}
template <int T> __global__ void __launch_bounds__(Cfg52<T>::NT, Cfg52<T>::MINB)
k_merkle_level52(const uint32_t* __restrict__ children, size_t n_children, uint32_t* __restrict__ parents, size_t n_parents,
                 int arity, uint32_t level, uint64_t label0, uint64_t label_stride, uint64_t pos_offset,
                 size_t child_col_stride, size_t parent_col_stride, Segs sg, int out_first, PoseidonDev P) {
    extern __shared__ __align__(16) unsigned char smem[];
    size_t q = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (q >= n_parents) return;
    const int col = blockIdx.y;
    StateRef52 S = state_ref52<T>(smem, blockDim.x, threadIdx.x);
    zero_lanes52(S, T);
    S.store(0, f52_from_u64((uint64_t)arity)); S.store(1, f52_from_u64(level)); S.store(2, f52_from_u64(q + pos_offset)); S.store(3, f52_from_u64(label0 + label_stride * col));
    size_t base = q * (size_t)arity;
    InChildren in;
    in.cnt = (int)((n_children - base < (size_t)arity) ? (n_children - base) : (size_t)arity);
    in.src = children + ((size_t)col * child_col_stride + base) * 8;
    const int my_seg = 2 * ((4 + in.cnt + 1 + (T - 2)) / (T - 1)) - 1;
    F52 out = f52_zero();
#pragma unroll 1
    for (int s = 0; s < sg.n; s++) {
        run_ops52<T>(blockDim.x, P.tbl52, P.prog, sg.first[s], sg.count[s], in);
        if (s == my_seg) out = S.load(0);
    }
    S.store(0, out);
    run_ops52<T>(blockDim.x, P.tbl52, P.prog, out_first, 1, in);
    emit_ext52(S.load(0), parents + ((size_t)col * parent_col_stride + q) * 8);
}
template <int T> __global__ void __launch_bounds__(Cfg52<T>::NT, Cfg52<T>::MINB)
k_pair_leaves52(const uint32_t* __restrict__ f, const uint32_t* __restrict__ cp, size_t n, int arity, uint64_t label, uint64_t pos_offset,
                uint32_t* __restrict__ out, size_t cp_div, Segs sg, int out_first, PoseidonDev P) {
    extern __shared__ __align__(16) unsigned char smem[];
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    StateRef52 S = state_ref52<T>(smem, blockDim.x, threadIdx.x);
    zero_lanes52(S, T);
    S.store(0, f52_from_u64((uint64_t)arity)); S.store(1, f52_from_u64(0xffffffffull)); S.store(2, f52_from_u64(i + pos_offset)); S.store(3, f52_from_u64(label));
    InPair in{f + i * 8, cp ? cp + (i / cp_div) * 8 : nullptr};
    run_segs52<T>(blockDim.x, P, sg, in);
    run_ops52<T>(blockDim.x, P.tbl52, P.prog, out_first, 1, in);
    emit_ext52(S.load(0), out + i * 8);
}

// ------------------------------------------------------------------ several runs in one launch (explicit segment arrays)
// Level j of every layer tree of a FRI commit phase, or the leaves of every layer, are hashed by ONE launch: the layers are
// independent (z_l is public, deep_ali/src/fri.rs:250), and per-tree launches left the machine with hundreds of sub-wave launches.
SMW_DEV int find_seg(const uint64_t g, const uint64_t* firsts, int stride_words, int n_segs) {
    int i = 0;
    while (i + 1 < n_segs && g >= firsts[(size_t)(i + 1) * stride_words]) i++;
    return i;
}
template <int T> __global__ void __launch_bounds__(Cfg52<T>::NT, Cfg52<T>::MINB)
k_merkle_level52_multi(const NodeSeg* __restrict__ segs, int n_segs, uint64_t total, Segs sg, int out_first, PoseidonDev P) {
    extern __shared__ __align__(16) unsigned char smem[];
    const uint64_t g = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (g >= total) return;
    const NodeSeg ns = segs[find_seg(g, &segs[0].first, sizeof(NodeSeg) / 8, n_segs)];
    const uint64_t q = g - ns.first;
    StateRef52 S = state_ref52<T>(smem, blockDim.x, threadIdx.x);
    zero_lanes52(S, T);
    S.store(0, f52_from_u64((uint64_t)ns.arity)); S.store(1, f52_from_u64(ns.level)); S.store(2, f52_from_u64(q + ns.pos_offset)); S.store(3, f52_from_u64(ns.label));
    const uint64_t base = q * ns.arity;
    InChildren in;
    in.cnt = (int)((ns.n_children - base < (uint64_t)ns.arity) ? (ns.n_children - base) : (uint64_t)ns.arity);
    in.src = ns.children + base * 8;
    const int my_seg = 2 * ((4 + in.cnt + 1 + (T - 2)) / (T - 1)) - 1;
    F52 out = f52_zero();
#pragma unroll 1
    for (int s = 0; s < sg.n; s++) {
        run_ops52<T>(blockDim.x, P.tbl52, P.prog, sg.first[s], sg.count[s], in);
        if (s == my_seg) out = S.load(0);
    }
    S.store(0, out);
    run_ops52<T>(blockDim.x, P.tbl52, P.prog, out_first, 1, in);
    emit_ext52(S.load(0), ns.parents + q * 8);
}
template <int T> __global__ void __launch_bounds__(Cfg52<T>::NT, Cfg52<T>::MINB)
k_pair_leaves52_multi(const LeafSeg* __restrict__ segs, int n_segs, uint64_t total, Segs sg, int out_first, PoseidonDev P) {
    extern __shared__ __align__(16) unsigned char smem[];
    const uint64_t g = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (g >= total) return;
    const LeafSeg ls = segs[find_seg(g, &segs[0].first, sizeof(LeafSeg) / 8, n_segs)];
    const uint64_t i = g - ls.first;
    StateRef52 S = state_ref52<T>(smem, blockDim.x, threadIdx.x);
    zero_lanes52(S, T);
    S.store(0, f52_from_u64((uint64_t)ls.arity)); S.store(1, f52_from_u64(0xffffffffull)); S.store(2, f52_from_u64(i + ls.pos_offset)); S.store(3, f52_from_u64(ls.label));
    InPair in{ls.f + i * 8, ls.cp ? ls.cp + (i / ls.cp_div) * 8 : nullptr};
    run_segs52<T>(blockDim.x, P, sg, in);
    run_ops52<T>(blockDim.x, P.tbl52, P.prog, out_first, 1, in);
    emit_ext52(S.load(0), ls.out + i * 8);
}

template <typename K> static int set_smem(smw_ctx* c, K kernel, size_t bytes) {
    SMW_CUDA(c, cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes));
    return SMW_OK;
}

} // namespace smw
