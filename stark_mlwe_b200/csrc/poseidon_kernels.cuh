// poseidon_kernels.cuh — kernel templates shared by the per-width translation units.
#pragma once
#include "smw_internal.h"
#include "poseidon.cuh"

namespace smw {

// launch shape per width: threads per CTA and CTAs per SM so that the per-thread state (T * 36 B)
// fills shared memory: t=9 -> 2 x 256, t=17 -> 2 x 160 (196 KB), larger widths 1 CTA.
template <int T> struct Cfg;
template <> struct Cfg<9>   { static constexpr int NT = 256, MINB = 2; };
template <> struct Cfg<17>  { static constexpr int NT = 160, MINB = 2; };
template <> struct Cfg<33>  { static constexpr int NT = 160, MINB = 1; };
template <> struct Cfg<65>  { static constexpr int NT = 96,  MINB = 1; };
template <> struct Cfg<129> { static constexpr int NT = 32,  MINB = 1; };

static PoseidonDev dev_view(const smw_params* p) {
    PoseidonDev d; d.t = p->t; d.rp = p->rp; d.lu_full = p->lu_full; d.rc_full = p->rc_full; d.rc_part = p->rc_part;
    d.sp_row = p->sp_row; d.sp_col = p->sp_col; d.lu_post = p->lu_post; return d;
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

// ------------------------------------------------------------------ batched permutation
template <int T, bool SPARSE> __global__ void __launch_bounds__(Cfg<T>::NT, Cfg<T>::MINB)
k_permute_batch(uint32_t* states, size_t batch, PoseidonDev P) {
    extern __shared__ __align__(16) unsigned char smem[];
    size_t b = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (b >= batch) return;
    StateRef S = state_ref<T>(smem, blockDim.x, threadIdx.x);
    uint32_t* st = states + b * T * 8;
#pragma unroll 1
    for (int i = 0; i < T; i++) { uint32_t w[8]; load_ext(st + i * 8, w); S.store(i, f29_from_ext(w)); }
    permute<T, SPARSE>(S, P);
#pragma unroll 1
    for (int i = 0; i < T; i++) { uint32_t w[8]; f29_to_ext(S.load(i), w); store_ext(st + i * 8, w); }
}

// ------------------------------------------------------------------ Merkle level (merkle/src/lib.rs:164-179)
// thread q of column c: parents[c][q] = hash_with_ds_dynamic([arity, level, q + pos_offset, label_c], children[c][q*arity ..])
template <int T> __global__ void __launch_bounds__(Cfg<T>::NT, Cfg<T>::MINB)
k_merkle_level(const uint32_t* __restrict__ children, size_t n_children, uint32_t* __restrict__ parents, size_t n_parents,
               int arity, uint32_t level, uint64_t label0, uint64_t label_stride, uint64_t pos_offset,
               size_t child_col_stride, size_t parent_col_stride, PoseidonDev P) {
    extern __shared__ __align__(16) unsigned char smem[];
    size_t q = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (q >= n_parents) return;
    const int col = blockIdx.y;
    StateRef S = state_ref<T>(smem, blockDim.x, threadIdx.x);
    Sponge<T, true> sp(S, P);
    sp.absorb(f29_from_u64((uint64_t)arity)); sp.absorb(f29_from_u64(level));
    sp.absorb(f29_from_u64(q + pos_offset)); sp.absorb(f29_from_u64(label0 + label_stride * col));
    size_t base = q * (size_t)arity;
    int cnt = (int)((n_children - base < (size_t)arity) ? (n_children - base) : (size_t)arity);
    const uint32_t* src = children + ((size_t)col * child_col_stride + base) * 8;
#pragma unroll 1
    for (int j = 0; j < cnt; j++) { uint32_t w[8]; load_ext(src + j * 8, w); sp.absorb(f29_from_ext(w)); }
    uint32_t w[8]; f29_to_ext(sp.finish(), w);
    store_ext(parents + ((size_t)col * parent_col_stride + q) * 8, w);
}
// pair leaves (merkle/src/lib.rs:382-400): leaf i = H(DS(arity, u32::MAX, i, label), [f_i, cp_{i / cp_div}])
template <int T> __global__ void __launch_bounds__(Cfg<T>::NT, Cfg<T>::MINB)
k_pair_leaves(const uint32_t* __restrict__ f, const uint32_t* __restrict__ cp, size_t n, int arity, uint64_t label, uint64_t pos_offset,
              uint32_t* __restrict__ out, size_t cp_div, PoseidonDev P) {
    extern __shared__ __align__(16) unsigned char smem[];
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    StateRef S = state_ref<T>(smem, blockDim.x, threadIdx.x);
    Sponge<T, true> sp(S, P);
    sp.absorb(f29_from_u64((uint64_t)arity)); sp.absorb(f29_from_u64(0xffffffffull));
    sp.absorb(f29_from_u64(i + pos_offset)); sp.absorb(f29_from_u64(label));
    uint32_t w[8];
    load_ext(f + i * 8, w); sp.absorb(f29_from_ext(w));
    if (cp) { load_ext(cp + (i / cp_div) * 8, w); sp.absorb(f29_from_ext(w)); } else sp.absorb(f29_zero());
    f29_to_ext(sp.finish(), w); store_ext(out + i * 8, w);
}
// one hash with arbitrary input length (poseidon/src/lib.rs:288-312), single thread
template <int T> __global__ void k_hash_ds_single(const uint32_t* ds, int nds, const uint32_t* in, size_t n, uint32_t* out, PoseidonDev P) {
    extern __shared__ __align__(16) unsigned char smem[];
    if (threadIdx.x != 0 || blockIdx.x != 0) return;
    StateRef S = state_ref<T>(smem, 1, 0);
    Sponge<T, true> sp(S, P);
    uint32_t w[8];
    for (int i = 0; i < nds; i++) { load_ext(ds + i * 8, w); sp.absorb(f29_from_ext(w)); }
    for (size_t i = 0; i < n; i++) { load_ext(in + i * 8, w); sp.absorb(f29_from_ext(w)); }
    f29_to_ext(sp.finish(), w); store_ext(out, w);
}

template <typename K> static int set_smem(smw_ctx* c, K kernel, size_t bytes) {
    SMW_CUDA(c, cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes));
    return SMW_OK;
}

} // namespace smw
