// poseidon_coop_kernels.cuh — kernels over the cooperative permutation (poseidon_coop52.cuh): Merkle nodes of launches
// below one wave (tree tops, the small layers of a FRI commit phase), the transcript (serial sponge, challenges), z sampling.
// Included by poseidon_T.cu for the widths that carry FP64-pipe tables (t = 9, 17).
#pragma once
#include "poseidon_kernels.cuh"
#include "poseidon_coop52.cuh"

namespace smw {

template <int T> struct CoopCfg;
template <> struct CoopCfg<9>  { static constexpr int G = 16; };      // two permutations per warp
template <> struct CoopCfg<17> { static constexpr int G = 32; };

// node g of a launch -> its segment and index inside it.  Explicit segments: linear scan over the (few) first-node offsets;
// uniform columns: segment = column.
SMW_DEV bool resolve_node(const LevelBatch& B, uint64_t g, NodeSeg& s, uint64_t& q) {
    if (g >= B.total) return false;
    if (B.segs) {
        int i = 0;
        while (i + 1 < B.n_segs && g >= B.segs[i + 1].first) i++;
        s = B.segs[i]; q = g - s.first;
    } else {
        const uint64_t c = g / B.u.n_parents; q = g - c * B.u.n_parents;
        s = B.u; s.children += c * B.child_col_stride * 8; s.parents += c * B.parent_col_stride * 8; s.label += c * B.label_stride;
    }
    return true;
}

SMW_DEV F52 raw_one_int52() { F52 r = f52_zero(); r.v[0] = 1.0; return r; }

// ------------------------------------------------------------------ Merkle nodes, one per lane group (merkle/src/lib.rs:164-179)
// parents[q] = hash_with_ds_dynamic([arity, level, q + pos_offset, label], children[q * arity ..]) (poseidon/src/lib.rs:288-312):
// lane i < t - 1 absorbs input i of every rate block (DS integers, children, the padding one), then the group permutes.
template <int T> __global__ void __launch_bounds__(32)
k_nodes_coop52(LevelBatch B, const uint32_t* __restrict__ tbl, CoopTab L) {
    constexpr int G = CoopCfg<T>::G, NG = 32 / G, R = T - 1;
    __shared__ __align__(16) double sh[NG][T * 6];
    const int group = threadIdx.x / G, lane = threadIdx.x % G;
    CoopCx<T, G> c{tbl, L, sh[group], lane, group};
    const uint64_t g = (uint64_t)blockIdx.x * NG + group;
    NodeSeg sg; uint64_t q = 0;
    const bool valid = resolve_node(B, g, sg, q);
    int cnt = 0; const uint32_t* src = nullptr;
    if (valid) {
        const uint64_t base = q * sg.arity;
        cnt = (int)((sg.n_children - base < (uint64_t)sg.arity) ? (sg.n_children - base) : (uint64_t)sg.arity);
        src = sg.children + base * 8;
    }
    const int nb = valid ? (4 + cnt + 1 + R - 1) / R : 0;
    int nb_max = nb;
#pragma unroll
    for (int d = G; d < 32; d <<= 1) nb_max = max(nb_max, __shfl_xor_sync(0xffffffffu, nb_max, d));   // the groups of a warp run in lockstep
    const F52 k_in = c.tab(L.k_in), k_rr = c.tab(L.k_rr);
    F52 s = f52_zero(), out = f52_zero();
#pragma unroll 1
    for (int b = 0; b < nb_max; b++) {
        const int e = b * R + lane;                        // input index of this lane in rate block b
        F52 raw = f52_zero(), cst = k_in, add = s;
        if (valid && lane < R) {
            if (e < 4) { cst = k_rr; raw = f52_from_u64(e == 0 ? (uint64_t)sg.arity : e == 1 ? (uint64_t)sg.level : e == 2 ? q + sg.pos_offset : sg.label); }
            else if (e < 4 + cnt) raw = raw_ext52(src + (size_t)(e - 4) * 8);
            else if (e == 4 + cnt) { raw = raw_one_int52(); cst = k_rr; }        // the padding one: 1 * 2^520 / 2^260
        }
        s = coop_mr(raw, cst, add);
        s = coop_perm<T, G>(c, s);
        if (b == nb - 1) out = s;
    }
    if (valid && lane == 0) emit_ext52(coop_mr(out, c.tab(L.k_out), f52_zero()), sg.parents + q * 8);
}

// batched permutation, one lane group per state (same result as k_permute_batch52; the latency form, used for small batches)
template <int T> __global__ void __launch_bounds__(32)
k_permute_coop52(uint32_t* states, size_t batch, const uint32_t* __restrict__ tbl, CoopTab L) {
    constexpr int G = CoopCfg<T>::G, NG = 32 / G;
    __shared__ __align__(16) double sh[NG][T * 6];
    const int group = threadIdx.x / G, lane = threadIdx.x % G;
    CoopCx<T, G> c{tbl, L, sh[group], lane, group};
    const size_t b = (size_t)blockIdx.x * NG + group;
    const bool act = b < batch && lane < T;
    F52 raw = f52_zero();
    if (act) raw = raw_ext52(states + (b * T + lane) * 8);
    F52 s = coop_mr(raw, c.tab(L.k_in), f52_zero());
    s = coop_perm<T, G>(c, s);
    const F52 e = coop_mr(s, c.tab(L.k_out), f52_zero());
    if (act) emit_ext52(e, states + (b * T + lane) * 8);
}

#if SMW_T == 17
// ------------------------------------------------------------------ transcript (transcript/src/lib.rs:55-101) on one warp
// Same contract as k_transcript_run: [fresh: state = 0, lane 16 = INIT tag (immA[0])] ; absorb immA[fresh..nA) ; absorb extB[0..nB) ;
// absorb immC[0..nC) ; [final permute] ; state -> `state` (if given), state[0] -> `out` (if given).  Immediates are 256-bit
// little-endian integers (reduced mod p by the multiplication with 2^520), extB elements are in the external Montgomery form.
__global__ void __launch_bounds__(32)
k_transcript_coop52(uint32_t* state, int fresh, int pos, const uint32_t* __restrict__ immA, int nA, const uint32_t* __restrict__ extB, size_t nB,
                    const uint32_t* __restrict__ immC, int nC, int final_permute, uint32_t* out, const uint32_t* __restrict__ tbl, CoopTab L) {
    __shared__ __align__(16) double sh[17 * 6];
    const int lane = threadIdx.x;
    CoopCx<17, 32> c{tbl, L, sh, lane, 0};
    const F52 k_in = c.tab(L.k_in), k_rr = c.tab(L.k_rr), zero = f52_zero();
    F52 s = zero;
    int a0 = 0;
    if (fresh) { F52 raw = zero; if (lane == 16) raw = raw_ext52(immA); s = coop_mr(raw, k_rr, zero); a0 = 1; pos = 0; }
    else { F52 raw = zero; if (lane < 17) raw = raw_ext52(state + lane * 8); s = coop_mr(raw, k_in, zero); }
    const size_t nA1 = (size_t)(nA - a0), total = nA1 + nB + (size_t)nC;
    size_t q = 0;
#pragma unroll 1
    while (q < total) {
        if (pos == 16) { s = coop_perm<17, 32>(c, s); pos = 0; }
        const int cnt = (int)((total - q < (size_t)(16 - pos)) ? (total - q) : (size_t)(16 - pos));
        F52 raw = zero, cst = k_in;
        if (lane >= pos && lane < pos + cnt) {
            const size_t i = q + (size_t)(lane - pos);
            if (i < nA1) { raw = raw_ext52(immA + ((size_t)a0 + i) * 8); cst = k_rr; }
            else if (i < nA1 + nB) raw = raw_ext52(extB + (i - nA1) * 8);
            else { raw = raw_ext52(immC + (i - nA1 - nB) * 8); cst = k_rr; }
        }
        s = coop_mr(raw, cst, s);
        pos += cnt; q += (size_t)cnt;
    }
    if (final_permute) s = coop_perm<17, 32>(c, s);
    const F52 e = coop_mr(s, c.tab(L.k_out), zero);
    if (lane < 17 && state) emit_ext52(e, state + lane * 8);
    if (lane == 0 && out) emit_ext52(e, out);
}

// ------------------------------------------------------------------ legacy trees (merkle/src/lib.rs:196-223, 373-375, 448-470, 487-584)
// out[g] = hash_with_ds(children of g, ds_tag) (poseidon/src/lib.rs:85-100: state[16] = ds_tag, children added into lanes 0.., ONE
// permutation, no padding) — one warp per node.  Children: mode 0 = src[16 g .. 16 g + 16) clipped to n_src; mode 1 = the pair
// (src[g], cp[g]); mode 2 = pool entries src[idx[row g + c]], c < cnt[g] (the recomputed parents of a verifier; n_src carries `row`).
__global__ void __launch_bounds__(32)
k_legacy_nodes_coop52(int mode, const uint32_t* __restrict__ src, const uint32_t* __restrict__ cp, size_t n_src, const uint64_t* __restrict__ idx,
                      const uint8_t* __restrict__ cnt, size_t n_out, const uint32_t* __restrict__ tag, uint32_t* __restrict__ out, const uint32_t* __restrict__ tbl, CoopTab L) {
    __shared__ __align__(16) double sh[17 * 6];
    const int lane = threadIdx.x; const size_t g = blockIdx.x;
    if (g >= n_out) return;
    CoopCx<17, 32> c{tbl, L, sh, lane, 0};
    F52 raw = f52_zero();
    if (lane == 16) raw = raw_ext52(tag);
    else if (mode == 0) { const size_t i = g * 16 + lane; if (lane < 16 && i < n_src) raw = raw_ext52(src + i * 8); }
    else if (mode == 1) { if (lane == 0) raw = raw_ext52(src + g * 8); else if (lane == 1) raw = raw_ext52(cp + g * 8); }
    else { if (lane < (int)cnt[g] && lane < 16) raw = raw_ext52(src + (size_t)idx[g * n_src + lane] * 8); }
    F52 s = coop_mr(raw, c.tab(L.k_in), f52_zero());
    s = coop_perm<17, 32>(c, s);
    const F52 e = coop_mr(s, c.tab(L.k_out), f52_zero());
    if (lane == 0) emit_ext52(e, out + g * 8);
}

// `count` one-shot transcript hashes that share label, tag and challenge and differ in their k field elements
// (tr_hash_fields_tagged, deep_ali/src/fri.rs:28-36, e.g. every index seed of the query phase, :184-191), one warp each: the
// pre-image template (tags and label words, 256-bit little-endian integers; all of it fits the 16-lane rate, so ONE permutation)
// with the item's k elements dropped into lanes lane0 .. lane0 + k - 1.
__global__ void __launch_bounds__(32)
k_tr_hash_batch_coop52(const uint32_t* __restrict__ tmpl_le256, int lane0, int k, const uint32_t* __restrict__ xs, size_t count, uint32_t* __restrict__ out,
                       const uint32_t* __restrict__ tbl, CoopTab L) {
    __shared__ __align__(16) double sh[17 * 6];
    const int lane = threadIdx.x; const size_t g = blockIdx.x;
    if (g >= count) return;
    CoopCx<17, 32> c{tbl, L, sh, lane, 0};
    F52 raw = f52_zero(), cst = c.tab(L.k_rr);
    if (lane >= lane0 && lane < lane0 + k) { raw = raw_ext52(xs + (g * (size_t)k + (size_t)(lane - lane0)) * 8); cst = c.tab(L.k_in); }
    else if (lane < 17) raw = raw_ext52(tmpl_le256 + (size_t)lane * 8);
    F52 s = coop_mr(raw, cst, f52_zero());
    s = coop_perm<17, 32>(c, s);
    const F52 e = coop_mr(s, c.tab(L.k_out), f52_zero());
    if (lane == 0) emit_ext52(e, out + g * 8);
}

// fri_sample_z_ell (deep_ali/src/fri.rs:59-82) for all L levels at once, one warp per level: the warp hashes the level's template
// (one permutation), lane 0 seeds ChaCha12 with the canonical bytes of the digest and takes the first u64 c with c != 0, c^n != 1.
__global__ void __launch_bounds__(32)
k_fri_sample_z_coop52(const uint32_t* __restrict__ tmpl_le256, const uint64_t* __restrict__ level_size, int nL, uint64_t seed_z, uint32_t* z_out,
                      const uint32_t* __restrict__ tbl, CoopTab L) {
    __shared__ __align__(16) double sh[17 * 6];
    const int lane = threadIdx.x, l = blockIdx.x;
    if (l >= nL) return;
    CoopCx<17, 32> c{tbl, L, sh, lane, 0};
    F52 raw = f52_zero();
    if (lane < 17) raw = raw_ext52(tmpl_le256 + ((size_t)l * 17 + lane) * 8);
    F52 s = coop_mr(raw, c.tab(L.k_rr), f52_zero());
    s = coop_perm<17, 32>(c, s);
    const F52 d = coop_mr(s, raw_one_int52(), f52_zero());          // x 2^260 * 1 / 2^260: the canonical integer of the digest (after the subtraction below)
    if (lane != 0) return;
    uint64_t dl[5]; f52_canonical_limbs(d, dl);
    uint32_t key[8]; f52_limbs_to32(dl, key);
    uint32_t w[8];
    const uint64_t level = level_size[2 * l], n = level_size[2 * l + 1];
    ChaChaRng rng; chacha_init(rng, key);
    F29 z = f29_zero(); bool found = false;
    for (int tries = 0; tries < 1000 && !found; tries++) {
        const uint64_t cc = chacha_next_u64(rng);
        if (cc == 0) continue;
        const F29 cand = f29_from_u64(cc);
        F29 pw = f29_const(C_ONE29);
        for (int b = 63; b >= 0; b--) { pw = f29_sqr(pw); if ((n >> b) & 1ull) pw = f29_mul(pw, cand); }
        const F29 diff = f29_sub_canon(f29_full_reduce(pw), f29_const(C_ONE29));
        if (!f29_is_zero_canon(diff)) { z = cand; found = true; }
    }
    if (!found) {   // fallbacks of fri.rs:75-79
        const F29 fb = f29_from_u64(seed_z + level + 7ull);
        F29 pw = f29_const(C_ONE29);
        for (int b = 63; b >= 0; b--) { pw = f29_sqr(pw); if ((n >> b) & 1ull) pw = f29_mul(pw, fb); }
        const F29 diff = f29_sub_canon(f29_full_reduce(pw), f29_const(C_ONE29));
        z = f29_is_zero_canon(diff) ? f29_from_u64(11) : fb;
    }
    f29_to_ext(z, w); store_ext(z_out + (size_t)l * 8, w);
}
#endif

} // namespace smw
