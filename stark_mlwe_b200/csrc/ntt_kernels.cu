// ntt_kernels.cu — radix-2 NTT / inverse NTT / coset LDE over Pallas Fr.
//
// Convention (fft/src/lib.rs:6-32 -> ark-poly 0.5 Radix2EvaluationDomain): fft takes coefficients in
// natural order and returns evals[i] = P(w^i) in natural order; ifft is its inverse including 1/n.
// LDE: c = ifft_n(col); c[i] *= g^i; zero-pad to beta*n; fft_{beta n}.
//
// Structure: a size-2^L transform is split into passes of <= 8 butterfly stages; a CTA stages a tile
// of 1024 elements (32 KB) in shared memory, gathered from global memory in runs of 2^c contiguous
// elements (>= 128 B) and runs its stages there.  The inverse transform is decimation-in-frequency
// (natural in, bit-reversed out) and the forward one decimation-in-time (bit-reversed in, natural
// out), so the LDE needs no bit-reversal pass at all: the coset scaling is applied in bit-reversed
// index space and the zero padding becomes a stride.  Only the stand-alone smw_ntt API pays for a
// bit-reversal permutation.  Twiddles w_N^i (i < N/2) are precomputed once per context.
#include "smw_internal.h"
#include "fr32.cuh"
#include "fr29.cuh"
#include <cstdlib>
#include <algorithm>
#include <vector>

namespace smw {

constexpr int TILE_LOG = 10;

// pow2[k] = primitive 2^k-th root of unity, k = 0..32 (ark-pallas FrConfig: generator 5, two-adicity 32)
__global__ void k_pow2_roots(uint32_t* pow2) {
    if (threadIdx.x != 0 || blockIdx.x != 0) return;
    // (p - 1) / 2^32
    const uint32_t e[7] = {0x8c46eb21u, 0x0994a8ddu, 0x224698fcu, 0u, 0u, 0u, 0x40000000u};
    Fr g = fr_from_u64(5), r = fr_const(C_R32);
    for (int k = 6; k >= 0; k--) for (int b = 31; b >= 0; b--) { r = fr_sqr(r); if ((e[k] >> b) & 1u) r = fr_mul(r, g); }
    for (int k = 32; k >= 0; k--) { fr_store(pow2 + k * 8, r); r = fr_sqr(r); }
}
// tw[i] = w_N^i, i < N/2
__global__ void k_twiddles(uint32_t* tw, int log_n, const uint32_t* pow2) {
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    size_t half = (size_t)1 << (log_n - 1);
    if (i >= half) return;
    Fr r = fr_const(C_R32);
    for (int b = 0; b < log_n - 1; b++) if ((i >> b) & 1) r = fr_mul(r, fr_load(pow2 + (log_n - b) * 8));
    fr_store(tw + i * 8, r);
}

struct PassArgs {
    int L;            // transform size 2^L
    int lo;           // lowest global bit covered by the tile's x field
    int span;         // number of x bits held in the tile
    int c;            // contiguous low bits held in the tile (0 when lo == 0)
    int s_begin, s_end;   // butterfly stages [s_begin, s_end) as global bit positions
    int tile_log;     // log2(elements per tile) = span + c
    int tw_shift;     // table index scale: log2(N_table / N)
    int inverse;      // use w^-1
    size_t col_stride;
};
SMW_DEV size_t tile_global(const PassArgs& a, size_t tile, int local) {
    int y = local & ((1 << a.c) - 1), x = local >> a.c;
    int lo_bits = a.lo - a.c;                       // tile bits that sit between y and x
    size_t t_lo = tile & (((size_t)1 << lo_bits) - 1), t_hi = tile >> lo_bits;
    return (t_hi << (a.lo + a.span)) | ((size_t)x << a.lo) | (t_lo << a.c) | (size_t)y;
}
SMW_DEV Fr twiddle(const uint32_t* tw, const PassArgs& a, size_t e /* exponent of w_N, < N/2 */) {
    size_t half = (size_t)1 << (a.L - 1);
    if (!a.inverse) return fr_ldg(tw + ((e << a.tw_shift) * 8));
    if (e == 0) return fr_const(C_R32);
    return fr_neg(fr_ldg(tw + (((half - e) << a.tw_shift) * 8)));     // w^-e = -w^(N/2 - e)
}
// DIF = true : stages run from s_end-1 down to s_begin, butterfly (u+v, (u-v)*w)
// DIF = false: stages run from s_begin up to s_end-1,   butterfly (u+v*w, u-v*w)
template <bool DIF> __global__ void __launch_bounds__(512) k_ntt_pass(uint32_t* data, const uint32_t* tw, PassArgs a) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    uint4* s_lo = reinterpret_cast<uint4*>(smem_raw);
    uint4* s_hi = s_lo + (1 << a.tile_log);
    uint32_t* col = data + (size_t)blockIdx.y * a.col_stride * 8;
    const int tile_elems = 1 << a.tile_log, nthr = blockDim.x;
    const size_t tile = blockIdx.x;
    for (int l = threadIdx.x; l < tile_elems; l += nthr) {
        const uint4* g = reinterpret_cast<const uint4*>(col + tile_global(a, tile, l) * 8);
        s_lo[l] = g[0]; s_hi[l] = g[1];
    }
    __syncthreads();
    const int nstages = a.s_end - a.s_begin;
    for (int st = 0; st < nstages; st++) {
        const int s = DIF ? (a.s_end - 1 - st) : (a.s_begin + st);
        const int lb = s - a.lo + a.c;                 // local bit position of global bit s
        for (int b = threadIdx.x; b < tile_elems / 2; b += nthr) {
            int i0 = ((b >> lb) << (lb + 1)) | (b & ((1 << lb) - 1)), i1 = i0 | (1 << lb);
            size_t gi = tile_global(a, tile, i0);
            size_t e = (gi & (((size_t)1 << s) - 1)) << (a.L - 1 - s);
            Fr w = twiddle(tw, a, e);
            uint4 a0 = s_lo[i0], a1 = s_hi[i0], b0 = s_lo[i1], b1 = s_hi[i1];
            Fr u, v;
            u.v[0] = a0.x; u.v[1] = a0.y; u.v[2] = a0.z; u.v[3] = a0.w; u.v[4] = a1.x; u.v[5] = a1.y; u.v[6] = a1.z; u.v[7] = a1.w;
            v.v[0] = b0.x; v.v[1] = b0.y; v.v[2] = b0.z; v.v[3] = b0.w; v.v[4] = b1.x; v.v[5] = b1.y; v.v[6] = b1.z; v.v[7] = b1.w;
            Fr x, y;
            if (DIF) { x = fr_add(u, v); y = fr_mul(fr_sub(u, v), w); }
            else { Fr vw = fr_mul(v, w); x = fr_add(u, vw); y = fr_sub(u, vw); }
            s_lo[i0] = make_uint4(x.v[0], x.v[1], x.v[2], x.v[3]); s_hi[i0] = make_uint4(x.v[4], x.v[5], x.v[6], x.v[7]);
            s_lo[i1] = make_uint4(y.v[0], y.v[1], y.v[2], y.v[3]); s_hi[i1] = make_uint4(y.v[4], y.v[5], y.v[6], y.v[7]);
        }
        __syncthreads();
    }
    for (int l = threadIdx.x; l < tile_elems; l += nthr) {
        uint4* g = reinterpret_cast<uint4*>(col + tile_global(a, tile, l) * 8);
        g[0] = s_lo[l]; g[1] = s_hi[l];
    }
}
// ------------------------------------------------------------------ the same pass on the carry-free multiplier (round 2)
// k_ntt_pass above spends ~850 instructions per butterfly: the 8 x 32-bit CIOS product compiles to IMAD.WIDE + IADD3 / IADD3.X carry
// pairs + shifts (322 IMAD, 248 IADD3, 91 SHF in its SASS).  Here a tile element is nine 29-bit limbs, LAZILY reduced, in the
// caller's scale (x 2^256): the product with a twiddle kept in the internal form (w 2^261) is 81 independent IMAD.WIDE into 17 column
// accumulators + a 45-IMAD.WIDE Montgomery reduction (-p^-1 = -1 mod 2^29: the quotient digit is a negation), fr29.cuh — and comes
// back in the caller's scale; additions are limb-wise, subtractions add a multiple of p whose limbs dominate the subtrahend's.
//   DIT  x = u + t, y = u + 4p - t, t = v w < 1.5 p : a value grows by <= 4p per stage, < 33p < 2^260 after <= 8 stages
//   DIF  x = u + v doubles per stage: every fourth stage a multiple of p read off the top limb is taken out (< 2p again, so a
//        value stays <= 32p); y = (u + 32p - v) w < 1.2 p
// A pass loads canonical elements and stores canonical elements (top-limb reduction + one conditional subtraction), so the
// kernels between the passes (expansion, bit reversal) and the caller see the same bytes as before: results are bit-identical.
constexpr int TW29 = 12;      // twiddle stride in words (9 used; 48 B keeps the 16-byte loads aligned)
struct Tile29 {
    uint4* q0; uint4* q1; uint32_t* w;
    SMW_DEV F29 load(int i) const { const uint4 a = q0[i], b = q1[i]; F29 r; r.v[0] = a.x; r.v[1] = a.y; r.v[2] = a.z; r.v[3] = a.w; r.v[4] = b.x; r.v[5] = b.y; r.v[6] = b.z; r.v[7] = b.w; r.v[8] = w[i]; return r; }
    SMW_DEV void store(int i, const F29& x) const { q0[i] = make_uint4(x.v[0], x.v[1], x.v[2], x.v[3]); q1[i] = make_uint4(x.v[4], x.v[5], x.v[6], x.v[7]); w[i] = x.v[8]; }
};
SMW_DEV F29 tw29_load(const uint32_t* tw, size_t idx) {
    const uint4* q = reinterpret_cast<const uint4*>(tw + idx * TW29);
    const uint4 a = __ldg(q), b = __ldg(q + 1); F29 r;
    r.v[0] = a.x; r.v[1] = a.y; r.v[2] = a.z; r.v[3] = a.w; r.v[4] = b.x; r.v[5] = b.y; r.v[6] = b.z; r.v[7] = b.w; r.v[8] = __ldg(reinterpret_cast<const uint32_t*>(q + 2));
    return r;
}
// u + K p - t with the limbs carried back below 2^29.  K p is written with a borrow lent to every limb (d_0 = e_0 + 2^29,
// d_i = e_i + 2^29 - 1 for 0 < i < 8, d_8 = e_8 - 1, e = the normalised limbs of K p), so that every d_i dominates t_i:
// t needs limbs < 2^29 and a top limb <= d_8, i.e. t < 3.9 p for K = 4 and t < 31.9 p for K = 32.
template <int K> SMW_DEV F29 f29_sub_lazy(const F29& u, const F29& t) {
    static_assert(K == 4 || K == 32, "tables exist for 4p and 32p");
    constexpr uint32_t D4[9]  = {0x20000004u, 0x28dd641fu, 0x34a8dd8bu, 0x34c7e04bu, 0x20000890u, 0x1fffffffu, 0x1fffffffu, 0x1fffffffu, 0x00ffffffu};
    constexpr uint32_t D32[9] = {0x20000020u, 0x26eb20ffu, 0x2546ec61u, 0x263f0264u, 0x2000448cu, 0x1fffffffu, 0x1fffffffu, 0x1fffffffu, 0x07ffffffu};
    F29 r; uint32_t c = 0;
#pragma unroll
    for (int i = 0; i < 8; i++) { const uint32_t s = u.v[i] + ((K == 4 ? D4[i] : D32[i]) - t.v[i]) + c; r.v[i] = s & M29; c = s >> 29; }
    r.v[8] = u.v[8] + ((K == 4 ? D4[8] : D32[8]) - t.v[8]) + c;
    return r;
}
// x - (q - 1) p with q = floor(x / 2^254) read off the top limb (0 when q = 0): the result is < 2 * 2^254 < 2p and >= 0
SMW_DEV F29 f29_top_reduce(const F29& x) {
    uint32_t q = x.v[8] >> 22; q = q ? q - 1u : 0u;
    const uint32_t pl[9] = {1u, P29_1, P29_2, P29_3, P29_4, 0u, 0u, 0u, P29_8};
    F29 r; long long c = 0;
#pragma unroll
    for (int i = 0; i < 8; i++) { const long long s = (long long)x.v[i] - (long long)((uint64_t)q * pl[i]) + c; r.v[i] = (uint32_t)s & M29; c = s >> 29; }
    r.v[8] = (uint32_t)((long long)x.v[8] - (long long)((uint64_t)q * pl[8]) + c);
    return r;
}
template <bool DIF> __global__ void __launch_bounds__(512) k_ntt_pass29(uint32_t* data, const uint32_t* __restrict__ tw, PassArgs a) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    const int tile_elems = 1 << a.tile_log, nthr = blockDim.x;
    Tile29 S{reinterpret_cast<uint4*>(smem_raw), reinterpret_cast<uint4*>(smem_raw) + tile_elems, reinterpret_cast<uint32_t*>(reinterpret_cast<uint4*>(smem_raw) + 2 * tile_elems)};
    uint32_t* col = data + (size_t)blockIdx.y * a.col_stride * 8;
    const size_t tile = blockIdx.x, half = (size_t)1 << (a.L - 1);
    for (int l = threadIdx.x; l < tile_elems; l += nthr) {
        uint32_t w[8]; const uint4* g = reinterpret_cast<const uint4*>(col + tile_global(a, tile, l) * 8);
        const uint4 x = g[0], y = g[1];
        w[0] = x.x; w[1] = x.y; w[2] = x.z; w[3] = x.w; w[4] = y.x; w[5] = y.y; w[6] = y.z; w[7] = y.w;
        S.store(l, f29_repack_from32(w));
    }
    __syncthreads();
    const int nstages = a.s_end - a.s_begin;
    for (int st = 0; st < nstages; st++) {
        const int s = DIF ? (a.s_end - 1 - st) : (a.s_begin + st);
        const int lb = s - a.lo + a.c;                 // local bit position of global bit s
        for (int b = threadIdx.x; b < tile_elems / 2; b += nthr) {
            const int i0 = ((b >> lb) << (lb + 1)) | (b & ((1 << lb) - 1)), i1 = i0 | (1 << lb);
            const size_t gi = tile_global(a, tile, i0);
            const size_t e = (size_t)(((uint32_t)gi & ((1u << s) - 1u)) << (a.L - 1 - s));     // s <= L - 1 <= 31: the masked index and the exponent (< N / 2) fit 32 bits
            // forward: w^e.  inverse: w^-e = -w^(N/2 - e) for e > 0 — the sign goes into the order of the subtraction.
            // (Requesting the twiddle one butterfly ahead was tried: 84.5 ms against 81.0 ms per cfg2 step, profiles/r02q_ntt_ab.txt —
            // four resident CTAs per SM already cover the L2 latency, the extra registers only cost occupancy.)
            const bool neg = a.inverse && e != 0;
            const F29 w = tw29_load(tw, (neg ? half - e : e) << a.tw_shift);
            const F29 u = S.load(i0), v = S.load(i1);
            if (DIF) {
                F29 x = f29_add(u, v);
                if ((st & 3) == 3) x = f29_top_reduce(x);                 // inputs <= 16p here; uniform over the CTA
                const F29 d = neg ? f29_sub_lazy<32>(v, u) : f29_sub_lazy<32>(u, v);
                Acc A; acc_mul(A, d, w);
                S.store(i0, x); S.store(i1, acc_reduce(A));
            } else {
                Acc A; acc_mul(A, v, w);
                const F29 t = acc_reduce(A);
                const F29 p = f29_add(u, t), m = f29_sub_lazy<4>(u, t);
                S.store(i0, neg ? m : p); S.store(i1, neg ? p : m);
            }
        }
        __syncthreads();
    }
    for (int l = threadIdx.x; l < tile_elems; l += nthr) {
        uint32_t w[8]; f29_repack_to32(f29_canonicalize(f29_top_reduce(S.load(l))), w);
        uint4* g = reinterpret_cast<uint4*>(col + tile_global(a, tile, l) * 8);
        g[0] = make_uint4(w[0], w[1], w[2], w[3]); g[1] = make_uint4(w[4], w[5], w[6], w[7]);
    }
}
// ext twiddles (w 2^256, canonical) -> internal form (w 2^261, canonical limbs), 48-byte stride
__global__ void k_twiddles_to29(const uint32_t* ext, size_t n, uint32_t* out) {
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    uint32_t w[8]; const uint4* g = reinterpret_cast<const uint4*>(ext + i * 8);
    const uint4 x = g[0], y = g[1];
    w[0] = x.x; w[1] = x.y; w[2] = x.z; w[3] = x.w; w[4] = y.x; w[5] = y.y; w[6] = y.z; w[7] = y.w;
    const F29 r = f29_canonicalize(f29_from_ext(w));
    uint32_t* q = out + i * TW29;
#pragma unroll
    for (int k = 0; k < 9; k++) q[k] = r.v[k];
    q[9] = 0; q[10] = 0; q[11] = 0;
}
static bool ntt_use29() { static int v = -1; if (v < 0) { const char* e = getenv("SMW_NTT"); v = (e && atoi(e) == 32) ? 0 : 1; } return v == 1; }   // SMW_NTT=32: the CIOS kernel (A/B)

SMW_DEV size_t bitrev(size_t x, int bits) { return bits == 0 ? 0 : (size_t)(__brevll((unsigned long long)x) >> (64 - bits)); }
// in-place bit-reversal permutation, optionally scaled (scale == nullptr: none)
__global__ void k_bitrev_scale(uint32_t* data, int L, const uint32_t* scale, size_t col_stride) {
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x, n = (size_t)1 << L;
    if (i >= n) return;
    uint32_t* col = data + (size_t)blockIdx.y * col_stride * 8;
    size_t j = bitrev(i, L);
    if (j < i) return;
    Fr a = fr_load(col + i * 8), b = fr_load(col + j * 8);
    if (scale) { Fr s = fr_load(scale); a = fr_mul(a, s); b = fr_mul(b, s); }
    fr_store(col + i * 8, b);
    if (j != i) fr_store(col + j * 8, a);
}
// g-power tables for the coset scaling: lo[i] = g^i / n (i < 1024), hi[j] = g^(1024 j) (j < n / 1024)
__global__ void k_gpow_tables(const uint32_t* shift, int log_n, uint32_t* lo, uint32_t* hi, size_t n_hi) {
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    Fr g = fr_load(shift);
    if (i < 1024) {
        Fr ninv = fr_inv(fr_from_u64(1ull << log_n));
        fr_store(lo + i * 8, fr_mul(fr_pow_u64(g, i), ninv));
    }
    if (i < n_hi) fr_store(hi + i * 8, fr_pow_u64(g, i << 10));
}
// coefficients (bit-reversed order, unscaled) of size n -> DIT input of size n << lb after its first lb stages:
// out[(i << lb) + r] = c_br[i] * g^bitrev(i) / n   for r < 2^lb
__global__ void k_lde_expand(const uint32_t* coef, int log_n, int lb, const uint32_t* gl, const uint32_t* gh, uint32_t* out,
                             size_t in_stride, size_t out_stride) {
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x, n = (size_t)1 << log_n;
    if (i >= n) return;
    const uint32_t* src = coef + (size_t)blockIdx.y * in_stride * 8;
    uint32_t* dst = out + (size_t)blockIdx.y * out_stride * 8;
    size_t k = bitrev(i, log_n);
    Fr s = fr_mul(fr_ldg(gl + (k & 1023) * 8), fr_ldg(gh + (k >> 10) * 8));
    Fr v = fr_mul(fr_load(src + i * 8), s);
    for (size_t r = 0; r < ((size_t)1 << lb); r++) fr_store(dst + ((i << lb) + r) * 8, v);
}
__global__ void k_copy_cols(const uint32_t* in, uint32_t* out, size_t n, size_t in_stride, size_t out_stride) {
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const uint4* s = reinterpret_cast<const uint4*>(in + ((size_t)blockIdx.y * in_stride + i) * 8);
    uint4* d = reinterpret_cast<uint4*>(out + ((size_t)blockIdx.y * out_stride + i) * 8);
    d[0] = s[0]; d[1] = s[1];
}

int ensure_twiddles(smw_ctx* c, int log_n) {
    if (log_n < 1) log_n = 1;
    if (c->d_twiddles && c->twiddle_log_n >= log_n) return SMW_OK;
    if (log_n > 32) return set_err(c, SMW_ERR_INVALID_ARG, "NTT size exceeds the field's two-adicity (2^32)");
    if (c->d_twiddles) { cudaStreamSynchronize(c->stream); cudaFree(c->d_twiddles); c->d_twiddles = nullptr; }
    size_t half = (size_t)1 << (log_n - 1);
    uint32_t *pow2, *ext;
    SMW_CUDA(c, cudaMalloc(&pow2, 33 * 32));
    SMW_CUDA(c, cudaMalloc(&ext, half * 32));
    k_pow2_roots<<<1, 32, 0, c->stream>>>(pow2);
    k_twiddles<<<(unsigned)((half + 255) / 256), 256, 0, c->stream>>>(ext, log_n, pow2);
    c->launches += 2;
    if (ntt_use29()) {        // the table the pass kernel reads: internal 29-bit form, 48-byte stride
        cudaError_t e = cudaMalloc(&c->d_twiddles, half * TW29 * 4);
        if (e != cudaSuccess) { cudaStreamSynchronize(c->stream); cudaFree(pow2); cudaFree(ext); return cuda_err(c, e, "cudaMalloc(twiddles)"); }
        k_twiddles_to29<<<(unsigned)((half + 255) / 256), 256, 0, c->stream>>>(ext, half, c->d_twiddles); c->launches++;
        SMW_CUDA(c, cudaStreamSynchronize(c->stream));
        cudaFree(ext);
    } else {
        SMW_CUDA(c, cudaStreamSynchronize(c->stream));
        c->d_twiddles = ext;
    }
    cudaFree(pow2);
    c->twiddle_log_n = log_n;
    return SMW_OK;
}
int launch_root_of_unity(smw_ctx* c, int log_n, uint32_t* d_out) {
    if (log_n < 0 || log_n > 32) return set_err(c, SMW_ERR_INVALID_ARG, "no root of unity of that order");
    uint32_t* pow2;
    SMW_CUDA(c, cudaMalloc(&pow2, 33 * 32));
    k_pow2_roots<<<1, 32, 0, c->stream>>>(pow2); c->launches++;
    SMW_CUDA(c, cudaMemcpyAsync(d_out, pow2 + log_n * 8, 32, cudaMemcpyDeviceToDevice, c->stream));
    SMW_CUDA(c, cudaStreamSynchronize(c->stream));
    cudaFree(pow2);
    return SMW_OK;
}

// run butterfly stages [s_from, L) (DIT, ascending) or all stages L-1..0 (DIF, descending) as tiled passes
static int run_passes(smw_ctx* c, uint32_t* d, int L, bool dif, int inverse, int s_from, int n_cols, size_t col_stride) {
    if (L == 0) return SMW_OK;
    int rc = ensure_twiddles(c, L); if (rc) return rc;
    // split bit positions [0, L) into passes: the lowest pass takes <= 10 bits with contiguous tiles, the others <= 8 bits + contiguous runs
    std::vector<std::pair<int, int>> ranges;   // [lo, hi)
    if (L <= TILE_LOG) ranges.push_back({0, L});
    else {
        int np = (L + 7) / 8, base = L / np, extra = L % np, lo = 0;
        for (int i = 0; i < np; i++) { int k = base + (i < extra ? 1 : 0); ranges.push_back({lo, lo + k}); lo += k; }
    }
    if (dif) std::reverse(ranges.begin(), ranges.end());
    for (auto& r : ranges) {
        PassArgs a; a.L = L; a.inverse = inverse; a.col_stride = col_stride; a.tw_shift = c->twiddle_log_n - L;
        a.s_begin = r.first; a.s_end = r.second;
        if (!dif) { if (a.s_end <= s_from) continue; if (a.s_begin < s_from) a.s_begin = s_from; }
        if (r.first == 0) { a.lo = 0; a.c = 0; a.span = L < TILE_LOG ? L : TILE_LOG; }
        else { a.lo = r.first; a.span = r.second - r.first; a.c = TILE_LOG - a.span; if (a.c > a.lo) a.c = a.lo; }
        a.tile_log = a.span + a.c;
        size_t tiles = ((size_t)1 << L) >> a.tile_log;
        static int cap = -1; if (cap < 0) { const char* e = getenv("SMW_NTT_THREADS"); cap = e ? atoi(e) : 128; if (cap < 32 || cap > 512) cap = 128; }   // threads per 1024-element tile: 128 (four butterflies per thread and stage; 512: 97 ms, 256: 82 ms, 128: 81 ms per cfg2 step, profiles/r02o_ntt_threads.txt — the stage barrier spans 4 warps instead of 16)
        int threads = (1 << a.tile_log) / 2; if (threads < 32) threads = 32; if (threads > cap) threads = cap;
        dim3 grid((unsigned)tiles, (unsigned)n_cols);
        LaunchScope ls(c, KC_NTT_PASS);
        if (ntt_use29()) {
            const size_t sm = (size_t)36 << a.tile_log;
            if (dif) k_ntt_pass29<true><<<grid, threads, sm, c->stream>>>(d, c->d_twiddles, a);
            else k_ntt_pass29<false><<<grid, threads, sm, c->stream>>>(d, c->d_twiddles, a);
        } else {
            const size_t sm = (size_t)32 << a.tile_log;
            if (dif) k_ntt_pass<true><<<grid, threads, sm, c->stream>>>(d, c->d_twiddles, a);
            else k_ntt_pass<false><<<grid, threads, sm, c->stream>>>(d, c->d_twiddles, a);
        }
    }
    SMW_CUDA(c, cudaGetLastError());
    return SMW_OK;
}

int launch_ntt(smw_ctx* c, uint32_t* d_buf, int log_n, int inverse, int n_cols, size_t col_stride) {
    if (log_n < 0 || log_n > 32) return set_err(c, SMW_ERR_INVALID_ARG, "bad NTT size");
    if (log_n == 0 || n_cols == 0) return SMW_OK;
    int rc = run_passes(c, d_buf, log_n, true, inverse, 0, n_cols, col_stride); if (rc) return rc;
    const size_t n = (size_t)1 << log_n;
    uint32_t *d_scale = nullptr;
    if (inverse) {   // 1/n: the g-power table kernel with g = pow2[0] = 1 gives lo[0] = 1/n; tables in a context scratch slot (no allocation, no sync per call)
        void* q; rc = ctx_scratch(c, smw_ctx::SLOT_NTT_SCALE, (size_t)(33 + 1025) * 32, &q); if (rc) return rc;
        uint32_t* pow2 = (uint32_t*)q; d_scale = pow2 + 33 * 8;
        k_pow2_roots<<<1, 32, 0, c->stream>>>(pow2);
        k_gpow_tables<<<4, 256, 0, c->stream>>>(pow2, log_n, d_scale, d_scale + 8 * 1024, 1);
        c->launches += 2;
    }
    k_bitrev_scale<<<dim3((unsigned)((n + 255) / 256), n_cols), 256, 0, c->stream>>>(d_buf, log_n, d_scale, col_stride);
    c->launches++;
    SMW_CUDA(c, cudaGetLastError());
    return SMW_OK;
}

int launch_lde(smw_ctx* c, const uint32_t* d_in, int log_n, int log_blowup, const uint32_t* d_shift, uint32_t* d_out, int n_cols) {
    if (log_n < 0 || log_blowup < 0 || log_n + log_blowup > 32) return set_err(c, SMW_ERR_INVALID_ARG, "bad LDE size");
    if (n_cols == 0) return SMW_OK;
    const size_t n = (size_t)1 << log_n, N = n << log_blowup;
    int rc = ensure_twiddles(c, log_n + log_blowup); if (rc) return rc;
    // 1. coefficients: inverse DIF in a scratch copy (bit-reversed, unscaled)
    uint32_t* d_coef; { void* q; int r0 = ctx_scratch(c, smw_ctx::SLOT_LDE_COEF, (size_t)n_cols * n * 32, &q); if (r0) return r0; d_coef = (uint32_t*)q; }
    { LaunchScope ls(c, KC_LDE_AUX); k_copy_cols<<<dim3((unsigned)((n + 255) / 256), n_cols), 256, 0, c->stream>>>(d_in, d_coef, n, n, n); }
    rc = run_passes(c, d_coef, log_n, true, 1, 0, n_cols, n); if (rc) return rc;
    // 2. coset scaling (g^k / n) + zero-pad-as-stride expansion
    size_t n_hi = (n + 1023) >> 10;
    uint32_t* d_g; { void* q; int r0 = ctx_scratch(c, smw_ctx::SLOT_LDE_G, (1024 + n_hi) * 32, &q); if (r0) return r0; d_g = (uint32_t*)q; }
    k_gpow_tables<<<(unsigned)((std::max<size_t>(1024, n_hi) + 255) / 256), 256, 0, c->stream>>>(d_shift, log_n, d_g, d_g + 8 * 1024, n_hi); c->launches++;
    { LaunchScope ls(c, KC_LDE_AUX); k_lde_expand<<<dim3((unsigned)((n + 255) / 256), n_cols), 256, 0, c->stream>>>(d_coef, log_n, log_blowup, d_g, d_g + 8 * 1024, d_out, n, N); }
    // 3. forward DIT, skipping the log_blowup stages already folded into the expansion
    rc = run_passes(c, d_out, log_n + log_blowup, false, 0, log_blowup, n_cols, N);
    return rc;
}

} // namespace smw
