// field_kernels.cu — element-wise and fold kernels in the external representation (fr32.cuh):
// FRI fold (deep_ali/src/fri.rs:85-102), s-layer broadcast (:123-143), the Merkle-folded sum-check's
// round coefficients and fold (channel/src/lib.rs:406-416, 648-655), the DEEP-ALI merge
// (deep_ali/src/lib.rs:60-105) and the canonical <-> Montgomery conversions.  All HBM-bound.
#include "smw_internal.h"
#include "fr32.cuh"
#include "chacha.cuh"
#include "fr52.cuh"
#include <algorithm>
#include <vector>

namespace smw {

// out[b] = sum_{t<m} f[b*m + t] * z^t.  One thread per output; z powers are rebuilt per thread by a
// running product (m <= 128 multiplies, against the m loads it needs anyway).
__global__ void k_fri_fold(const uint32_t* __restrict__ f, size_t n_out, const uint32_t* __restrict__ z, size_t m, uint32_t* __restrict__ out) {
    size_t b = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (b >= n_out) return;
    Fr zz = fr_ldg(z), zp = fr_const(C_R32), acc = fr_zero();
    const uint32_t* src = f + b * m * 8;
    for (size_t t = 0; t < m; t++) {
        Fr x = fr_load(src + t * 8);
        acc = fr_add(acc, t == 0 ? x : fr_mul(x, zp));
        zp = (t == 0) ? zz : fr_mul(zp, zz);
    }
    fr_store(out + b * 8, acc);
}
__global__ void k_broadcast_div(const uint32_t* __restrict__ src, size_t n, size_t m, uint32_t* __restrict__ out) {
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    fr_store(out + i * 8, fr_load(src + (i / m) * 8));
}
__global__ void k_to_mont(const uint32_t* __restrict__ in, uint32_t* __restrict__ out, size_t n, int inverse) {
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    Fr x = fr_load(in + i * 8);
    fr_store(out + i * 8, inverse ? fr_to_canonical(x) : fr_from_canonical(x));
}
// c0 = sum a, c1 = sum (b - a) over pairs (a, b) = (layer[2j], layer[2j+1]); block tree reduction, then a
// second launch over the per-block partials (exact field sums, so the order does not matter)
__global__ void __launch_bounds__(256) k_mf_coeffs(const uint32_t* __restrict__ layer, size_t n_pairs, uint32_t* __restrict__ partial, int pairs_are_c0c1) {
    __shared__ uint32_t sh[2][256][8];
    Fr c0 = fr_zero(), c1 = fr_zero();
    for (size_t j = (size_t)blockIdx.x * blockDim.x + threadIdx.x; j < n_pairs; j += (size_t)gridDim.x * blockDim.x) {
        Fr a = fr_load(layer + (2 * j) * 8), b = fr_load(layer + (2 * j + 1) * 8);
        if (pairs_are_c0c1) { c0 = fr_add(c0, a); c1 = fr_add(c1, b); }
        else { c0 = fr_add(c0, a); c1 = fr_add(c1, fr_sub(b, a)); }
    }
    for (int k = 0; k < 8; k++) { sh[0][threadIdx.x][k] = c0.v[k]; sh[1][threadIdx.x][k] = c1.v[k]; }
    __syncthreads();
    for (int stride = 128; stride > 0; stride >>= 1) {
        if ((int)threadIdx.x < stride) {
            for (int w = 0; w < 2; w++) {
                Fr x, y;
                for (int k = 0; k < 8; k++) { x.v[k] = sh[w][threadIdx.x][k]; y.v[k] = sh[w][threadIdx.x + stride][k]; }
                x = fr_add(x, y);
                for (int k = 0; k < 8; k++) sh[w][threadIdx.x][k] = x.v[k];
            }
        }
        __syncthreads();
    }
    if (threadIdx.x == 0) for (int w = 0; w < 2; w++) for (int k = 0; k < 8; k++) partial[((size_t)blockIdx.x * 2 + w) * 8 + k] = sh[w][0][k];
}
// next[j] = (1 - r) * cur[2j] + r * cur[2j+1]
__global__ void k_mf_fold(const uint32_t* __restrict__ cur, size_t n_out, const uint32_t* __restrict__ r, uint32_t* __restrict__ next) {
    size_t j = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= n_out) return;
    Fr rr = fr_ldg(r), om = fr_sub(fr_const(C_R32), rr);
    Fr a = fr_load(cur + (2 * j) * 8), b = fr_load(cur + (2 * j + 1) * 8);
    fr_store(next + j * 8, fr_add(fr_mul(om, a), fr_mul(rr, b)));
}
// DEEP-ALI merge, f0 only: f0[j] = (a*s + e - t)[j] / (w^j - z).  Batch inversion per thread over a strip of
// CH consecutive j (Montgomery's trick: one field inversion per CH elements instead of one per element).
template <int CH> __global__ void k_deep_ali_merge(const uint32_t* __restrict__ a, const uint32_t* __restrict__ s, const uint32_t* __restrict__ e, const uint32_t* __restrict__ t,
                                                   size_t n, const uint32_t* __restrict__ omega, const uint32_t* __restrict__ z, uint32_t* __restrict__ f0) {
    size_t strip = (size_t)blockIdx.x * blockDim.x + threadIdx.x, j0 = strip * CH;
    if (j0 >= n) return;
    Fr w = fr_ldg(omega), zz = fr_ldg(z);
    Fr wj = fr_pow_u64(w, j0);
    Fr d[CH], pre[CH];
    Fr run = fr_const(C_R32);
    int cnt = 0;
    for (int k = 0; k < CH && j0 + k < n; k++, cnt++) { d[k] = fr_sub(wj, zz); pre[k] = run; run = fr_mul(run, d[k]); wj = fr_mul(wj, w); }
    Fr inv = fr_inv(run);
    for (int k = cnt - 1; k >= 0; k--) {
        Fr dinv = fr_mul(inv, pre[k]); inv = fr_mul(inv, d[k]);
        size_t j = j0 + k;
        Fr phi = fr_sub(fr_add(fr_mul(fr_load(a + j * 8), fr_load(s + j * 8)), fr_load(e + j * 8)), fr_load(t + j * 8));
        fr_store(f0 + j * 8, fr_mul(phi, dinv));
    }
}

__global__ void k_gather(const uint32_t* __restrict__ src, const uint64_t* __restrict__ idx, size_t n, uint32_t* __restrict__ out) {
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    fr_store(out + i * 8, fr_load(src + idx[i] * 8));
}
// data[i] *= g^i: two-level power tables (lo[i] = g^i, i < 1024; hi[j] = g^(1024 j)) built per call, then one pass.
// The twiddle step of the cross-GPU four-step NTT (g = w_N^rank) and any coset scaling.
__global__ void k_pow_tables(const uint32_t* __restrict__ g, uint32_t* __restrict__ lo, uint32_t* __restrict__ hi, size_t n_hi) {
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    Fr gg = fr_ldg(g);
    if (i < 1024) fr_store(lo + i * 8, fr_pow_u64(gg, i));
    if (i < n_hi) fr_store(hi + i * 8, fr_pow_u64(gg, i << 10));
}
__global__ void k_scale_powers(uint32_t* __restrict__ data, size_t n, const uint32_t* __restrict__ lo, const uint32_t* __restrict__ hi) {
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    Fr s = fr_mul(fr_ldg(lo + (i & 1023) * 8), fr_ldg(hi + (i >> 10) * 8));
    fr_store(data + i * 8, fr_mul(fr_load(data + i * 8), s));
}
// in-place DFT of length G = 2^LG across `stride`-spaced elements, one thread per offset j < count:
// (data[a * stride + j])_a  ->  (sum_a data[a * stride + j] * w_G^(a k))_k.  G <= 16: the cross-rank step of the four-step NTT.
template <int LG> __global__ void k_dft_strided(uint32_t* __restrict__ data, size_t stride, size_t count, const uint32_t* __restrict__ wpow /* w_G^i, i < G */) {
    constexpr int G = 1 << LG;
    size_t j = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= count) return;
    Fr x[G];
#pragma unroll
    for (int a = 0; a < G; a++) x[a] = fr_load(data + ((size_t)a * stride + j) * 8);
#pragma unroll 1
    for (int k = 0; k < G; k++) {
        Fr acc = x[0];
#pragma unroll
        for (int a = 1; a < G; a++) {
            int e = (a * k) & (G - 1);
            acc = fr_add(acc, e == 0 ? x[a] : fr_mul(x[a], fr_ldg(wpow + e * 8)));
        }
        fr_store(data + ((size_t)k * stride + j) * 8, acc);      // safe: every input of this thread is already in registers
    }
}
int launch_scale_powers(smw_ctx* c, uint32_t* d_data, size_t n, const uint32_t* d_g) {
    if (n == 0) return SMW_OK;
    size_t n_hi = (n + 1023) >> 10;
    uint32_t* d_t; { void* q; int rc = ctx_scratch(c, smw_ctx::SLOT_LDE_G, (1024 + n_hi) * 32, &q); if (rc) return rc; d_t = (uint32_t*)q; }
    k_pow_tables<<<(unsigned)((std::max<size_t>(1024, n_hi) + 255) / 256), 256, 0, c->stream>>>(d_g, d_t, d_t + 8 * 1024, n_hi); c->launches++;
    { LaunchScope ls(c, KC_LDE_AUX); k_scale_powers<<<(unsigned)((n + 255) / 256), 256, 0, c->stream>>>(d_data, n, d_t, d_t + 8 * 1024); }
    SMW_CUDA(c, cudaGetLastError());
    return SMW_OK;
}
__global__ void k_fr_pow(const uint32_t* base, unsigned long long e, uint32_t* out) {
    if (threadIdx.x != 0 || blockIdx.x != 0) return;
    fr_store(out, fr_pow_u64(fr_ldg(base), e));
}
int launch_fr_pow(smw_ctx* c, const uint32_t* d_base, uint64_t e, uint32_t* d_out) {
    k_fr_pow<<<1, 32, 0, c->stream>>>(d_base, (unsigned long long)e, d_out); c->launches++;
    SMW_CUDA(c, cudaGetLastError());
    return SMW_OK;
}
int launch_dft_strided(smw_ctx* c, uint32_t* d_data, int log_g, size_t stride, size_t count, const uint32_t* d_wpow) {
    if (count == 0 || log_g == 0) return SMW_OK;
    unsigned grid = (unsigned)((count + 127) / 128);
    LaunchScope ls(c, KC_NTT_PASS);
    switch (log_g) {
        case 1: k_dft_strided<1><<<grid, 128, 0, c->stream>>>(d_data, stride, count, d_wpow); break;
        case 2: k_dft_strided<2><<<grid, 128, 0, c->stream>>>(d_data, stride, count, d_wpow); break;
        case 3: k_dft_strided<3><<<grid, 128, 0, c->stream>>>(d_data, stride, count, d_wpow); break;
        case 4: k_dft_strided<4><<<grid, 128, 0, c->stream>>>(d_data, stride, count, d_wpow); break;
        default: return set_err(c, SMW_ERR_INVALID_ARG, "strided DFT length must be 2..16");
    }
    SMW_CUDA(c, cudaGetLastError());
    return SMW_OK;
}
// index_from_seed (deep_ali/src/fri.rs:178-183): StdRng::from_seed(canonical LE bytes of the seed).next_u64() & (n_pow2 - 1)
__global__ void k_index_from_seed(const uint32_t* __restrict__ seeds, size_t n, uint64_t mask, uint64_t* __restrict__ out) {
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    Fr c = fr_to_canonical(fr_load(seeds + i * 8));
    ChaChaRng g; chacha_init(g, c.v);
    out[i] = chacha_next_u64(g) & mask;
}
// ali_sample_z_beta_fs after its hash (deep_ali/src/fri.rs:515-533): beta = first draw; z = first later draw c with c != 0 and
// c^n0 != 1 (at most 1000 tries), else roots_seed + 17 if that is not an n0-th root of unity, else 19
__global__ void k_ali_sample(const uint32_t* fused, const uint32_t* roots_seed, uint64_t n0, uint32_t* z_out, uint32_t* beta_out) {
    if (threadIdx.x != 0 || blockIdx.x != 0) return;
    Fr c = fr_to_canonical(fr_load(fused));
    ChaChaRng g; chacha_init(g, c.v);
    fr_store(beta_out, fr_from_u64(chacha_next_u64(g)));
    const Fr one = fr_const(C_R32);
    for (int tries = 0; tries < 1000; tries++) {
        uint64_t v = chacha_next_u64(g);
        if (v == 0) continue;
        Fr cand = fr_from_u64(v);
        if (!fr_eq(fr_pow_u64(cand, n0), one)) { fr_store(z_out, cand); return; }
    }
    Fr fb = fr_add(fr_load(roots_seed), fr_from_u64(17));
    fr_store(z_out, fr_eq(fr_pow_u64(fb, n0), one) ? fr_from_u64(19) : fb);
}
int launch_index_from_seed(smw_ctx* c, const uint32_t* d_seeds, size_t n, uint64_t n_pow2, uint64_t* d_out) {
    if (n == 0) return SMW_OK;
    k_index_from_seed<<<(unsigned)((n + 63) / 64), 64, 0, c->stream>>>(d_seeds, n, n_pow2 - 1, d_out); c->launches++;
    SMW_CUDA(c, cudaGetLastError());
    return SMW_OK;
}
int launch_ali_sample(smw_ctx* c, const uint32_t* d_fused, const uint32_t* d_roots_seed, uint64_t n0, uint32_t* d_z, uint32_t* d_beta) {
    k_ali_sample<<<1, 32, 0, c->stream>>>(d_fused, d_roots_seed, n0, d_z, d_beta); c->launches++;
    SMW_CUDA(c, cudaGetLastError());
    return SMW_OK;
}
// self-test of the FP64-pipe arithmetic (fr52.cuh): out[i] = sum_{j < k} a[i k + j] * b[i k + j], external form in and out
__global__ void k_f52_dot(const uint32_t* __restrict__ a, const uint32_t* __restrict__ b, size_t n_out, int k, uint32_t* __restrict__ out) {
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_out) return;
    const F52 kin = f52_const(C_KIN52), kout = f52_const(C_KOUT52);
    Acc52 A; acc52_zero(A);
    for (int j = 0; j < k; j++) {
        Fr x = fr_load(a + (i * k + j) * 8), y = fr_load(b + (i * k + j) * 8);
        F52 X = f52_mul(f52_repack_from32(x.v), kin), Y = f52_mul(f52_repack_from32(y.v), kin);
        uint64_t yl[5]; f52_canonical_limbs(Y, yl);                 // table operands are canonical (< p): keep the bound of the production path
        acc52_mac(A, X, f52_from_limbs(yl));
    }
    F52 r = acc52_reduce_add2(A, (uint32_t)k, f52_zero(), f52_zero());
    uint64_t l[5]; f52_canonical_limbs(f52_mul(r, kout), l);
    Fr o; f52_limbs_to32(l, o.v); fr_store(out + i * 8, o);
}
int launch_f52_dot(smw_ctx* c, const uint32_t* d_a, const uint32_t* d_b, size_t n_out, int k, uint32_t* d_out) {
    if (n_out == 0) return SMW_OK;
    k_f52_dot<<<(unsigned)((n_out + 63) / 64), 64, 0, c->stream>>>(d_a, d_b, n_out, k, d_out); c->launches++;
    SMW_CUDA(c, cudaGetLastError());
    return SMW_OK;
}
int launch_gather(smw_ctx* c, const uint32_t* d_src, const uint64_t* d_idx, size_t n, uint32_t* d_out) {
    if (n == 0) return SMW_OK;
    k_gather<<<(unsigned)((n + 127) / 128), 128, 0, c->stream>>>(d_src, d_idx, n, d_out); c->launches++;
    SMW_CUDA(c, cudaGetLastError()); return SMW_OK;
}
int launch_fri_fold(smw_ctx* c, const uint32_t* d_f, size_t n, const uint32_t* d_z, size_t m, uint32_t* d_out) {
    size_t n_out = n / m; if (n_out == 0) return SMW_OK;
    { LaunchScope ls(c, KC_FOLD); k_fri_fold<<<(unsigned)((n_out + 127) / 128), 128, 0, c->stream>>>(d_f, n_out, d_z, m, d_out); }
    SMW_CUDA(c, cudaGetLastError()); return SMW_OK;
}
int launch_broadcast_div(smw_ctx* c, const uint32_t* d_src, size_t n, size_t m, uint32_t* d_out) {
    if (n == 0) return SMW_OK;
    k_broadcast_div<<<(unsigned)((n + 255) / 256), 256, 0, c->stream>>>(d_src, n, m, d_out); c->launches++;
    SMW_CUDA(c, cudaGetLastError()); return SMW_OK;
}
int launch_to_mont(smw_ctx* c, const uint32_t* d_in, uint32_t* d_out, size_t n, int inverse) {
    if (n == 0) return SMW_OK;
    k_to_mont<<<(unsigned)((n + 255) / 256), 256, 0, c->stream>>>(d_in, d_out, n, inverse); c->launches++;
    SMW_CUDA(c, cudaGetLastError()); return SMW_OK;
}
int launch_mf_coeffs(smw_ctx* c, const uint32_t* d_layer, size_t n, uint32_t* d_c0c1) {
    size_t pairs = n / 2;
    unsigned blocks = (unsigned)std::min<size_t>((pairs + 255) / 256, 1024); if (blocks == 0) blocks = 1;
    uint32_t* d_part; SMW_CUDA(c, cudaMalloc(&d_part, (size_t)blocks * 2 * 32));
    k_mf_coeffs<<<blocks, 256, 0, c->stream>>>(d_layer, pairs, d_part, 0);
    k_mf_coeffs<<<1, 256, 0, c->stream>>>(d_part, blocks, d_c0c1, 1);
    c->launches += 2;
    SMW_CUDA(c, cudaStreamSynchronize(c->stream));
    cudaFree(d_part);
    return SMW_OK;
}
int launch_mf_fold(smw_ctx* c, const uint32_t* d_cur, size_t n, const uint32_t* d_r, uint32_t* d_next) {
    size_t n_out = n / 2; if (n_out == 0) return SMW_OK;
    { LaunchScope ls(c, KC_FOLD); k_mf_fold<<<(unsigned)((n_out + 255) / 256), 256, 0, c->stream>>>(d_cur, n_out, d_r, d_next); }
    SMW_CUDA(c, cudaGetLastError()); return SMW_OK;
}
// deep_ali/src/lib.rs:77,97 asserts !is_in_domain(z, n): z^n == 1 would make a denominator w^j - z vanish
__global__ void k_z_in_domain(const uint32_t* z, unsigned long long n, int* flag) {
    if (threadIdx.x != 0 || blockIdx.x != 0) return;
    const Fr zn = fr_pow_u64(fr_ldg(z), n), one = fr_const(C_R32);
    bool eq = true;
    for (int i = 0; i < 8; i++) eq = eq && zn.v[i] == one.v[i];
    *flag = eq ? 1 : 0;
}
int launch_deep_ali_merge(smw_ctx* c, const uint32_t* a, const uint32_t* s, const uint32_t* e, const uint32_t* t, size_t n, const uint32_t* d_omega, const uint32_t* d_z, uint32_t* d_f0) {
    constexpr int CH = 8;
    size_t strips = (n + CH - 1) / CH; if (strips == 0) return SMW_OK;
    {
        int* d_flag; SMW_CUDA(c, cudaMalloc(&d_flag, 4));
        k_z_in_domain<<<1, 32, 0, c->stream>>>(d_z, (unsigned long long)n, d_flag); c->launches++;
        int flag = 0; cudaError_t er = cudaMemcpyAsync(&flag, d_flag, 4, cudaMemcpyDeviceToHost, c->stream);
        if (er == cudaSuccess) er = cudaStreamSynchronize(c->stream);
        cudaFree(d_flag);
        if (er != cudaSuccess) return cuda_err(c, er, "k_z_in_domain");
        if (flag) return set_err(c, SMW_ERR_Z_IN_DOMAIN, "z must be outside H");
    }
    k_deep_ali_merge<CH><<<(unsigned)((strips + 127) / 128), 128, 0, c->stream>>>(a, s, e, t, n, d_omega, d_z, d_f0); c->launches++;
    SMW_CUDA(c, cudaGetLastError()); return SMW_OK;
}


// ---- integer-pipe micro-benchmark: the issue rate of the accumulating IMAD.WIDE.U32 (32x32+64 -> 64), 16 independent
// chains per thread, 8 x 256 threads per SM.  Its result is the denominator of the "fraction of the integer-multiply
// roofline" figures.  Four of the operands change every iteration (a LOP3 each, on the other pipe) so that ptxas can
// neither hoist the products out of the loop nor split the instruction into IMAD.WIDE (RZ) + IADD3 pairs — round 1a's
// version had loop-invariant operands and measured 64-bit ADDS instead (tools/imad_peak2.cu is the stand-alone, wider
// version of this measurement; tools/check_imad_sass.py prints the SASS mix of each timed loop).
__global__ void __launch_bounds__(256) k_diag_mac(uint32_t* out, const uint32_t* in, long long* cyc, int iters) {
    uint64_t w[16]; uint32_t a[4], b[8];
#pragma unroll
    for (int i = 0; i < 16; i++) w[i] = reinterpret_cast<const uint64_t*>(in)[(threadIdx.x + 31 * i) & 511];
#pragma unroll
    for (int i = 0; i < 4; i++) a[i] = in[(threadIdx.x + 40 + i) & 1023] & 0x1fffffffu;
#pragma unroll
    for (int i = 0; i < 8; i++) b[i] = in[(threadIdx.x + 80 + i) & 1023] & 0x1fffffffu;
    long long t0 = clock64();
#pragma unroll 1
    for (int it = 0; it < iters; it++) {
#pragma unroll
        for (int i = 0; i < 16; i++) w[i] += (uint64_t)a[i & 3] * b[i >> 2 | (i & 3) << 1];
#pragma unroll
        for (int i = 0; i < 4; i++) a[i] ^= it;
    }
    long long t1 = clock64();
    uint32_t r = 0;
#pragma unroll
    for (int i = 0; i < 16; i++) r ^= (uint32_t)w[i] ^ (uint32_t)(w[i] >> 32);
    out[(size_t)blockIdx.x * blockDim.x + threadIdx.x] = r;
    if (threadIdx.x == 0) cyc[blockIdx.x] = t1 - t0;
}
// the same measurement for the FP64 pipe: 16 independent DFMA.RZ chains per thread (what fr52.cuh issues)
__global__ void __launch_bounds__(256) k_diag_dfma(uint32_t* out, const uint32_t* in, long long* cyc, int iters) {
    double d[16];
#pragma unroll
    for (int i = 0; i < 16; i++) d[i] = 1.0 + 1e-9 * (double)(in[(threadIdx.x + 31 * i) & 1023] & 0xffff);
    const double x = 1.0 - 1e-9 * (double)(in[threadIdx.x & 1023] & 0xff), y = 1e-7;
    long long t0 = clock64();
#pragma unroll 1
    for (int it = 0; it < iters; it++) {
#pragma unroll
        for (int i = 0; i < 16; i++) d[i] = __fma_rz(d[i], x, y);
    }
    long long t1 = clock64();
    uint32_t r = 0;
#pragma unroll
    for (int i = 0; i < 16; i++) r ^= (uint32_t)__double_as_longlong(d[i]) ^ (uint32_t)(__double_as_longlong(d[i]) >> 32);
    out[(size_t)blockIdx.x * blockDim.x + threadIdx.x] = r;
    if (threadIdx.x == 0) cyc[blockIdx.x] = t1 - t0;
}
int launch_diag_mac_rate(smw_ctx* c, double* mac_per_s, double* winstr) { return launch_diag_rate(c, 0, mac_per_s, winstr); }
int launch_diag_rate(smw_ctx* c, int fp64, double* mac_per_s, double* winstr) { return launch_diag_rate_ex(c, fp64, mac_per_s, winstr, nullptr, nullptr, nullptr); }
// also reports the SM clock the GPU held DURING the measurement (in-kernel clock64 ticks / event time), the SM count and the
// nominal maximum clock, so that lanes_per_s = warp_instr_per_clk_per_sm x 32 x n_sm x sm_hz_during can be re-derived by the reader
int launch_diag_rate_ex(smw_ctx* c, int fp64, double* mac_per_s, double* winstr, double* sm_hz_during, double* n_sm_out, double* sm_max_hz) {
    cudaDeviceProp prop; SMW_CUDA(c, cudaGetDeviceProperties(&prop, c->device));
    // 2 CTAs x 256 threads per SM (all resident, so the in-kernel clock and the event clock agree); ~40 ms per repetition
    // so that the clocks settle under this load
    const int nsm = prop.multiProcessorCount, per_sm = 2, block = 256, iters = 1 << 19, grid = nsm * per_sm;
    uint32_t *out, *in; long long* cyc;
    SMW_CUDA(c, cudaMalloc(&out, (size_t)grid * block * 4)); SMW_CUDA(c, cudaMalloc(&cyc, (size_t)grid * 8)); SMW_CUDA(c, cudaMalloc(&in, 4096 * 4));
    std::vector<uint32_t> hin(4096); for (int i = 0; i < 4096; i++) hin[i] = 0x9E3779B9u * (uint32_t)(i + 1) ^ ((uint32_t)i << 7);
    SMW_CUDA(c, cudaMemcpyAsync(in, hin.data(), 4096 * 4, cudaMemcpyHostToDevice, c->stream));
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    double best = 0, best_w = 0, best_hz = 0;
    std::vector<long long> h(grid);
    for (int rep = 0; rep < 4; rep++) {
        cudaEventRecord(e0, c->stream);
        if (fp64) k_diag_dfma<<<grid, block, 0, c->stream>>>(out, in, cyc, iters); else k_diag_mac<<<grid, block, 0, c->stream>>>(out, in, cyc, iters);
        c->launches++;
        cudaEventRecord(e1, c->stream);
        SMW_CUDA(c, cudaStreamSynchronize(c->stream));
        float ms = 0; cudaEventElapsedTime(&ms, e0, e1);
        cudaMemcpy(h.data(), cyc, (size_t)grid * 8, cudaMemcpyDeviceToHost);
        double avg = 0; for (int i = 0; i < grid; i++) avg += (double)h[i]; avg /= grid;
        double rate = (double)grid * block * (double)iters * 16 / (ms * 1e-3);
        if (rep > 0 && rate > best) { best = rate; best_w = per_sm * (block / 32.0) * (double)iters * 16 / avg; best_hz = avg / (ms * 1e-3); }
    }
    cudaEventDestroy(e0); cudaEventDestroy(e1); cudaFree(out); cudaFree(cyc); cudaFree(in);
    if (mac_per_s) *mac_per_s = best;
    if (winstr) *winstr = best_w;
    if (sm_hz_during) *sm_hz_during = best_hz;
    if (n_sm_out) *n_sm_out = (double)nsm;
    if (sm_max_hz) *sm_max_hz = (double)prop.clockRate * 1e3;
    return SMW_OK;
}

// ---------------------------------------------------------------------------------------- allocation-free round coefficients
// smw_sumcheck_round_coeffs for callers that may not allocate (CUDA graph capture, tight serving loops, harnesses that forbid
// cudaMalloc).  No counterpart in the reference.  Same two launches as
// launch_mf_coeffs above, but the per-block partial sums live in a file-scope device array instead of a cudaMalloc'ed buffer:
//   * SMW_NOALLOC_MAX_BLOCKS x 2 x 32 B = 64 KiB, held by every GPU that loads the library until the process exits;
//   * every call on a device shares it: calls that overlap on the GPU (different streams / contexts) race;
//   * nothing in it has to start at zero — the first launch overwrites exactly the `blocks` entries the second one reads, so
//     what an earlier (possibly larger) call left behind is never read.
// In host code `g_noalloc_partials` is only a stand-in for the symbol, not a device address: the address comes from
// cudaGetSymbolAddress, once per device, with that device current.
constexpr unsigned SMW_NOALLOC_MAX_BLOCKS = 1024;
constexpr size_t SMW_NOALLOC_MAX_N = (size_t)1 << 26;          // elements per call (2 GiB of input): BASELINE's largest layer
static __device__ __align__(256) uint32_t g_noalloc_partials[SMW_NOALLOC_MAX_BLOCKS * 2 * 8];

static int noalloc_partials_ptr(smw_ctx* c, uint32_t** out) {
    constexpr int MAX_DEV = 64;
    static std::mutex mu;
    static uint32_t* ptr[MAX_DEV] = {};
    if (c->device < 0 || c->device >= MAX_DEV) return set_err(c, SMW_ERR_INVALID_ARG, "smw_sumcheck_round_coeffs_noalloc: device index beyond 63");
    std::lock_guard<std::mutex> lk(mu);
    if (!ptr[c->device]) {
        void* p = nullptr;
        SMW_CUDA(c, cudaGetSymbolAddress(&p, g_noalloc_partials));          // the device selected by the caller (cudaSetDevice) is current
        ptr[c->device] = (uint32_t*)p;
    }
    *out = ptr[c->device];
    return SMW_OK;
}

} // namespace smw

int32_t smw_sumcheck_round_coeffs_noalloc(smw_ctx* c, const void* layer_dev, size_t n, void* c0c1_out_dev) {
    using namespace smw;
    if (!c) return SMW_ERR_INVALID_ARG;
    if (!layer_dev || !c0c1_out_dev || n == 0 || (n & 1)) return set_err(c, SMW_ERR_INVALID_ARG, "smw_sumcheck_round_coeffs_noalloc: device pointers and an even, non-zero length are required");
    if (n > SMW_NOALLOC_MAX_N) return set_err(c, SMW_ERR_INVALID_ARG, "smw_sumcheck_round_coeffs_noalloc: n exceeds the fixed scratch's limit of 2^26 elements (use smw_sumcheck_round_coeffs)");
    SMW_CUDA(c, cudaSetDevice(c->device));
    uint32_t* d_part = nullptr;
    int rc = noalloc_partials_ptr(c, &d_part); if (rc) return rc;
    const size_t pairs = n / 2;
    unsigned blocks = (unsigned)std::min<size_t>((pairs + 255) / 256, SMW_NOALLOC_MAX_BLOCKS); if (blocks == 0) blocks = 1;
    k_mf_coeffs<<<blocks, 256, 0, c->stream>>>((const uint32_t*)layer_dev, pairs, d_part, 0);
    k_mf_coeffs<<<1, 256, 0, c->stream>>>(d_part, blocks, (uint32_t*)c0c1_out_dev, 1);
    c->launches += 2;
    SMW_CUDA(c, cudaGetLastError());
    return SMW_OK;
}
