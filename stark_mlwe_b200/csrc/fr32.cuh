// fr32.cuh — Pallas Fr in the external representation (8 x u32, Montgomery R = 2^256, canonical),
// for the HBM-bound kernels (NTT butterflies, folds, element-wise maps) where a value is touched by
// one or two multiplications between a load and a store, so staying in the caller's representation
// beats converting to the 29-bit form of fr29.cuh.  Same memory image as ark-ff Fp256 (field/src/lib.rs:13).
#pragma once
#include <cstdint>

namespace smw {

#ifndef SMW_DEV
#define SMW_DEV __device__ __forceinline__
#endif

struct Fr { uint32_t v[8]; };

constexpr uint32_t P32_0 = 0x00000001u, P32_1 = 0x8c46eb21u, P32_2 = 0x0994a8ddu, P32_3 = 0x224698fcu, P32_7 = 0x40000000u;
__device__ __constant__ const uint32_t C_P32[8]  = {0x00000001u, 0x8c46eb21u, 0x0994a8ddu, 0x224698fcu, 0u, 0u, 0u, 0x40000000u};
__device__ __constant__ const uint32_t C_R32[8]  = {0xfffffffdu, 0x5b2b3e9cu, 0xe3420567u, 0x992c350bu, 0xffffffffu, 0xffffffffu, 0xffffffffu, 0x3fffffffu}; // 2^256 mod p (Montgomery one)
__device__ __constant__ const uint32_t C_RR32[8] = {0x0000000fu, 0xfc9678ffu, 0x891a16e3u, 0x67bb433du, 0x04ccf590u, 0x7fae2310u, 0x7ccfdaa9u, 0x096d41afu}; // 2^512 mod p

SMW_DEV Fr fr_load(const uint32_t* p) { const uint4* q = reinterpret_cast<const uint4*>(p); uint4 a = q[0], b = q[1]; Fr r; r.v[0] = a.x; r.v[1] = a.y; r.v[2] = a.z; r.v[3] = a.w; r.v[4] = b.x; r.v[5] = b.y; r.v[6] = b.z; r.v[7] = b.w; return r; }
SMW_DEV Fr fr_ldg(const uint32_t* p) { const uint4* q = reinterpret_cast<const uint4*>(p); uint4 a = __ldg(q), b = __ldg(q + 1); Fr r; r.v[0] = a.x; r.v[1] = a.y; r.v[2] = a.z; r.v[3] = a.w; r.v[4] = b.x; r.v[5] = b.y; r.v[6] = b.z; r.v[7] = b.w; return r; }
SMW_DEV void fr_store(uint32_t* p, const Fr& x) { uint4* q = reinterpret_cast<uint4*>(p); q[0] = make_uint4(x.v[0], x.v[1], x.v[2], x.v[3]); q[1] = make_uint4(x.v[4], x.v[5], x.v[6], x.v[7]); }
SMW_DEV Fr fr_const(const uint32_t* c) { Fr r;
#pragma unroll
    for (int i = 0; i < 8; i++) r.v[i] = c[i];
    return r; }
SMW_DEV Fr fr_zero() { Fr r;
#pragma unroll
    for (int i = 0; i < 8; i++) r.v[i] = 0;
    return r; }
SMW_DEV bool fr_is_zero(const Fr& a) { uint32_t o = 0;
#pragma unroll
    for (int i = 0; i < 8; i++) o |= a.v[i];
    return o == 0; }
SMW_DEV bool fr_eq(const Fr& a, const Fr& b) { uint32_t o = 0;
#pragma unroll
    for (int i = 0; i < 8; i++) o |= a.v[i] ^ b.v[i];
    return o == 0; }

// r = a - p if a >= p (a < 2p)
SMW_DEV Fr fr_csub_p(const Fr& a) {
    Fr d; uint32_t borrow = 0;
#pragma unroll
    for (int i = 0; i < 8; i++) {
        uint64_t t = (uint64_t)a.v[i] - C_P32[i] - borrow;
        d.v[i] = (uint32_t)t; borrow = (uint32_t)(t >> 63);
    }
    Fr r;
#pragma unroll
    for (int i = 0; i < 8; i++) r.v[i] = borrow ? a.v[i] : d.v[i];
    return r;
}
SMW_DEV Fr fr_add(const Fr& a, const Fr& b) {
    Fr s; uint32_t c = 0;
#pragma unroll
    for (int i = 0; i < 8; i++) { uint64_t t = (uint64_t)a.v[i] + b.v[i] + c; s.v[i] = (uint32_t)t; c = (uint32_t)(t >> 32); }
    return fr_csub_p(s);   // p < 2^255: no carry out of 256 bits
}
SMW_DEV Fr fr_sub(const Fr& a, const Fr& b) {
    Fr d; uint32_t borrow = 0;
#pragma unroll
    for (int i = 0; i < 8; i++) { uint64_t t = (uint64_t)a.v[i] - b.v[i] - borrow; d.v[i] = (uint32_t)t; borrow = (uint32_t)(t >> 63); }
    uint32_t mask = 0u - borrow, c = 0;
#pragma unroll
    for (int i = 0; i < 8; i++) { uint64_t t = (uint64_t)d.v[i] + (C_P32[i] & mask) + c; d.v[i] = (uint32_t)t; c = (uint32_t)(t >> 32); }
    return d;
}
SMW_DEV Fr fr_neg(const Fr& a) { return fr_sub(fr_zero(), a); }
// Montgomery product (CIOS, 32-bit words; -p^-1 mod 2^32 = 0xffffffff, p limbs 4..6 are zero)
SMW_DEV Fr fr_mul(const Fr& a, const Fr& b) {
    uint32_t t[10];
#pragma unroll
    for (int i = 0; i < 10; i++) t[i] = 0;
#pragma unroll
    for (int i = 0; i < 8; i++) {
        uint64_t c = 0;
#pragma unroll
        for (int j = 0; j < 8; j++) { c += (uint64_t)a.v[j] * b.v[i] + t[j]; t[j] = (uint32_t)c; c >>= 32; }
        c += t[8]; t[8] = (uint32_t)c; t[9] = (uint32_t)(c >> 32);
        uint32_t m = 0u - t[0];
        c = ((uint64_t)m + t[0]) >> 32;                       // m * p0 (= m) + t0: low word is zero
        c += (uint64_t)m * P32_1 + t[1]; t[0] = (uint32_t)c; c >>= 32;
        c += (uint64_t)m * P32_2 + t[2]; t[1] = (uint32_t)c; c >>= 32;
        c += (uint64_t)m * P32_3 + t[3]; t[2] = (uint32_t)c; c >>= 32;
        c += t[4]; t[3] = (uint32_t)c; c >>= 32;
        c += t[5]; t[4] = (uint32_t)c; c >>= 32;
        c += t[6]; t[5] = (uint32_t)c; c >>= 32;
        c += (uint64_t)m * P32_7 + t[7]; t[6] = (uint32_t)c; c >>= 32;
        c += t[8]; t[7] = (uint32_t)c; c >>= 32;
        t[8] = t[9] + (uint32_t)c;
    }
    Fr r;
#pragma unroll
    for (int i = 0; i < 8; i++) r.v[i] = t[i];
    // a, b < p  =>  result < 2p < 2^256, so t[8] == 0
    return fr_csub_p(r);
}
SMW_DEV Fr fr_sqr(const Fr& a) { return fr_mul(a, a); }
SMW_DEV Fr fr_from_u64(uint64_t x) { Fr a = fr_zero(); a.v[0] = (uint32_t)x; a.v[1] = (uint32_t)(x >> 32); return fr_mul(a, fr_const(C_RR32)); }
SMW_DEV Fr fr_pow_u64(const Fr& a, uint64_t e) { Fr r = fr_const(C_R32); for (int b = 63; b >= 0; b--) { r = fr_sqr(r); if ((e >> b) & 1ull) r = fr_mul(r, a); } return r; }
SMW_DEV Fr fr_inv(const Fr& a) {   // a^(p-2)
    const uint32_t e[8] = {0xffffffffu, 0x8c46eb20u, 0x0994a8ddu, 0x224698fcu, 0u, 0u, 0u, 0x40000000u};
    Fr r = fr_const(C_R32);
    for (int k = 7; k >= 0; k--) for (int b = 31; b >= 0; b--) { r = fr_sqr(r); if ((e[k] >> b) & 1u) r = fr_mul(r, a); }
    return r;
}
// canonical integer (< 2^256, reduced here) <-> Montgomery form
SMW_DEV Fr fr_from_canonical(const Fr& x) { Fr y = x; y = fr_csub_p(y); y = fr_csub_p(y); y = fr_csub_p(y); return fr_mul(y, fr_const(C_RR32)); }
SMW_DEV Fr fr_to_canonical(const Fr& a) { Fr one = fr_zero(); one.v[0] = 1; return fr_mul(a, one); }

} // namespace smw
