// fr52.cuh — Pallas Fr arithmetic on the FP64 pipe, for the throughput-bound Poseidon kernels (t = 9, 17).
//
// Why: tools/dfma_peak.cu (profiles/r01n_dfma_vs_imad_product_block.jsonl) shows that a saturated IMAD.WIDE stream is
// throttled inside the SM (the kernel's clock counter advances at 1.49 GHz of wall time at 10 warps/SM), while the same
// partial-product block written with DFMA keeps 1.87 GHz: +26 % products per second.  A field element is 5 limbs of 52
// bits held as exact doubles; a limb product a_i * b_j < 2^104 is split without any integer multiply (Emmart, Zheng,
// Weems, "Faster modular exponentiation using double precision floating point arithmetic on the GPU", ARITH 2018):
//        hi = fma_rz(a, b, 2^104)                 -> bit pattern 0x467<<52 | floor(ab / 2^52)
//        lo = fma_rz(a, b, (2^104 + 2^52) - hi)   -> bit pattern 0x433<<52 | (ab mod 2^52)
// and the raw bit patterns are accumulated as 64-bit integers into 10 column accumulators; the exponent bits add up to a
// constant per term that is removed once per dot product.  Montgomery radix 2^260, values kept lazily reduced (< 2^260,
// limbs < 2^52), one reduction per dot product — same lazy scheme as fr29.cuh, same programs, exact arithmetic, so the
// digests are bit-identical.  The algorithm was checked against a big-integer model before it was written in CUDA.
//
// Replaces ark-ff 0.5.0 Fp256<MontBackend<ark_pallas::FrConfig,4>> mul/add as called from poseidon/src/lib.rs:24-29,219-272.
#pragma once
#include <cstdint>

namespace smw {

#ifndef SMW_DEV
#define SMW_DEV __device__ __forceinline__
#endif

struct F52 { double v[5]; };
struct Acc52 { uint64_t c[10]; };

constexpr uint64_t M52 = 0x000fffffffffffffull;
constexpr uint64_t EXP52 = 0x4330000000000000ull;     // bit pattern of 2^52: (EXP52 | x) is the double 2^52 + x for x < 2^52
constexpr uint64_t EXP104 = 0x4670000000000000ull;    // bit pattern of 2^104
// p = 2^254 + c in radix 2^52: limb 3 is zero, limb 4 is 2^46;  -p^-1 mod 2^52
constexpr uint64_t P52_0 = 0x6eb2100000001ull, P52_1 = 0xfc0994a8dd8c4ull, P52_2 = 0x224698ull, P52_4 = 0x400000000000ull;
constexpr uint64_t PINV52 = 0x6eb20ffffffffull;
// conversion constants (canonical integers, radix 2^52): 2^264, 2^256, 2^520, 2^260 mod p
__device__ __constant__ const uint64_t C_KIN52[5]  = {0xa6720fffffc01ull, 0xd5b6f132ac709ull, 0xfffff7707e2a8ull, 0xfffffffffffffull, 0x3fffffffffffull};
__device__ __constant__ const uint64_t C_KOUT52[5] = {0xb3e9cfffffffdull, 0x0be34205675b2ull, 0xfffffff992c35ull, 0xfffffffffffffull, 0x3fffffffffffull};
__device__ __constant__ const uint64_t C_KRR52[5]  = {0x9033b00000edbull, 0x1bb79a7bf7503ull, 0xcf59062c70f21ull, 0xa97fae231004cull, 0x2d41af7ccfdaull};
__device__ __constant__ const uint64_t C_KONE52[5] = {0xc22e0ffffffc1ull, 0xf9a46a717a7a8ull, 0xffffff790a059ull, 0xfffffffffffffull, 0x3fffffffffffull};

SMW_DEV double u52_to_double(uint64_t x) { return __longlong_as_double((long long)(x | EXP52)) - 0x1p52; }      // exact for x < 2^52
SMW_DEV uint64_t double_to_u52(double d) { return (uint64_t)__double_as_longlong(d + 0x1p52) & M52; }            // exact integer d < 2^52
SMW_DEV F52 f52_zero() { F52 r;
#pragma unroll
    for (int i = 0; i < 5; i++) r.v[i] = 0.0;
    return r; }
SMW_DEV F52 f52_from_limbs(const uint64_t l[5]) { F52 r;
#pragma unroll
    for (int i = 0; i < 5; i++) r.v[i] = u52_to_double(l[i]);
    return r; }
SMW_DEV F52 f52_const(const uint64_t* c) { uint64_t l[5];
#pragma unroll
    for (int i = 0; i < 5; i++) l[i] = c[i];
    return f52_from_limbs(l); }

SMW_DEV void acc52_zero(Acc52& A) {
#pragma unroll
    for (int i = 0; i < 10; i++) A.c[i] = 0;
}
// A += a * b: 25 x (DFMA.RZ, DADD, DFMA.RZ) on the FP64 pipe + 50 64-bit integer adds on the ALU pipe
SMW_DEV void acc52_mac(Acc52& A, const F52& a, const F52& b) {
#pragma unroll
    for (int i = 0; i < 5; i++)
#pragma unroll
        for (int j = 0; j < 5; j++) {
            const double hi = __fma_rz(a.v[i], b.v[j], 0x1p104);
            const double lo = __fma_rz(a.v[i], b.v[j], (0x1p104 + 0x1p52) - hi);
            A.c[i + j + 1] += (uint64_t)__double_as_longlong(hi);
            A.c[i + j] += (uint64_t)__double_as_longlong(lo);
        }
}
SMW_DEV void acc52_mul(Acc52& A, const F52& a, const F52& b) { acc52_zero(A); acc52_mac(A, a, b); }
// A = a * a with 15 limb products instead of 25: an off-diagonal product a_i a_j (i < j) is split once and its two bit patterns are
// added TWICE (a doubled operand would leave the 2^104 window of the split).  Every column receives exactly as many low / high
// patterns as in acc52_mul, so the exponent bookkeeping of acc52_reduce_add2 (n_terms = 1) is unchanged.
SMW_DEV void acc52_sqr(Acc52& A, const F52& a) {
    acc52_zero(A);
#pragma unroll
    for (int i = 0; i < 5; i++)
#pragma unroll
        for (int j = i; j < 5; j++) {
            const double hi = __fma_rz(a.v[i], a.v[j], 0x1p104);
            const double lo = __fma_rz(a.v[i], a.v[j], (0x1p104 + 0x1p52) - hi);
            const uint64_t h = (uint64_t)__double_as_longlong(hi), l = (uint64_t)__double_as_longlong(lo);
            if (i == j) { A.c[i + j + 1] += h; A.c[i + j] += l; }
            else { A.c[i + j + 1] += h + h; A.c[i + j] += l + l; }
        }
}

// Montgomery reduction of an accumulator of `n_terms` products, plus two addends on top of the reduced value:
// returns s + t + (A + M p) / 2^260, limbs < 2^52 (the caller's bounds keep the value below 2^260).
// Exponent bits: one term leaves nlo[k] low parts (0x433 << 52) and nhi[k] high parts (0x467 << 52) in column k, the five
// reduction steps add rlo[k] / rhi[k] more, and the two addends two low parts in columns 5..9.  Only bits 52..63 are touched
// (mod 2^64), so ALL of them are taken out up front with one 32-bit multiply-add on the high word of each column; every
// column is exact again by the time its carry is taken (all its contributions arrive at steps <= its own index).
SMW_DEV F52 acc52_reduce_add2(Acc52& A, uint32_t n_terms, const F52& s, const F52& t) {
    constexpr uint32_t nlo[10] = {1, 2, 3, 4, 5, 4, 3, 2, 1, 0}, nhi[10] = {0, 1, 2, 3, 4, 5, 4, 3, 2, 1};
    constexpr uint32_t rlo[10] = {1, 2, 3, 3, 3, 2, 1, 0, 0, 0}, rhi[10] = {0, 1, 2, 3, 3, 3, 2, 1, 0, 0};
#pragma unroll
    for (int k = 0; k < 10; k++) {
        const uint32_t per_term = (nlo[k] * 0x433u + nhi[k] * 0x467u) << 20;
        const uint32_t fixed = (rlo[k] * 0x433u + rhi[k] * 0x467u + (k >= 5 ? 2u * 0x433u : 0u)) << 20;
        const uint32_t h = (uint32_t)(A.c[k] >> 32) - (per_term * n_terms + fixed);
        A.c[k] = (A.c[k] & 0xffffffffull) | ((uint64_t)h << 32);
    }
#pragma unroll
    for (int i = 0; i < 5; i++)
        A.c[5 + i] += (uint64_t)__double_as_longlong(s.v[i] + 0x1p52) + (uint64_t)__double_as_longlong(t.v[i] + 0x1p52);
    const double p0 = (double)P52_0, p1 = (double)P52_1, p2 = (double)P52_2, pinv = (double)PINV52;
#pragma unroll
    for (int i = 0; i < 5; i++) {
        const double tl = u52_to_double(A.c[i] & M52);
        const double mh = __fma_rz(tl, pinv, 0x1p104);
        const double ml = __fma_rz(tl, pinv, (0x1p104 + 0x1p52) - mh);              // 2^52 + m,  m = -A.c[i] / p mod 2^52
        const uint64_t m = (uint64_t)__double_as_longlong(ml) & M52;
        const double md = ml - 0x1p52;
        const double h0 = __fma_rz(md, p0, 0x1p104), l0 = __fma_rz(md, p0, (0x1p104 + 0x1p52) - h0);
        const double h1 = __fma_rz(md, p1, 0x1p104), l1 = __fma_rz(md, p1, (0x1p104 + 0x1p52) - h1);
        const double h2 = __fma_rz(md, p2, 0x1p104), l2 = __fma_rz(md, p2, (0x1p104 + 0x1p52) - h2);
        A.c[i] += (uint64_t)__double_as_longlong(l0);                                                             // low 52 bits are now zero
        A.c[i + 1] += (uint64_t)__double_as_longlong(h0) + (uint64_t)__double_as_longlong(l1) + (A.c[i] >> 52);
        A.c[i + 2] += (uint64_t)__double_as_longlong(h1) + (uint64_t)__double_as_longlong(l2);
        A.c[i + 3] += (uint64_t)__double_as_longlong(h2);
        A.c[i + 4] += (m & 63ull) << 46;                                                                          // m * 2^46 = (m >> 6) 2^52 + (m & 63) 2^46
        A.c[i + 5] += m >> 6;
    }
    F52 r; uint64_t carry = 0;
#pragma unroll
    for (int k = 0; k < 5; k++) {
        const uint64_t v = A.c[5 + k] + carry;
        r.v[k] = u52_to_double(k < 4 ? (v & M52) : v);
        carry = v >> 52;
    }
    return r;
}
// ---- one product, no addends: the form the S-boxes of the tensor-core epilogues use (poseidon_tc.cuh) -------------------------
// The exponent bits that ONE product's patterns and the five reduction steps will add to each column are known at compile time, so the
// accumulator STARTS at minus that amount (a constant folded into the first addition) instead of having it subtracted from the high
// word of every column before the reduction; with no addends on top, the 3-input additions of two 2^52-biased zeros go as well.
// ~60 ALU-pipe instructions less per x^5 (the ALU pipe is the busiest pipe of those epilogues, profiles/r04a_ncu_tc_stage_summary.txt).
// (low 52 bits of x) as an exact double.  ptxas has no three-input logic instruction with TWO immediates, so `(hi & 0xfffff) | 0x43300000` written with
// constants compiles to two LOP3 on the ALU pipe (the busiest pipe of the S-box epilogues).  The exponent word is therefore read from a device variable
// the compiler cannot fold: one LOP3 (a & imm | register).  ~30 ALU instructions less per x^5.
static __device__ uint32_t g_exp52_hi = 0x43300000u;
SMW_DEV double u52_to_double_m(uint64_t x, uint32_t exp_hi) {
    const uint32_t hi = ((uint32_t)(x >> 32) & 0xfffffu) | exp_hi;
    return __longlong_as_double((long long)(((uint64_t)hi << 32) | (x & 0xffffffffull))) - 0x1p52;
}
#ifndef SMW_SBOX_M46_FP
#define SMW_SBOX_M46_FP 0
#endif
// the top limb of a reduced value: its column can still carry exponent bits above bit 52 only if the bookkeeping is wrong, so it is taken as is
SMW_DEV double u52_to_double_top(uint64_t v) { return u52_to_double(v); }
SMW_DEV void acc52_init_biased(Acc52& A) {
    constexpr uint32_t nlo[10] = {1, 2, 3, 4, 5, 4, 3, 2, 1, 0}, nhi[10] = {0, 1, 2, 3, 4, 5, 4, 3, 2, 1};
#if SMW_SBOX_M46_FP
    // the 2^46 limb of p is split on the FP64 pipe like the others: one more low pattern in columns 4..8, one more high pattern in columns 5..9
    constexpr uint32_t rlo[10] = {1, 2, 3, 3, 4, 3, 2, 1, 1, 0}, rhi[10] = {0, 1, 2, 3, 3, 4, 3, 2, 1, 1};
#else
    constexpr uint32_t rlo[10] = {1, 2, 3, 3, 3, 2, 1, 0, 0, 0}, rhi[10] = {0, 1, 2, 3, 3, 3, 2, 1, 0, 0};
#endif
#pragma unroll
    for (int k = 0; k < 10; k++) {
        const uint32_t bias = ((nlo[k] + rlo[k]) * 0x433u + (nhi[k] + rhi[k]) * 0x467u) << 20;
        A.c[k] = (uint64_t)(0u - bias) << 32;
    }
}
SMW_DEV void acc52_mul_b(Acc52& A, const F52& a, const F52& b) { acc52_init_biased(A); acc52_mac(A, a, b); }
SMW_DEV void acc52_sqr_b(Acc52& A, const F52& a) {
    acc52_init_biased(A);
#pragma unroll
    for (int i = 0; i < 5; i++)
#pragma unroll
        for (int j = i; j < 5; j++) {
            const double hi = __fma_rz(a.v[i], a.v[j], 0x1p104);
            const double lo = __fma_rz(a.v[i], a.v[j], (0x1p104 + 0x1p52) - hi);
            const uint64_t h = (uint64_t)__double_as_longlong(hi), l = (uint64_t)__double_as_longlong(lo);
            if (i == j) { A.c[i + j + 1] += h; A.c[i + j] += l; }
            else { A.c[i + j + 1] += h + h; A.c[i + j] += l + l; }
        }
}
// (A + M p) / 2^260 for an accumulator started by acc52_init_biased and holding one product; limbs < 2^52
SMW_DEV F52 acc52_reduce_b(Acc52& A, uint32_t exp_hi) {
    const double p0 = (double)P52_0, p1 = (double)P52_1, p2 = (double)P52_2, pinv = (double)PINV52;
#pragma unroll
    for (int i = 0; i < 5; i++) {
        const double tl = u52_to_double_m(A.c[i], exp_hi);
        const double mh = __fma_rz(tl, pinv, 0x1p104);
        const double ml = __fma_rz(tl, pinv, (0x1p104 + 0x1p52) - mh);
        const double md = ml - 0x1p52;
        const double h0 = __fma_rz(md, p0, 0x1p104), l0 = __fma_rz(md, p0, (0x1p104 + 0x1p52) - h0);
        const double h1 = __fma_rz(md, p1, 0x1p104), l1 = __fma_rz(md, p1, (0x1p104 + 0x1p52) - h1);
        const double h2 = __fma_rz(md, p2, 0x1p104), l2 = __fma_rz(md, p2, (0x1p104 + 0x1p52) - h2);
        A.c[i] += (uint64_t)__double_as_longlong(l0);
        A.c[i + 1] += (uint64_t)__double_as_longlong(h0) + (uint64_t)__double_as_longlong(l1) + (A.c[i] >> 52);
        A.c[i + 2] += (uint64_t)__double_as_longlong(h1) + (uint64_t)__double_as_longlong(l2);
        A.c[i + 3] += (uint64_t)__double_as_longlong(h2);
#if SMW_SBOX_M46_FP
        const double h4 = __fma_rz(md, 0x1p46, 0x1p104), l4 = __fma_rz(md, 0x1p46, (0x1p104 + 0x1p52) - h4);      // m 2^46 = (m >> 6) 2^52 + (m & 63) 2^46
        A.c[i + 4] += (uint64_t)__double_as_longlong(l4);
        A.c[i + 5] += (uint64_t)__double_as_longlong(h4);
#else
        const uint64_t m = (uint64_t)__double_as_longlong(ml) & M52;
        A.c[i + 4] += (m & 63ull) << 46;
        A.c[i + 5] += m >> 6;
#endif
    }
    F52 r; uint64_t carry = 0;
#pragma unroll
    for (int k = 0; k < 5; k++) {
        const uint64_t v = A.c[5 + k] + carry;
        r.v[k] = k < 4 ? u52_to_double_m(v, exp_hi) : u52_to_double_top(v);
        carry = v >> 52;
    }
    return r;
}

// The same reduction with a SHORTER DEPENDENCY CHAIN, for the cooperative (latency-bound, one lone warp) kernels: the quotient
// digit m = -c_i / p mod 2^52 and the low halves of m * p_j come from 64-bit integer multiplies (IMAD.WIDE + 2 IMAD, ~5 cycles each,
// running beside the FP64 pipe), only the high halves floor(m p_j / 2^52) stay on DFMA.RZ.  Per step the chain to the next column is
// ~9 integer operations instead of ~14 operations of which 8 are dependent FP64 instructions; the instruction count is higher
// (~225 against ~160), which is why the throughput-bound batch interpreter keeps acc52_reduce_add2.  Same value, bit for bit.
SMW_DEV F52 acc52_reduce_add2_lat(Acc52& A, uint32_t n_terms, const F52& s, const F52& t) {
    constexpr uint32_t nlo[10] = {1, 2, 3, 4, 5, 4, 3, 2, 1, 0}, nhi[10] = {0, 1, 2, 3, 4, 5, 4, 3, 2, 1};
    constexpr uint32_t rhi[10] = {0, 1, 2, 3, 3, 3, 2, 1, 0, 0};          // high-half patterns (0x467 << 52) the five steps add to column k; low halves are plain integers here
#pragma unroll
    for (int k = 0; k < 10; k++) {
        const uint32_t per_term = (nlo[k] * 0x433u + nhi[k] * 0x467u) << 20;
        const uint32_t fixed = (rhi[k] * 0x467u + (k >= 5 ? 2u * 0x433u : 0u)) << 20;
        const uint32_t h = (uint32_t)(A.c[k] >> 32) - (per_term * n_terms + fixed);
        A.c[k] = (A.c[k] & 0xffffffffull) | ((uint64_t)h << 32);
    }
#pragma unroll
    for (int i = 0; i < 5; i++)
        A.c[5 + i] += (uint64_t)__double_as_longlong(s.v[i] + 0x1p52) + (uint64_t)__double_as_longlong(t.v[i] + 0x1p52);
    const double p0 = (double)P52_0, p1 = (double)P52_1, p2 = (double)P52_2;
#pragma unroll
    for (int i = 0; i < 5; i++) {
        const uint64_t m = (A.c[i] * PINV52) & M52;                       // only the low 52 bits of the column matter
        const double md = u52_to_double(m);
        const uint64_t h0 = (uint64_t)__double_as_longlong(__fma_rz(md, p0, 0x1p104)), h1 = (uint64_t)__double_as_longlong(__fma_rz(md, p1, 0x1p104)),
                       h2 = (uint64_t)__double_as_longlong(__fma_rz(md, p2, 0x1p104));
        const uint64_t l0 = (m * P52_0) & M52, l1 = (m * P52_1) & M52, l2 = (m * P52_2) & M52;
        A.c[i] += l0;                                                                      // low 52 bits are now zero
        A.c[i + 1] += h0 + l1 + (A.c[i] >> 52);
        A.c[i + 2] += h1 + l2;
        A.c[i + 3] += h2;
        A.c[i + 4] += (m & 63ull) << 46;                                                   // m * 2^46 = (m >> 6) 2^52 + (m & 63) 2^46
        A.c[i + 5] += m >> 6;
    }
    F52 r; uint64_t carry = 0;
#pragma unroll
    for (int k = 0; k < 5; k++) {
        const uint64_t v = A.c[5 + k] + carry;
        r.v[k] = u52_to_double(k < 4 ? (v & M52) : v);
        carry = v >> 52;
    }
    return r;
}
SMW_DEV F52 f52_mul(const F52& a, const F52& b) { Acc52 A; acc52_mul(A, a, b); return acc52_reduce_add2(A, 1, f52_zero(), f52_zero()); }

// ---- boundary conversions -------------------------------------------------------------------
// external: 8 x u32 little-endian words (any 256-bit integer) -> raw 52-bit limbs
SMW_DEV F52 f52_repack_from32(const uint32_t w[8]) {
    const uint64_t x0 = (uint64_t)w[0] | ((uint64_t)w[1] << 32), x1 = (uint64_t)w[2] | ((uint64_t)w[3] << 32),
                   x2 = (uint64_t)w[4] | ((uint64_t)w[5] << 32), x3 = (uint64_t)w[6] | ((uint64_t)w[7] << 32);
    uint64_t l[5];
    l[0] = x0 & M52; l[1] = ((x0 >> 52) | (x1 << 12)) & M52; l[2] = ((x1 >> 40) | (x2 << 24)) & M52; l[3] = ((x2 >> 28) | (x3 << 36)) & M52; l[4] = x3 >> 16;
    return f52_from_limbs(l);
}
// value < 2p with limbs < 2^52 -> canonical integer limbs
SMW_DEV void f52_canonical_limbs(const F52& a, uint64_t l[5]) {
    const uint64_t pl[5] = {P52_0, P52_1, P52_2, 0ull, P52_4};
    uint64_t d[5], borrow = 0;
#pragma unroll
    for (int i = 0; i < 5; i++) {
        l[i] = double_to_u52(a.v[i]);
        const uint64_t t = l[i] - pl[i] - borrow;
        borrow = t >> 63;                          // limbs < 2^52: a negative difference sets the top bit
        d[i] = (i < 4) ? (t & M52) : t;
    }
#pragma unroll
    for (int i = 0; i < 5; i++) l[i] = borrow ? l[i] : d[i];
}
SMW_DEV void f52_limbs_to32(const uint64_t l[5], uint32_t w[8]) {
    const uint64_t x0 = l[0] | (l[1] << 52), x1 = (l[1] >> 12) | (l[2] << 40), x2 = (l[2] >> 24) | (l[3] << 28), x3 = (l[3] >> 36) | (l[4] << 16);
    w[0] = (uint32_t)x0; w[1] = (uint32_t)(x0 >> 32); w[2] = (uint32_t)x1; w[3] = (uint32_t)(x1 >> 32);
    w[4] = (uint32_t)x2; w[5] = (uint32_t)(x2 >> 32); w[6] = (uint32_t)x3; w[7] = (uint32_t)(x3 >> 32);
}
SMW_DEV F52 f52_from_u64(uint64_t x) { uint64_t l[5] = {x & M52, x >> 52, 0, 0, 0}; return f52_from_limbs(l); }   // raw integer, not yet in Montgomery form

} // namespace smw
