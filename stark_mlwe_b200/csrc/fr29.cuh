// fr29.cuh — Pallas Fr arithmetic for the Poseidon kernels, designed around what the B200 integer
// pipe actually does (tools/imad_peak.cu, profiles/r01_imad_peak.md):
//   * IMAD.WIDE.U32 (32x32+64 -> 64, no carry)  issues at ~2.2 warp-instr/clk/SM,
//   * the carry-predicate form IMAD.WIDE.U32.X (what mad.lo.cc/madc.hi.cc chains become) at ~1.3,
//   * IMAD.HI at ~1.5.
// So the multiplier avoids carry chains altogether: a field element is 9 limbs of 29 bits, partial
// products are summed with plain IMAD.WIDE into 17 independent 64-bit column accumulators (a column
// takes <= 9 products < 2^58 per term, so 6 terms fit before a carry-save normalisation), and a dot
// product of many terms pays for ONE Montgomery reduction.  The Montgomery radix is R' = 2^261, which
// leaves 7 bits of headroom above p ~ 2^254: values are kept lazily reduced (< 2^261) and no
// conditional subtraction is needed anywhere inside a permutation.
//
// Replaces ark-ff 0.5.0 Fp256<MontBackend<ark_pallas::FrConfig,4>> mul/add/square as called from
// poseidon/src/lib.rs:24-29,219-272.  The external representation (what a Rust caller holds in
// memory: 4 x u64 Montgomery limbs, R = 2^256, canonical) is converted at the kernel boundary by
// f29_from_ext / f29_to_ext; results are bit-identical because the arithmetic is exact.
#pragma once
#include <cstdint>

namespace smw {

#define SMW_DEV __device__ __forceinline__

constexpr uint32_t M29 = 0x1fffffffu;
// p = 2^254 + c in radix 2^29: limbs 5..7 are zero, limb 0 is 1, limb 8 is 2^22; -p^-1 mod 2^29 = 2^29-1
constexpr uint32_t P29_1 = 0x02375908u, P29_2 = 0x052a3763u, P29_3 = 0x0d31f813u, P29_4 = 0x00000224u, P29_8 = 0x00400000u;

// external form (8 x u32, Montgomery R = 2^256) of 1: 2^256 mod p
__device__ __constant__ const uint32_t C_R32W[8] = {0xfffffffdu, 0x5b2b3e9cu, 0xe3420567u, 0x992c350bu, 0xffffffffu, 0xffffffffu, 0xffffffffu, 0x3fffffffu};

struct F29 { uint32_t v[9]; };
struct Acc { uint64_t c[17]; };

__device__ __constant__ const uint32_t C_ONE29[9]  = {0x1fffff81u, 0x068ad507u, 0x100e85dau, 0x1435ee7eu, 0x1ffeefefu, 0x1fffffffu, 0x1fffffffu, 0x1fffffffu, 0x003fffffu}; // 2^261 mod p
__device__ __constant__ const uint32_t C_RR29[9]   = {0x00003b6au, 0x02b1b550u, 0x1027888au, 0x1ea4ed96u, 0x0418ad7au, 0x000999ebu, 0x17fae231u, 0x1e67ed54u, 0x003506bdu}; // 2^522 mod p
__device__ __constant__ const uint32_t C_KIN29[9]  = {0x1ffff001u, 0x0ca6d907u, 0x01b40647u, 0x0db0c57eu, 0x1fddbb8bu, 0x1fffffffu, 0x1fffffffu, 0x1fffffffu, 0x003fffffu}; // 2^266 mod p
__device__ __constant__ const uint32_t C_KOUT29[9] = {0x1ffffffdu, 0x1959f4e7u, 0x108159d6u, 0x186a17c6u, 0x1ffff992u, 0x1fffffffu, 0x1fffffffu, 0x1fffffffu, 0x003fffffu}; // 2^256 mod p
__device__ __constant__ const uint32_t C_P29[9]    = {0x00000001u, 0x02375908u, 0x052a3763u, 0x0d31f813u, 0x00000224u, 0x00000000u, 0x00000000u, 0x00000000u, 0x00400000u};

// Written as plain C on purpose: ptxas turns `mul.wide.u32 + add.s64` into ONE accumulating IMAD.WIDE.U32, whereas the
// PTX `mad.wide.u32` spelled through inline asm is split into IMAD.WIDE.U32 (RZ addend) + IADD3 + IADD3.X, which makes
// the ALU pipe, not the multiplier, the bottleneck of every dot product (profiles/r01b_sass_mac_forms.md).
SMW_DEV uint64_t madw(uint32_t a, uint32_t b, uint64_t c) {
#ifdef SMW_MADW_ASM
    uint64_t r;
    asm("mad.wide.u32 %0, %1, %2, %3;" : "=l"(r) : "r"(a), "r"(b), "l"(c));
    return r;
#else
    return (uint64_t)a * b + c;
#endif
}

SMW_DEV F29 f29_const(const uint32_t* c) { F29 r;
#pragma unroll
    for (int i = 0; i < 9; i++) r.v[i] = c[i];
    return r; }
SMW_DEV F29 f29_zero() { F29 r;
#pragma unroll
    for (int i = 0; i < 9; i++) r.v[i] = 0;
    return r; }

SMW_DEV void acc_zero(Acc& A) {
#pragma unroll
    for (int i = 0; i < 17; i++) A.c[i] = 0;
}
// A += a * b : 81 IMAD.WIDE, 17 independent accumulation chains
SMW_DEV void acc_mac(Acc& A, const F29& a, const F29& b) {
#pragma unroll
    for (int i = 0; i < 9; i++)
#pragma unroll
        for (int j = 0; j < 9; j++) A.c[i + j] = madw(a.v[i], b.v[j], A.c[i + j]);
}
// A = a * b : the same 81 products without a prior value (a column's first product is an IMAD.WIDE with RZ addend)
SMW_DEV void acc_mul(Acc& A, const F29& a, const F29& b) {
#pragma unroll
    for (int i = 0; i < 9; i++)
#pragma unroll
        for (int j = 0; j < 9; j++) {
            if (i == 0 || j == 8) A.c[i + j] = (uint64_t)a.v[i] * b.v[j];
            else A.c[i + j] = madw(a.v[i], b.v[j], A.c[i + j]);
        }
}
// A += a * a : 45 IMAD.WIDE (off-diagonal terms use the doubled operand, < 2^30)
SMW_DEV void acc_sqr(Acc& A, const F29& a) {
    uint32_t d[9];
#pragma unroll
    for (int i = 0; i < 9; i++) d[i] = a.v[i] << 1;
#pragma unroll
    for (int i = 0; i < 9; i++) {
        A.c[2 * i] = madw(a.v[i], a.v[i], A.c[2 * i]);
#pragma unroll
        for (int j = i + 1; j < 9; j++) A.c[i + j] = madw(d[i], a.v[j], A.c[i + j]);
    }
}
// carry-save normalisation: columns 0..15 back below 2^29 + 2^35, value unchanged (no ripple).  Column 16 is the
// top: it keeps its own high part (a8*b8 < 2^51 per term leaves it thousands of terms of headroom).
SMW_DEV void acc_normalize(Acc& A) {
    uint64_t hi_prev = 0;
#pragma unroll
    for (int i = 0; i < 16; i++) {
        uint64_t hi = A.c[i] >> 29;
        A.c[i] = (A.c[i] & (uint64_t)M29) + hi_prev;
        hi_prev = hi;
    }
    A.c[16] += hi_prev;
}
// Montgomery reduction: returns (A + M*p) / 2^261, limbs < 2^29 (top limb holds what is left).
// Per step: m = -col mod 2^29 (p0 = 1, -p^-1 = -1), 4 IMAD.WIDE for p1..p4 and one for p8 = 2^22.
SMW_DEV F29 acc_reduce(Acc& A) {
    uint64_t carry = 0;
#pragma unroll
    for (int i = 0; i < 9; i++) {
        uint64_t v = A.c[i] + carry;
        uint32_t m = (0u - (uint32_t)v) & M29;
        carry = (v + m) >> 29;
        A.c[i + 1] = madw(m, P29_1, A.c[i + 1]);
        A.c[i + 2] = madw(m, P29_2, A.c[i + 2]);
        A.c[i + 3] = madw(m, P29_3, A.c[i + 3]);
        A.c[i + 4] = madw(m, P29_4, A.c[i + 4]);
        A.c[i + 8] = madw(m, P29_8, A.c[i + 8]);
    }
    F29 r;
#pragma unroll
    for (int i = 0; i < 8; i++) {
        uint64_t v = A.c[9 + i] + carry;
        r.v[i] = (uint32_t)v & M29;
        carry = v >> 29;
    }
    r.v[8] = (uint32_t)carry;
    return r;
}
// same, plus s added on top of the reduced value: returns s + (A + M*p)/2^261
SMW_DEV F29 acc_reduce_add(Acc& A, const F29& s) {
#pragma unroll
    for (int i = 0; i < 8; i++) A.c[9 + i] += s.v[i];
    F29 r = acc_reduce(A);
    r.v[8] += s.v[8];
    return r;
}
// same with two addends (old lane + round constant)
SMW_DEV F29 acc_reduce_add2(Acc& A, const F29& s, const F29& t) {
#pragma unroll
    for (int i = 0; i < 8; i++) A.c[9 + i] += (uint64_t)(s.v[i] + t.v[i]);
    F29 r = acc_reduce(A);
    r.v[8] += s.v[8] + t.v[8];
    return r;
}
SMW_DEV F29 f29_mul(const F29& a, const F29& b) { Acc A; acc_zero(A); acc_mac(A, a, b); return acc_reduce(A); }
SMW_DEV F29 f29_sqr(const F29& a) { Acc A; acc_zero(A); acc_sqr(A, a); return acc_reduce(A); }
// lazy add: limbs carry-normalised so every limb but the top stays < 2^29
SMW_DEV F29 f29_add(const F29& a, const F29& b) {
    F29 r; uint32_t c = 0;
#pragma unroll
    for (int i = 0; i < 8; i++) { uint32_t t = a.v[i] + b.v[i] + c; r.v[i] = t & M29; c = t >> 29; }
    r.v[8] = a.v[8] + b.v[8] + c;
    return r;
}
// x^5 = x * (x^2)^2  (poseidon/src/lib.rs:24-29)
SMW_DEV F29 f29_sbox5(const F29& x) { F29 x2 = f29_sqr(x); F29 x4 = f29_sqr(x2); return f29_mul(x, x4); }

// ---- boundary conversions -------------------------------------------------------------------
// external: 8 x u32 little-endian limbs, Montgomery R = 2^256, canonical (ark-ff memory image)
SMW_DEV F29 f29_repack_from32(const uint32_t w[8]) {
    F29 r;
#pragma unroll
    for (int k = 0; k < 9; k++) {
        const int bit = 29 * k, wi = bit >> 5, sh = bit & 31;
        uint32_t lo = w[wi], hi = (wi + 1 < 8) ? w[wi + 1] : 0u;
        r.v[k] = (sh == 0 ? lo : __funnelshift_r(lo, hi, sh)) & M29;
    }
    return r;
}
SMW_DEV void f29_repack_to32(const F29& a, uint32_t w[8]) {   // a canonical (< p), limbs < 2^29
#pragma unroll
    for (int i = 0; i < 8; i++) {
        // word i = bits [32i, 32i+32): limb k0 = (32i)/29 at offset (32i)%29, plus the next limb(s)
        const int bit = 32 * i, k0 = bit / 29, off = bit % 29;
        uint32_t x = a.v[k0] >> off;
        x |= a.v[k0 + 1] << (29 - off);
        if (58 - off < 32 && k0 + 2 < 9) x |= a.v[k0 + 2] << (58 - off);
        w[i] = x;
    }
}
SMW_DEV F29 f29_from_ext(const uint32_t w[8]) {   // x*2^256 (canonical) -> x*2^261 (lazy)
    return f29_mul(f29_repack_from32(w), f29_const(C_KIN29));
}
// value < 2p with limbs < 2^29 (top limb small) -> canonical
SMW_DEV F29 f29_canonicalize(const F29& a) {
    // d = a - p with borrow chain in radix 2^29
    F29 d; uint32_t borrow = 0;
#pragma unroll
    for (int i = 0; i < 9; i++) {
        uint32_t pi = (i == 0) ? 1u : (i == 1) ? P29_1 : (i == 2) ? P29_2 : (i == 3) ? P29_3 : (i == 4) ? P29_4 : (i == 8) ? P29_8 : 0u;
        uint32_t t = a.v[i] - pi - borrow;
        borrow = (t >> 31);                 // limbs < 2^30 here, so a negative difference sets bit 31
        d.v[i] = (i < 8) ? (t & M29) : t;
    }
    F29 r;
#pragma unroll
    for (int i = 0; i < 9; i++) r.v[i] = borrow ? a.v[i] : d.v[i];
    return r;
}
SMW_DEV void f29_to_ext(const F29& a, uint32_t w[8]) {   // x*2^261 (lazy, < 2^261) -> x*2^256 canonical
    F29 t = f29_mul(a, f29_const(C_KOUT29));    // < a*K/2^261 + p < 2p
    t = f29_canonicalize(t);
    f29_repack_to32(t, w);
}
// small integers (DS labels: arity, level, position, tree_label) -> internal form
SMW_DEV F29 f29_from_u64(uint64_t x) {
    F29 a = f29_zero();
    a.v[0] = (uint32_t)x & M29; a.v[1] = (uint32_t)(x >> 29) & M29; a.v[2] = (uint32_t)(x >> 58);
    return f29_mul(a, f29_const(C_RR29));
}

// ---- init-time helpers (parameter derivation only; not on the hashing path) -------------------
// any lazy value (< 2^261) -> canonical representative, same Montgomery form
SMW_DEV F29 f29_full_reduce(const F29& a) { return f29_canonicalize(f29_mul(a, f29_const(C_ONE29))); }
SMW_DEV bool f29_is_zero_canon(const F29& a) { uint32_t o = 0;
#pragma unroll
    for (int i = 0; i < 9; i++) o |= a.v[i];
    return o == 0; }
// canonical a, b -> canonical a - b
SMW_DEV F29 f29_sub_canon(const F29& a, const F29& b) {
    // a + (p - b), then one conditional subtraction
    F29 nb; uint32_t borrow = 0;
#pragma unroll
    for (int i = 0; i < 9; i++) { uint32_t t = C_P29[i] - b.v[i] - borrow; borrow = t >> 31; nb.v[i] = (i < 8) ? (t & M29) : t; }
    return f29_canonicalize(f29_add(a, nb));
}
// canonical a != 0 -> canonical a^-1 (Fermat, exponent p - 2)
SMW_DEV F29 f29_inv_canon(const F29& a) {
    // p - 2 in radix 2^29: limb0 = 2^29 - 1 with a borrow out of limb 1 (p0 = 1)
    const uint32_t e[9] = {0x1fffffffu, P29_1 - 1u, P29_2, P29_3, P29_4, 0u, 0u, 0u, P29_8};
    F29 r = f29_const(C_ONE29);
    for (int k = 8; k >= 0; k--) {
        int nb = (k == 8) ? 23 : 29;
        for (int bit = nb - 1; bit >= 0; bit--) {
            r = f29_sqr(r);
            if ((e[k] >> bit) & 1u) r = f29_mul(r, a);
        }
    }
    return f29_full_reduce(r);
}

} // namespace smw
