// chacha.cuh — rand 0.8.5 `StdRng` (rand_chacha 0.3.1 ChaCha12Rng) on the device: key = 32-byte seed,
// 64-bit block counter from 0, stream id 0; next_u64 = two consecutive output words (lo, hi).
// Used by fri_sample_z_ell (deep_ali/src/fri.rs:64-71), which draws a handful of u64 per level.
#pragma once
#include <cstdint>

namespace smw {

struct ChaChaRng { uint32_t key[8]; uint64_t counter; uint32_t buf[16]; int idx; };

__device__ __forceinline__ uint32_t cc_rotl(uint32_t x, int n) { return (x << n) | (x >> (32 - n)); }
#define SMW_QR(a, b, c, d) a += b; d ^= a; d = cc_rotl(d, 16); c += d; b ^= c; b = cc_rotl(b, 12); a += b; d ^= a; d = cc_rotl(d, 8); c += d; b ^= c; b = cc_rotl(b, 7);
__device__ inline void chacha12_block(const uint32_t key[8], uint64_t ctr, uint32_t out[16]) {
    uint32_t s[16] = {0x61707865u, 0x3320646eu, 0x79622d32u, 0x6b206574u, key[0], key[1], key[2], key[3], key[4], key[5], key[6], key[7],
                      (uint32_t)ctr, (uint32_t)(ctr >> 32), 0u, 0u};
    uint32_t x[16];
    for (int i = 0; i < 16; i++) x[i] = s[i];
    for (int r = 0; r < 6; r++) {
        SMW_QR(x[0], x[4], x[8], x[12]) SMW_QR(x[1], x[5], x[9], x[13]) SMW_QR(x[2], x[6], x[10], x[14]) SMW_QR(x[3], x[7], x[11], x[15])
        SMW_QR(x[0], x[5], x[10], x[15]) SMW_QR(x[1], x[6], x[11], x[12]) SMW_QR(x[2], x[7], x[8], x[13]) SMW_QR(x[3], x[4], x[9], x[14])
    }
    for (int i = 0; i < 16; i++) out[i] = x[i] + s[i];
}
#undef SMW_QR
__device__ inline void chacha_init(ChaChaRng& g, const uint32_t key[8]) { for (int i = 0; i < 8; i++) g.key[i] = key[i]; g.counter = 0; g.idx = 16; }
__device__ inline uint64_t chacha_next_u64(ChaChaRng& g) {   // the 64-word buffer of BlockRng is 4 consecutive blocks, and draws stay word-pair aligned
    if (g.idx >= 16) { chacha12_block(g.key, g.counter, g.buf); g.counter++; g.idx = 0; }
    uint64_t v = (uint64_t)g.buf[g.idx] | ((uint64_t)g.buf[g.idx + 1] << 32);
    g.idx += 2;
    return v;
}

} // namespace smw
