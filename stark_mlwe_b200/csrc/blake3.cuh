// blake3.cuh — single-chunk BLAKE3 (default hash mode, <= 1024 input bytes) on the device.
// Used only by the Poseidon constant derivation (utils/src/lib.rs:16-22 fr_from_hash): every input
// there is tag || u64le indices || seed, well under one chunk.
#pragma once
#include <cstdint>

namespace smw {

__device__ __forceinline__ uint32_t b3_rotr(uint32_t x, int n) { return (x >> n) | (x << (32 - n)); }

__device__ inline void b3_compress(const uint32_t cv[8], const uint32_t blk[16], uint32_t blen, uint32_t flags, uint32_t out[8]) {
    const uint32_t IV[8] = {0x6A09E667u, 0xBB67AE85u, 0x3C6EF372u, 0xA54FF53Au, 0x510E527Fu, 0x9B05688Cu, 0x1F83D9ABu, 0x5BE0CD19u};
    const unsigned char PERM[16] = {2, 6, 3, 10, 7, 0, 4, 13, 1, 11, 12, 5, 9, 14, 15, 8};
    uint32_t s[16], m[16], t[16];
    for (int i = 0; i < 8; i++) s[i] = cv[i];
    for (int i = 0; i < 4; i++) s[8 + i] = IV[i];
    s[12] = 0; s[13] = 0; s[14] = blen; s[15] = flags;   // chunk counter 0: single chunk
    for (int i = 0; i < 16; i++) m[i] = blk[i];
#define B3G(a, b, c, d, x, y)                                                                   \
    s[a] = s[a] + s[b] + (x); s[d] = b3_rotr(s[d] ^ s[a], 16); s[c] = s[c] + s[d]; s[b] = b3_rotr(s[b] ^ s[c], 12); \
    s[a] = s[a] + s[b] + (y); s[d] = b3_rotr(s[d] ^ s[a], 8);  s[c] = s[c] + s[d]; s[b] = b3_rotr(s[b] ^ s[c], 7);
    for (int r = 0; r < 7; r++) {
        B3G(0, 4, 8, 12, m[0], m[1]) B3G(1, 5, 9, 13, m[2], m[3]) B3G(2, 6, 10, 14, m[4], m[5]) B3G(3, 7, 11, 15, m[6], m[7])
        B3G(0, 5, 10, 15, m[8], m[9]) B3G(1, 6, 11, 12, m[10], m[11]) B3G(2, 7, 8, 13, m[12], m[13]) B3G(3, 4, 9, 14, m[14], m[15])
        for (int i = 0; i < 16; i++) t[i] = m[PERM[i]];
        for (int i = 0; i < 16; i++) m[i] = t[i];
    }
#undef B3G
    for (int i = 0; i < 8; i++) out[i] = s[i] ^ s[i + 8];
}

// digest words (little-endian) of msg[0..len), len <= 1024
__device__ inline void blake3_chunk(const unsigned char* msg, int len, uint32_t out[8]) {
    const uint32_t IV[8] = {0x6A09E667u, 0xBB67AE85u, 0x3C6EF372u, 0xA54FF53Au, 0x510E527Fu, 0x9B05688Cu, 0x1F83D9ABu, 0x5BE0CD19u};
    uint32_t cv[8];
    for (int i = 0; i < 8; i++) cv[i] = IV[i];
    int nblocks = len == 0 ? 1 : (len + 63) / 64;
    for (int b = 0; b < nblocks; b++) {
        uint32_t w[16];
        int off = b * 64, bl = len - off > 64 ? 64 : len - off;
        for (int i = 0; i < 16; i++) {
            uint32_t x = 0;
            for (int k = 0; k < 4; k++) { int p = 4 * i + k; if (p < bl) x |= (uint32_t)msg[off + p] << (8 * k); }
            w[i] = x;
        }
        uint32_t flags = (b == 0 ? 1u : 0u) | (b == nblocks - 1 ? (2u | 8u) : 0u);   // CHUNK_START, CHUNK_END | ROOT
        uint32_t o[8];
        b3_compress(cv, w, (uint32_t)bl, flags, o);
        for (int i = 0; i < 8; i++) cv[i] = o[i];
    }
    for (int i = 0; i < 8; i++) out[i] = cv[i];
}

} // namespace smw
