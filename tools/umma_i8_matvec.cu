// tools/umma_i8_matvec.cu — bounded experiment: the dense constant-matrix product of a Poseidon full round,
//      y[p][i] = sum_j M[i][j] * x[p][j]            (t = 17, entries and state words < 2^256, exact integers, 17 terms per output)
// as an INT8 tensor-core GEMM on sm_100a (tcgen05.mma kind::i8, accumulators in TMEM), measured against the 25-DFMA limb
// product of run_ops52 (csrc/poseidon52.cuh).  Micro-benchmark only: nothing in libsmw.so links or calls this file.
//
// Formulation.  A word is 32 unsigned 8-bit limbs.  A (M x K, K-major) holds one permutation per row, K = (j, a) = 17 x 32 = 544
// bytes; B (N x K, K-major) is the Toeplitz expansion of the constant matrix, N = (i, c) = 17 x 64 output columns:
//      B[(i, c)][(j, a)] = limb_{c - a}(M[i][j])   (0 when c - a is outside 0..31)
// so that D[p][(i, c)] = sum_j sum_a x[p][j][a] * M[i][j][c - a] is column c (weight 2^(8c)) of the exact 512-bit row product;
// a column is a sum of <= 17 * 32 products < 2^16, < 2^25.1 — exact in the int32 accumulator.  One k-step of the MMA (K = 32
// bytes) is exactly one input word j.  The epilogue reads a word's 64 columns from TMEM, packs them into ten 52-bit limb
// columns and (mode 1) runs the library's own Montgomery reduction (fr52.cuh acc52_reduce_add2) — the 17 products and their
// 50-instruction additions per term are what the tensor core replaces.
//
// Kernel: persistent, one CTA per SM, 192 threads.  warp 0: one lane streams A tiles and B blocks with cp.async.bulk (the host
// writes both in the UMMA no-swizzle K-major core-matrix layout, so a block is one contiguous bulk copy); warp 1: allocates
// TMEM, one lane issues the MMAs (M = 128, N = 256 / 64, K = 32) and commits to mbarriers; warps 2-5: epilogue, one TMEM lane
// (= one permutation) per thread, double-buffered against the MMA through two 256-column accumulator stages.
// Every mbarrier wait is bounded: a protocol error ends the kernel with an error flag instead of hanging the box.
//
// Build:  nvcc -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -lineinfo -I stark_mlwe_b200/csrc -o tools/_bin/umma_i8 tools/umma_i8_matvec.cu
// Run:    tools/_bin/umma_i8 [tiles_per_cta=16] [dump.bin]      (dump: inputs + tile 0 outputs, checked by tools/umma_i8_check.py)
#include <cstdio>
#include <cstdlib>
#include <cstdint>
#include <cstring>
#include <vector>
#include <cuda_runtime.h>
#include "fr52.cuh"

#define CHK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { printf("{\"error\": \"%s at line %d\"}\n", cudaGetErrorString(e_), __LINE__); return 1; } } while (0)

namespace {
constexpr int T = 17, LIMBS = 32, KTOT = T * LIMBS;          // 544 bytes per row of A
constexpr int KSTEPS = T;                                    // one MMA k-step (32 bytes) per input word
constexpr int WCOLS = 64;                                    // output columns per word (63 used)
constexpr int TILE_M = 128;                                  // permutations per tile = TMEM lanes
constexpr int NCHUNK = 5;                                    // N chunks per tile: 4 x 256 columns (4 words each) + 1 x 64 (1 word)
constexpr int A_KSTEP_BYTES = TILE_M * 32;                   // 4096: [k16 (2)][m8 (16)][8 rows x 16 B]
constexpr int A_TILE_BYTES = KSTEPS * A_KSTEP_BYTES;         // 69 632
constexpr int B_BLOCK_MAX = 256 * 32;                        // 8192: [k16 (2)][n8 (32)][8 rows x 16 B]
#ifndef KSUB
#define KSUB 1
#endif
constexpr int NSTAGE = 8 / KSUB;                             // B stages in flight, KSUB k-steps (one bulk copy, one commit) each
constexpr int NTHREADS = 192;
__host__ __device__ constexpr int chunk_n(int c) { return c < 4 ? 256 : 64; }
__host__ __device__ constexpr int chunk_words(int c) { return c < 4 ? 4 : 1; }
__host__ __device__ constexpr size_t chunk_offset(int c) { return (size_t)c * KSTEPS * B_BLOCK_MAX; }     // chunk 4's blocks are 2048 B each, after four full chunks
constexpr size_t B_TOTAL_BYTES = 4 * (size_t)KSTEPS * B_BLOCK_MAX + (size_t)KSTEPS * 64 * 32;

struct Smem {
    alignas(1024) uint8_t a[2][A_TILE_BYTES];
    alignas(1024) uint8_t b[NSTAGE][KSUB * B_BLOCK_MAX];
    alignas(8) uint64_t b_full[NSTAGE], b_empty[NSTAGE], a_full[2], a_empty[2], acc_full[2], acc_empty[2];
    uint32_t tmem_base;
};

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t* b, uint32_t count) { asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" :: "r"(smem_u32(b)), "r"(count)); }
__device__ __forceinline__ void mbar_expect_tx(uint64_t* b, uint32_t bytes) { asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" :: "r"(smem_u32(b)), "r"(bytes) : "memory"); }
__device__ __forceinline__ void mbar_arrive(uint64_t* b) { asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" :: "r"(smem_u32(b)) : "memory"); }
// bounded wait (non-blocking test, ~2^24 polls = a fraction of a second): a protocol bug ends the kernel instead of hanging it
__device__ __forceinline__ bool mbar_wait(uint64_t* b, uint32_t parity) {
    const uint32_t addr = smem_u32(b);
    for (uint32_t spin = 0; spin < (1u << 24); spin++) {
        uint32_t ok;
        asm volatile("{\n\t.reg .pred p;\n\tmbarrier.test_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}" : "=r"(ok) : "r"(addr), "r"(parity) : "memory");
        if (ok) return true;
    }
    return false;
}
__device__ __forceinline__ void bulk_g2s(void* dst, const void* src, uint32_t bytes, uint64_t* bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 :: "r"(smem_u32(dst)), "l"(src), "r"(bytes), "r"(smem_u32(bar)) : "memory");
}
// UMMA shared-memory descriptor, K-major, no swizzle: 8 rows x 16 B core matrices; LBO = byte distance between the two 16-byte
// K halves of one k-step, SBO = byte distance between consecutive 8-row groups (cute/arch/mma_sm100_desc.hpp SmemDescriptor)
__device__ __forceinline__ uint64_t umma_desc(uint32_t saddr, uint32_t lbo, uint32_t sbo) {
    return (uint64_t)((saddr & 0x3ffffu) >> 4) | ((uint64_t)(lbo >> 4) << 16) | ((uint64_t)(sbo >> 4) << 32) | (1ull << 46);
}
// instruction descriptor, kind::i8: D = s32 (bits 4-5 = 2), A = B = unsigned 8 bit (formats 0), both K-major, N >> 3 at bit 17, M >> 4 at bit 24
__host__ __device__ constexpr uint32_t umma_idesc(int m, int n) { return (2u << 4) | ((uint32_t)(n >> 3) << 17) | ((uint32_t)(m >> 4) << 24); }
__device__ __forceinline__ void umma_i8(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
    asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\ttcgen05.mma.cta_group::1.kind::i8 [%0], %1, %2, %3, p;\n\t}"
                 :: "r"(tmem_d), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate) : "memory");
}
// timing-only variant: the same operands read as e4m3 with an f32 accumulator (kind::f8f6f4) — tells whether the int8 kind is paced differently
__device__ __forceinline__ void umma_f8(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
    asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\ttcgen05.mma.cta_group::1.kind::f8f6f4 [%0], %1, %2, %3, p;\n\t}"
                 :: "r"(tmem_d), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate) : "memory");
}
__device__ __forceinline__ void umma_commit(uint64_t* bar) { asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" :: "r"(smem_u32(bar)) : "memory"); }
__device__ __forceinline__ bool elect_one() {
    uint32_t pred;
    asm volatile("{\n\t.reg .b32 rx;\n\t.reg .pred px;\n\telect.sync rx|px, 0xffffffff;\n\tselp.b32 %0, 1, 0, px;\n\t}" : "=r"(pred));
    return pred != 0;
}
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
// 16 consecutive 32-bit TMEM columns of this thread's lane
__device__ __forceinline__ void tmem_ld16(uint32_t taddr, uint32_t* r) {
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
                   "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
                 : "r"(taddr));
}
__device__ __forceinline__ void tmem_ld_wait() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }

// 64 int32 columns (weights 2^(8k), each < 2^26) -> ten lazily carried 52-bit limb columns (weights 2^(52 l)), each < 2^60.
// Column 63 is structurally zero (the highest limb pair is 31 + 31).  All shifts and masks are compile-time constants.
__device__ __forceinline__ void columns_to_limbs(const uint32_t* c, uint64_t* acc) {
#pragma unroll
    for (int l = 0; l < 10; l++) acc[l] = 0;
#pragma unroll
    for (int k = 0; k < 63; k++) {
        const int bit = 8 * k, l = bit / 52, sh = bit % 52;
        const uint64_t v = c[k];
        if (sh <= 30) acc[l] += v << sh;
        else { acc[l] += (v & ((1ull << (52 - sh)) - 1ull)) << sh; acc[l + 1] += v >> (52 - sh); }
    }
}

struct Args {
    const uint8_t* a_tiles; int n_a_tiles;      // A tiles, UMMA core-matrix layout, A_TILE_BYTES each
    const uint8_t* b_blocks;                    // Toeplitz blocks [chunk][k-step], UMMA core-matrix layout
    int tiles_per_cta, mode, flags;   // mode 0: columns -> limbs; 1: + Montgomery reduction (fr52); 2: MMA only (epilogue reads one column)
    uint64_t* checksum;                         // [grid * 128]
    uint64_t* raw_out; uint64_t* red_out;       // tile 0: [128][17][10] limbs, [128][17][5] double bit patterns
    int* err;
};

__global__ void __launch_bounds__(NTHREADS, 1) k_matvec_umma(Args g) {
    extern __shared__ __align__(1024) uint8_t smem_raw[];
    Smem& S = *reinterpret_cast<Smem*>(smem_raw);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (threadIdx.x == 0) {
        for (int i = 0; i < NSTAGE; i++) { mbar_init(&S.b_full[i], 1); mbar_init(&S.b_empty[i], 1); }
        for (int i = 0; i < 2; i++) { mbar_init(&S.a_full[i], 1); mbar_init(&S.a_empty[i], 1); mbar_init(&S.acc_full[i], 1); mbar_init(&S.acc_empty[i], 128); }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 1) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" :: "r"(smem_u32(&S.tmem_base)), "r"(512u) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = S.tmem_base;

    if (warp == 0 && lane == 0) {
        // ---------------- producer: A tile, then its 5 x 17 B blocks, through the stage ring
        uint32_t bit = 0;
        for (int it = 0; it < g.tiles_per_cta; it++) {
            const int ab = it & 1;
            if (!mbar_wait(&S.a_empty[ab], ((it >> 1) & 1) ^ 1)) { atomicExch(g.err, 1); break; }
            const size_t tile = ((size_t)blockIdx.x + (size_t)it * gridDim.x) % (size_t)g.n_a_tiles;
            mbar_expect_tx(&S.a_full[ab], A_TILE_BYTES);
            bulk_g2s(S.a[ab], g.a_tiles + tile * A_TILE_BYTES, A_TILE_BYTES, &S.a_full[ab]);
            bool bad = false;
            for (int c = 0; c < NCHUNK && !bad; c++) {
                const uint32_t bytes = (uint32_t)chunk_n(c) * 32u;
                for (int s = 0; s < KSTEPS; s += KSUB, bit++) {
                    const int st = bit % NSTAGE, ks = min(KSUB, KSTEPS - s);
                    if (!mbar_wait(&S.b_empty[st], ((bit / NSTAGE) & 1) ^ 1)) { atomicExch(g.err, 2); bad = true; break; }
                    if (g.flags & 2) { mbar_arrive(&S.b_full[st]); continue; }
                    mbar_expect_tx(&S.b_full[st], bytes * ks);
                    bulk_g2s(S.b[st], g.b_blocks + ((g.flags & 1) ? (size_t)blockIdx.x * B_TOTAL_BYTES : 0) + chunk_offset(c) + (size_t)s * bytes, bytes * ks, &S.b_full[st]);
                }
            }
            if (bad) break;
        }
    } else if (warp == 1) {
        // ---------------- MMA issuer: the WHOLE warp runs the loop converged (descriptors stay in uniform registers), one elected lane issues
        uint32_t bit = 0, accit = 0;
        const bool leader = elect_one();
        for (int it = 0; it < g.tiles_per_cta; it++) {
            const int ab = it & 1;
            if (!mbar_wait(&S.a_full[ab], (it >> 1) & 1)) { if (leader) atomicExch(g.err, 3); break; }
            tc_fence_after();
            const uint32_t a_lo = (smem_u32(S.a[ab]) & 0x3ffffu) >> 4;
            bool bad = false;
            for (int c = 0; c < NCHUNK && !bad; c++, accit++) {
                const int as = accit & 1, n = chunk_n(c);
                if (!mbar_wait(&S.acc_empty[as], ((accit >> 1) & 1) ^ 1)) { if (leader) atomicExch(g.err, 4); bad = true; break; }
                tc_fence_after();
                const uint32_t idesc = umma_idesc(TILE_M, n);
                const uint32_t a_hi = (uint32_t)((TILE_M * 16) >> 4) << 16, hi = (128u >> 4) | (1u << 14), b_hi_lbo = (uint32_t)((n * 16) >> 4) << 16;
                for (int s = 0; s < KSTEPS; s += KSUB, bit++) {
                    const int st = bit % NSTAGE, ks = min(KSUB, KSTEPS - s);
                    if (!mbar_wait(&S.b_full[st], (bit / NSTAGE) & 1)) { if (leader) atomicExch(g.err, 5); bad = true; break; }
                    tc_fence_after();
                    const uint32_t b_lo = (smem_u32(S.b[st]) & 0x3ffffu) >> 4;
                    if (leader) {
#pragma unroll
                        for (int k = 0; k < KSUB; k++) {
                            if (k < ks) {
                                const uint64_t ad = (uint64_t)((a_lo + (uint32_t)(s + k) * (A_KSTEP_BYTES >> 4)) | a_hi) | ((uint64_t)hi << 32);
                                const uint64_t bd = (uint64_t)((b_lo + (uint32_t)k * (uint32_t)(n * 2)) | b_hi_lbo) | ((uint64_t)hi << 32);
                                umma_i8(tmem + (uint32_t)as * 256u, ad, bd, idesc, (s + k) > 0 ? 1u : 0u);
                            }
                        }
                        umma_commit(&S.b_empty[st]);
                    }
                    __syncwarp();
                }
                if (!bad && leader) umma_commit(&S.acc_full[as]);
                __syncwarp();
            }
            if (bad) break;
            if (leader) umma_commit(&S.a_empty[ab]);
            __syncwarp();
        }
    } else if (warp >= 2) {
        // ---------------- epilogue: TMEM lane quarter warp % 4, one permutation per thread
        const int q = warp & 3, row = 32 * q + lane;
        uint32_t accit = 0; uint64_t chk = 0; bool bad = false;
        for (int it = 0; it < g.tiles_per_cta && !bad; it++) {
            const bool dump = (blockIdx.x == 0 && it == 0 && g.raw_out != nullptr);
            for (int c = 0; c < NCHUNK; c++, accit++) {
                const int as = accit & 1;
                bool ok = mbar_wait(&S.acc_full[as], (accit >> 1) & 1);
                ok = __all_sync(0xffffffffu, ok);
                if (!ok) { if (lane == 0) atomicExch(g.err, 6); bad = true; break; }
                tc_fence_after();
                for (int w = 0; w < chunk_words(c); w++) {
                    const uint32_t taddr = tmem + ((uint32_t)(32 * q) << 16) + (uint32_t)as * 256u + (uint32_t)w * WCOLS;
                    if (g.mode == 2) { uint32_t r[16]; tmem_ld16(taddr, r); tmem_ld_wait(); chk ^= r[0] + r[15]; continue; }
                    uint32_t col[64];
                    tmem_ld16(taddr, col); tmem_ld16(taddr + 16, col + 16); tmem_ld16(taddr + 32, col + 32); tmem_ld16(taddr + 48, col + 48);
                    tmem_ld_wait();
                    uint64_t acc[10];
                    columns_to_limbs(col, acc);
                    const int word = 4 * c + w;
                    if (g.mode == 0 || dump) {
                        uint64_t l[10];
#pragma unroll
                        for (int i = 0; i < 10; i++) l[i] = acc[i];
#pragma unroll
                        for (int i = 0; i < 9; i++) { l[i + 1] += l[i] >> 52; l[i] &= smw::M52; }
#pragma unroll
                        for (int i = 0; i < 10; i++) chk ^= l[i] * (uint64_t)(2 * i + 3);
                        if (dump) for (int i = 0; i < 10; i++) g.raw_out[((size_t)row * T + word) * 10 + i] = l[i];
                    }
                    if (g.mode == 1) {
                        smw::Acc52 A;
#pragma unroll
                        for (int i = 0; i < 10; i++) A.c[i] = acc[i];
                        const smw::F52 r = smw::acc52_reduce_add2(A, 0, smw::f52_zero(), smw::f52_zero());
#pragma unroll
                        for (int i = 0; i < 5; i++) chk ^= (uint64_t)__double_as_longlong(r.v[i]) * (uint64_t)(2 * i + 5);
                        if (dump) for (int i = 0; i < 5; i++) g.red_out[((size_t)row * T + word) * 5 + i] = (uint64_t)__double_as_longlong(r.v[i]);
                    }
                }
                tc_fence_before();
                mbar_arrive(&S.acc_empty[as]);
            }
        }
        g.checksum[(size_t)blockIdx.x * TILE_M + row] = chk;
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 1) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" :: "r"(tmem), "r"(512u) : "memory");
}
} // namespace

// ------------------------------------------------------------------ host
static uint64_t rng_state = 0x9e3779b97f4a7c15ull;
static uint64_t rnd64() { rng_state ^= rng_state << 13; rng_state ^= rng_state >> 7; rng_state ^= rng_state << 17; return rng_state; }

// UMMA no-swizzle K-major core-matrix position of element (row, k) inside one k-step block of `rows` rows x 32 bytes
static size_t core_pos(int rows, int row, int k) { return (size_t)(k / 16) * rows * 16 + (size_t)(row / 8) * 128 + (size_t)(row % 8) * 16 + (size_t)(k % 16); }

int main(int argc, char** argv) {
    const int tiles_per_cta = argc > 1 ? atoi(argv[1]) : 16;
    const char* dump_path = argc > 2 ? argv[2] : nullptr;
    cudaDeviceProp prop; CHK(cudaGetDeviceProperties(&prop, 0));
    const int grid = prop.multiProcessorCount, n_a_tiles = 2 * grid;
    // constant matrix: 17 x 17 entries of 32 limbs (any 256-bit values: exactness of the integer product is what is checked;
    // top limb < 0x40 so that an entry is < 2^254 like a canonical Pallas Fr constant)
    std::vector<uint8_t> mlimb((size_t)T * T * LIMBS), x((size_t)n_a_tiles * TILE_M * T * LIMBS);
    for (auto& v : mlimb) v = (uint8_t)rnd64();
    for (int e = 0; e < T * T; e++) mlimb[(size_t)e * LIMBS + 31] &= 0x3f;
    for (auto& v : x) v = (uint8_t)rnd64();
    // corner cases in tile 0: all-ones words and zero words
    for (int j = 0; j < T; j++) for (int a = 0; a < LIMBS; a++) { x[((size_t)0 * T + j) * LIMBS + a] = 0xff; x[((size_t)1 * T + j) * LIMBS + a] = 0; }
    // A tiles: [tile][k-step j][core layout over (row, a)]
    std::vector<uint8_t> a_tiles((size_t)n_a_tiles * A_TILE_BYTES);
    for (int t = 0; t < n_a_tiles; t++)
        for (int r = 0; r < TILE_M; r++)
            for (int j = 0; j < T; j++)
                for (int a = 0; a < LIMBS; a++)
                    a_tiles[(size_t)t * A_TILE_BYTES + (size_t)j * A_KSTEP_BYTES + core_pos(TILE_M, r, a)] = x[(((size_t)t * TILE_M + r) * T + j) * LIMBS + a];
    // B blocks: [chunk][k-step j][core layout over (n = (word in chunk, column c), a)]
    std::vector<uint8_t> b_blocks(B_TOTAL_BYTES, 0);
    for (int c = 0; c < NCHUNK; c++)
        for (int j = 0; j < T; j++) {
            const int n_rows = chunk_n(c);
            uint8_t* blk = b_blocks.data() + chunk_offset(c) + (size_t)j * n_rows * 32;
            for (int n = 0; n < n_rows; n++) {
                const int i = 4 * c + n / WCOLS, col = n % WCOLS;
                for (int a = 0; a < LIMBS; a++) {
                    const int d = col - a;
                    blk[core_pos(n_rows, n, a)] = (i < T && d >= 0 && d < LIMBS) ? mlimb[((size_t)i * T + j) * LIMBS + d] : 0;
                }
            }
        }
    uint8_t *d_a, *d_b; uint64_t *d_chk, *d_raw, *d_red; int* d_err;
    CHK(cudaMalloc(&d_a, a_tiles.size())); CHK(cudaMalloc(&d_b, b_blocks.size() * (size_t)grid));
    CHK(cudaMalloc(&d_chk, (size_t)grid * TILE_M * 8)); CHK(cudaMalloc(&d_raw, (size_t)TILE_M * T * 10 * 8)); CHK(cudaMalloc(&d_red, (size_t)TILE_M * T * 5 * 8));
    CHK(cudaMalloc(&d_err, 4));
    CHK(cudaMemcpy(d_a, a_tiles.data(), a_tiles.size(), cudaMemcpyHostToDevice)); for (int i = 0; i < grid; i++) CHK(cudaMemcpy(d_b + (size_t)i * b_blocks.size(), b_blocks.data(), b_blocks.size(), cudaMemcpyHostToDevice));
    CHK(cudaMemset(d_err, 0, 4)); CHK(cudaMemset(d_raw, 0, (size_t)TILE_M * T * 10 * 8)); CHK(cudaMemset(d_red, 0, (size_t)TILE_M * T * 5 * 8));
    const size_t smem = sizeof(Smem) + 1024;
    CHK(cudaFuncSetAttribute(k_matvec_umma, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    cudaEvent_t e0, e1; CHK(cudaEventCreate(&e0)); CHK(cudaEventCreate(&e1));
    printf("{\"tool\": \"umma_i8_matvec\", \"device\": \"%s\", \"sms\": %d, \"tiles_per_cta\": %d, \"smem_bytes\": %zu, \"b_stream_bytes_per_tile\": %zu", prop.name, grid, tiles_per_cta, smem, B_TOTAL_BYTES);
    // correctness pass: mode 1 with the dump of tile 0 (raw limbs and reduced values)
    {
        Args g{d_a, n_a_tiles, d_b, 1, 1, 0, d_chk, d_raw, d_red, d_err};
        k_matvec_umma<<<grid, NTHREADS, smem>>>(g);
        CHK(cudaDeviceSynchronize());
        int err = 0; CHK(cudaMemcpy(&err, d_err, 4, cudaMemcpyDeviceToHost));
        printf(", \"protocol_error\": %d", err);
        if (err) { printf("}\n"); return 2; }
        if (dump_path) {
            std::vector<uint64_t> raw((size_t)TILE_M * T * 10), red((size_t)TILE_M * T * 5);
            CHK(cudaMemcpy(raw.data(), d_raw, raw.size() * 8, cudaMemcpyDeviceToHost)); CHK(cudaMemcpy(red.data(), d_red, red.size() * 8, cudaMemcpyDeviceToHost));
            FILE* f = fopen(dump_path, "wb");
            if (!f) { printf(", \"dump\": \"cannot open\"}\n"); return 3; }
            fwrite(mlimb.data(), 1, mlimb.size(), f); fwrite(x.data(), 1, (size_t)TILE_M * T * LIMBS, f);
            fwrite(raw.data(), 8, raw.size(), f); fwrite(red.data(), 8, red.size(), f); fclose(f);
        }
    }
    // timing: three epilogue modes, best of 3 after one warm-up launch each
    const char* names[6] = {"limbs", "limbs_and_montgomery_reduction", "mma_only", "mma_only_private_b_copies", "mma_only_no_b_loads", "mma_only_no_b_loads_kind_f8f6f4"};
    for (int v = 0; v < 6; v++) {
        const int mode = v < 3 ? v : 2, flags = v == 3 ? 1 : v == 4 ? 2 : v == 5 ? 6 : 0;
        Args g{d_a, n_a_tiles, d_b, tiles_per_cta, mode, flags, d_chk, nullptr, nullptr, d_err};
        k_matvec_umma<<<grid, NTHREADS, smem>>>(g);
        CHK(cudaDeviceSynchronize());
        float best = 1e30f;
        for (int rep = 0; rep < 3; rep++) {
            CHK(cudaEventRecord(e0)); k_matvec_umma<<<grid, NTHREADS, smem>>>(g); CHK(cudaEventRecord(e1)); CHK(cudaEventSynchronize(e1));
            float ms; CHK(cudaEventElapsedTime(&ms, e0, e1)); if (ms < best) best = ms;
        }
        int err = 0; CHK(cudaMemcpy(&err, d_err, 4, cudaMemcpyDeviceToHost));
        const double matvecs = (double)grid * tiles_per_cta * TILE_M;
        printf(", \"%s\": {\"ms\": %.4f, \"matvec_per_s\": %.4e, \"word_products_per_s\": %.4e, \"int8_mac_per_s\": %.4e, \"protocol_error\": %d}",
               names[v], best, matvecs / (best * 1e-3), matvecs * T * T / (best * 1e-3), matvecs * (double)KTOT * T * WCOLS / (best * 1e-3), err);
    }
    printf("}\n");
    return 0;
}
