// poseidon_tc.cuh — the Poseidon permutation of the arity-16 Merkle node hash (t = 17) with every linear layer on the INT8 tensor
// cores (tcgen05.mma kind::i8, accumulators in TMEM) and only the S-boxes and Montgomery reductions on the FP64 pipe.
//
// Why it is exact.  A state word is an integer < 2^256 (a lazily reduced representative in the 2^260-Montgomery form of fr52.cuh),
// written as 32 unsigned bytes.  For a constant matrix W (n_out x n_in field elements) the B operand holds REDUCED constants:
//      B[(i, c)][(j, a)] = byte_c( (W[i][j] * 2^(8a)) mod p ),   c, a = 0..31
// so that the int32 accumulator column (i, c) = sum_j sum_a byte_a(x_j) * byte_c(W_ij 2^(8a) mod p) and
//      y_i = sum_c column(i, c) * 2^(8c)  <  2^274,     y_i = sum_j W_ij x_j  (mod p)
// — 32 columns per output word (a plain Toeplitz expansion of the 512-bit product needs 63 and is half zeros: twice the MMA work,
// twice the TMEM reads).  A column is a sum of <= 25 * 32 byte products < 2^16, < 2^26: exact in int32.  The epilogue packs the 32
// columns into 52-bit limbs and folds the 20 bits above 2^254 back with p = 2^254 + c:  y = low + p - top * c  in [0, 2^255) — integer
// instructions only, no Montgomery reduction (the constants carry no Montgomery factor: the map is linear in the representatives).
// S-boxes stay Montgomery products of fr52.cuh, so the values agree mod p with poseidon/src/lib.rs:219-272 and the emitted
// digests are bit-identical (tests/test_gpu_poseidon_tc.py).
//
// Round structure (tools/tc_model.py is the big-integer model of all of this; checked against the oracle's permutation):
//   full rounds     x = sbox(v + rc);  v' = M x                 -> one stage per round: GEMM 17 -> 17 words, epilogue reduce + rc' + x^5
//   partial rounds  v' = Mh v + m0 sbox(v_0 + c_k),  Mh = M with column 0 zeroed, m0 = column 0 of M.  G rounds form a group:
//                   the S-box inputs of a group are  s_k = e0^T Mh^k v + sum_{m<k} gamma_{k-1-m} x_m + c_k   (gamma_d = e0^T Mh^d m0)
//                   so a group is: [chain kernel] G dependent S-boxes per permutation, each with a <= G-1 term correction, from the
//                   G "dots" e0^T Mh^k v;  [stage] GEMM (17 + G) -> (17 + G) words giving the next group's v and dots.
// Every GEMM input is an S-box output (< 2^254.3) or a stage output (< 2^255), so 32 bytes per word always suffice; the lazily
// reduced S-box inputs (< 2^260) never enter a GEMM.
//
// Data: "byte tiles" of 128 permutations, one 4 KB k-step per word in the UMMA no-swizzle K-major core-matrix layout
// ([16-byte half][8-row group][row][16 bytes]) — what the stage kernel's bulk copies bring into shared memory unchanged.
#pragma once
#include "poseidon_kernels.cuh"
#include "poseidon_coop_kernels.cuh"      // resolve_node

namespace smw {
namespace tc {

constexpr int TILE_M = 128;               // permutations per tile = TMEM lanes
constexpr int KSTEP_BYTES = TILE_M * 32;  // one word of every permutation of a tile
constexpr int WCOLS = 32;                 // accumulator columns per output word
constexpr int PAIR = 1;                   // one tile per CTA at a time (two tiles sharing every B block measured slower: profiles/r02u_tc_variants.txt)
#ifndef SMW_TC_SLOTS
#define SMW_TC_SLOTS 6
#endif
#ifndef SMW_TC_WPS
#define SMW_TC_WPS 1
#endif
constexpr int EPI_SLOTS = SMW_TC_SLOTS;   // word slots per TMEM lane quarter (one epilogue warp each)
constexpr int WPS = SMW_TC_WPS;           // words per slot and chunk: slot s owns words s, s + EPI_SLOTS, ... of every chunk
constexpr int CHUNK_WORDS = EPI_SLOTS * WPS;              // output words per MMA (N = 32 x that <= 256 columns)
constexpr int B_BLOCK_MAX = CHUNK_WORDS * WCOLS * 32;     // 8192 bytes: one k-step of one chunk
constexpr int MAX_KW = 25;                // k-steps (input words) per tile: 17 state words + a group of up to 8 S-box outputs
#ifndef SMW_TC_KSUB
#define SMW_TC_KSUB 9
#endif
// k-steps per ring stage: one bulk copy, one wait and one commit for KSUB MMAs.  The hand-over of a stage (mbarrier wait, fence, commit), not the L2
// stream or the bytes in flight, paces the stages without S-boxes: cfg2 per step 799.8 ms at KSUB = 2 (ten stages), 733.0 at 4 (four stages; five stages: 741.6),
// 731.1 at 5, 729.5 at 6, 725.5 at 9 (two stages of 54 KB: a chunk's 25 k-steps in three hand-overs) - profiles/r04m_tc_ring_variants.txt, r04n_tc_ring_stage_size.txt
constexpr int KSUB = SMW_TC_KSUB;
#ifdef SMW_TC_NSTAGE
constexpr int NSTAGE = SMW_TC_NSTAGE;
#else
constexpr int NSTAGE = 112 * 1024 / (KSUB * B_BLOCK_MAX);   // ring stages: what 227 KB leave beside the resident input words (KSUB = 9: two)
#endif
static_assert(NSTAGE >= 2, "the matrix ring needs two stages");
constexpr int EPI_WARPS = 4 * EPI_SLOTS;
constexpr int STAGE_THREADS = 64 + 32 * EPI_WARPS;
constexpr int ACC_COLS = CHUNK_WORDS * WCOLS;             // <= 256 TMEM columns per accumulator stage: [word][32]
static_assert(ACC_COLS <= 256 && ACC_COLS % 16 == 0 && STAGE_THREADS <= 1024, "stage shape");

struct StageSmem {
    alignas(1024) uint8_t a[PAIR][MAX_KW * KSTEP_BYTES];  // 100 KB per tile: its input words
    alignas(1024) uint8_t b[NSTAGE][KSUB * B_BLOCK_MAX];
    alignas(8) uint64_t b_full[NSTAGE], b_empty[NSTAGE], a_full, a_empty, acc_full[2], acc_empty[2];
    uint32_t tmem_base;
};

SMW_DEV uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
SMW_DEV void mbar_init(uint64_t* b, uint32_t count) { asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" :: "r"(smem_u32(b)), "r"(count)); }
SMW_DEV void mbar_expect_tx(uint64_t* b, uint32_t bytes) { asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" :: "r"(smem_u32(b)), "r"(bytes) : "memory"); }
SMW_DEV void mbar_arrive(uint64_t* b) { asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" :: "r"(smem_u32(b)) : "memory"); }
#ifndef SMW_TC_WAIT_HINT_NS
#define SMW_TC_WAIT_HINT_NS 20000u
#endif
// bounded wait: mbarrier.try_wait suspends the thread in hardware for up to the hinted time (the producer and MMA warps wait most of
// the time: without the hint their polling loop took ~12 % of the issue slots of two schedulers, profiles/r03b); ~2 s of wall time without the phase
// completing is a protocol error that ends the kernel with a flag instead of hanging the GPU
SMW_DEV bool mbar_wait(uint64_t* b, uint32_t parity) {
    const uint32_t addr = smem_u32(b);
    uint64_t t0 = 0;
#pragma unroll 1
    for (uint32_t spin = 0; ; spin++) {
        uint32_t ok;
        asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2, %3;\n\tselp.u32 %0, 1, 0, p;\n\t}" : "=r"(ok) : "r"(addr), "r"(parity), "r"(SMW_TC_WAIT_HINT_NS) : "memory");
        if (ok) return true;
        if ((spin & 63) == 63) {
            uint64_t now; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(now));
            if (t0 == 0) t0 = now; else if (now - t0 > 2000000000ull) return false;
        }
    }
}
#ifndef SMW_TC_PREFETCH
#define SMW_TC_PREFETCH 1
#endif
// asks L2 for a contiguous global range ahead of the bulk copy that will bring it into shared memory
SMW_DEV void bulk_prefetch_l2(const void* src, uint32_t bytes) {
    asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" :: "l"(src), "r"(bytes) : "memory");
}
SMW_DEV void bulk_g2s(void* dst, const void* src, uint32_t bytes, uint64_t* bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 :: "r"(smem_u32(dst)), "l"(src), "r"(bytes), "r"(smem_u32(bar)) : "memory");
}
// UMMA shared-memory descriptor, K-major, no swizzle, version 1 (sm_100): start >> 4, LBO (distance of the two 16-byte K halves) >> 4
// at bit 16, SBO (distance of consecutive 8-row groups) >> 4 at bit 32
SMW_DEV uint64_t umma_desc(uint32_t saddr, uint32_t lbo, uint32_t sbo) {
    return (uint64_t)((saddr & 0x3ffffu) >> 4) | ((uint64_t)(lbo >> 4) << 16) | ((uint64_t)(sbo >> 4) << 32) | (1ull << 46);
}
// kind::i8 instruction descriptor: D = s32 (bits 4-5 = 2), A and B unsigned 8-bit, K-major; N >> 3 at bit 17, M >> 4 at bit 24
SMW_DEV uint32_t umma_idesc(int m, int n) { return (2u << 4) | ((uint32_t)(n >> 3) << 17) | ((uint32_t)(m >> 4) << 24); }
SMW_DEV void umma_i8(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
    asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\ttcgen05.mma.cta_group::1.kind::i8 [%0], %1, %2, %3, p;\n\t}"
                 :: "r"(tmem_d), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate) : "memory");
}
SMW_DEV void umma_commit(uint64_t* bar) { asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" :: "r"(smem_u32(bar)) : "memory"); }
SMW_DEV bool elect_one() {
    uint32_t pred;
    asm volatile("{\n\t.reg .b32 rx;\n\t.reg .pred px;\n\telect.sync rx|px, 0xffffffff;\n\tselp.b32 %0, 1, 0, px;\n\t}" : "=r"(pred));
    return pred != 0;
}
SMW_DEV void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
SMW_DEV void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
SMW_DEV void tmem_ld16(uint32_t taddr, uint32_t* r) {
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
                   "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
                 : "r"(taddr));
}
SMW_DEV void tmem_ld_wait() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }

// p = 2^254 + c, c in 26-bit pieces (126 bits): top * c_k < 2^47 is one IMAD.WIDE
constexpr uint32_t PC26_0 = 0x0000001u, PC26_1 = 0x1bac840u, PC26_2 = 0x28dd8c4u, PC26_3 = 0x3f02652u, PC26_4 = 0x0224698u;
// 16 int32 columns k0 .. k0+15 (weights 2^(8k), each < 2^26) added into lazily carried 52-bit limb columns (each stays < 2^60).
// A column lands at bit sh = 8k mod 52 of limb 8k / 52: for sh < 32 that is ONE multiply-add on the FMA pipe (IMAD.WIDE.U32 c, 2^sh,
// acc: shift and 64-bit accumulation in one instruction — the ALU pipe is the busiest pipe of the epilogue); for sh >= 32 the part
// below bit 52 goes into the high word of the limb and the rest into the next limb.
template <int K0> SMW_DEV void columns_to_limbs16(const uint32_t* c, uint64_t* acc) {
#pragma unroll
    for (int i = 0; i < 16; i++) {
        const int k = K0 + i;
        const int bit = 8 * k, l = bit / 52, sh = bit % 52;
        if (sh < 32) acc[l] += (uint64_t)c[i] * (uint64_t)(1u << sh);
        else {
            const uint32_t lo = c[i] & ((1u << (52 - sh)) - 1u);                    // 52 - sh <= 20 bits, shifted to bit sh - 32 of the high word
            acc[l] += (uint64_t)(lo << (sh - 32)) << 32;
            acc[l + 1] += (uint64_t)(c[i] >> (52 - sh));
        }
    }
}
// six limb columns of y < 2^274 (+ the limbs of two addends a, b given as exact doubles < 2^52 each) -> limbs of
//      low254(y) + p - top(y) * c + a + b,    top = y >> 254 < 2^21
// carried to limbs < 2^52 (the top limb takes what is left).  Integer instructions only.
SMW_DEV F52 fold_top_add2(uint64_t* acc, const F52& a, const F52& b) {
#pragma unroll
    for (int i = 0; i < 5; i++) { acc[i + 1] += acc[i] >> 52; acc[i] &= M52; }
    const uint32_t top = (uint32_t)((acc[4] >> 46) | (acc[5] << 6));
    const uint64_t t0 = (uint64_t)top * PC26_0, t1 = (uint64_t)top * PC26_1, t2 = (uint64_t)top * PC26_2, t3 = (uint64_t)top * PC26_3, t4 = (uint64_t)top * PC26_4;
    int64_t r[5];
    r[0] = (int64_t)(acc[0] + P52_0) - (int64_t)(t0 + ((t1 & 0x3ffffffull) << 26));
    r[1] = (int64_t)(acc[1] + P52_1) - (int64_t)((t1 >> 26) + t2 + ((t3 & 0x3ffffffull) << 26));
    r[2] = (int64_t)(acc[2] + P52_2) - (int64_t)((t3 >> 26) + t4);
    r[3] = (int64_t)acc[3];
    r[4] = (int64_t)((acc[4] & ((1ull << 46) - 1ull)) + P52_4);
#pragma unroll
    for (int i = 0; i < 5; i++) r[i] += (int64_t)double_to_u52(a.v[i]) + (int64_t)double_to_u52(b.v[i]);
    F52 out;
#pragma unroll
    for (int i = 0; i < 4; i++) { r[i + 1] += r[i] >> 52; out.v[i] = u52_to_double((uint64_t)r[i] & M52); }      // arithmetic shift: borrows travel up
    out.v[4] = u52_to_double((uint64_t)r[4]);
    return out;
}

// byte-tile address of word `kstep` of row `row` (first 16-byte half; the second half is 2048 bytes further)
SMW_DEV size_t tile_word_offset(int kstep, int row) { return (size_t)kstep * KSTEP_BYTES + (size_t)(row >> 3) * 128 + (size_t)(row & 7) * 16; }
// value < 2^256 with limbs < 2^52 -> 32 little-endian bytes at the two halves of a tile word
SMW_DEV void store_word_bytes(uint8_t* tile, int kstep, int row, const F52& x) {
    uint64_t l[5];
#pragma unroll
    for (int i = 0; i < 5; i++) l[i] = double_to_u52(x.v[i]);
    const uint64_t x0 = l[0] | (l[1] << 52), x1 = (l[1] >> 12) | (l[2] << 40), x2 = (l[2] >> 24) | (l[3] << 28), x3 = (l[3] >> 36) | (l[4] << 16);
    uint8_t* q = tile + tile_word_offset(kstep, row);
    *reinterpret_cast<uint4*>(q) = make_uint4((uint32_t)x0, (uint32_t)(x0 >> 32), (uint32_t)x1, (uint32_t)(x1 >> 32));
    *reinterpret_cast<uint4*>(q + TILE_M * 16) = make_uint4((uint32_t)x2, (uint32_t)(x2 >> 32), (uint32_t)x3, (uint32_t)(x3 >> 32));
}
// x^5 (poseidon/src/lib.rs:24-29): two squarings with 15 limb products each and one general product
SMW_DEV F52 sbox52(const F52& s) {
    Acc52 A;
    const uint32_t exp_hi = g_exp52_hi;
    acc52_sqr_b(A, s); const F52 x2 = acc52_reduce_b(A, exp_hi);
    acc52_sqr_b(A, x2); const F52 x4 = acc52_reduce_b(A, exp_hi);
    acc52_mul_b(A, x4, s); return acc52_reduce_b(A, exp_hi);
}
// a canonical external-form element (x * 2^256 mod p, < p) -> limbs of 16 * that (= x * 2^260 mod p up to a multiple of p, < 2^259)
SMW_DEV F52 ext_times16(const uint32_t* src) {
    uint32_t w[8]; load_ext(src, w);
    const uint64_t x0 = (uint64_t)w[0] | ((uint64_t)w[1] << 32), x1 = (uint64_t)w[2] | ((uint64_t)w[3] << 32),
                   x2 = (uint64_t)w[4] | ((uint64_t)w[5] << 32), x3 = (uint64_t)w[6] | ((uint64_t)w[7] << 32);
    const uint64_t y0 = x0 << 4, y1 = (x1 << 4) | (x0 >> 60), y2 = (x2 << 4) | (x1 >> 60), y3 = (x3 << 4) | (x2 >> 60), y4 = x3 >> 60;
    uint64_t l[5];
    l[0] = y0 & M52; l[1] = ((y0 >> 52) | (y1 << 12)) & M52; l[2] = ((y1 >> 40) | (y2 << 24)) & M52; l[3] = ((y2 >> 28) | (y3 << 36)) & M52; l[4] = (y3 >> 16) | (y4 << 48);
    return f52_from_limbs(l);
}

// leaves of one or several FRI layers hashed by one launch: explicit runs (device array) or a single run `u`
struct LeafRuns { const LeafSeg* segs; int n_segs; LeafSeg u; uint64_t total; };
SMW_DEV bool resolve_leaf(const LeafRuns& R, uint64_t g, LeafSeg& ls, uint64_t& i) {
    if (g >= R.total) return false;
    if (R.segs) { ls = R.segs[find_seg(g, &R.segs[0].first, sizeof(LeafSeg) / 8, R.n_segs)]; i = g - ls.first; }
    else { ls = R.u; i = g; }
    return true;
}

// ------------------------------------------------------------------ the GEMM stage
enum { EPI_WORDS = 0, EPI_SBOX = 1, EPI_ABSORB_SBOX = 2, EPI_EMIT = 3, EPI_EMIT_LEAF = 4 };
struct StageArgs {
    const uint8_t* in; size_t in_tile_stride; int kw;            // input byte tiles: kw k-steps are read from each
    const uint8_t* bmat; int nw;                                 // Toeplitz blocks [chunk][k-step] of the stage's matrix, nw output words
    uint8_t* out; size_t out_tile_stride; int out_koff;          // output byte tiles: word w -> k-step out_koff + w
    int nbytes;                                                  // EPI_WORDS: words < nbytes leave as bytes, the others as dots
    double* dots; size_t dots_stride;                            // dots[(w - nbytes) * 5 + limb][perm], row stride dots_stride
    const uint32_t* tbl; int rc_idx;                             // tbl52 table and element index of the constant added to word 0 (rc_idx + w for word w); < 0: none
    uint64_t n_tiles;
    LevelBatch B; uint64_t node0;                                // EPI_ABSORB_SBOX / EPI_EMIT: the nodes of this batch are node0 .. of the level launch
    LeafRuns leaves;                                             // EPI_EMIT_LEAF: the items are leaves node0 .. of these runs (digest -> run.out + 32 i)
    int absorb_first;                                            // EPI_ABSORB_SBOX: sponge element index of lane 0 in this rate block (16 for the second block)
    int* err;
};

template <int MODE>
__global__ void __launch_bounds__(STAGE_THREADS, 1) k_tc_stage(const StageArgs g) {
    extern __shared__ __align__(1024) uint8_t smem_raw[];
    StageSmem& S = *reinterpret_cast<StageSmem*>(smem_raw);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int n_chunks = (g.nw + CHUNK_WORDS - 1) / CHUNK_WORDS;
    if (threadIdx.x == 0) {
        for (int i = 0; i < NSTAGE; i++) { mbar_init(&S.b_full[i], 1); mbar_init(&S.b_empty[i], 1); }
        mbar_init(&S.a_full, 1); mbar_init(&S.a_empty, 1);
        for (int i = 0; i < 2; i++) { mbar_init(&S.acc_full[i], 1); mbar_init(&S.acc_empty[i], EPI_WARPS); }
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
    const uint64_t n_pairs = (g.n_tiles + PAIR - 1) / PAIR;
    const uint64_t my_pairs = n_pairs > blockIdx.x ? (n_pairs - blockIdx.x + gridDim.x - 1) / gridDim.x : 0;

    if (warp == 0 && lane == 0) {
        // ---------------- producer: the input words of the pair's tiles, then the matrix blocks of every (chunk, k-step) through the ring
        uint32_t bit = 0;
        for (uint64_t it = 0; it < my_pairs; it++) {
            const uint64_t tile0 = (blockIdx.x + it * gridDim.x) * PAIR;
            const int nt = (int)min((uint64_t)PAIR, g.n_tiles - tile0);
            // The input words of a tile are resident for ALL of its chunks, so the copy of the next tile can only start when the last MMA of this
            // one has retired, and the tile comes from HBM (the batch's tiles are far larger than L2): ~2-3 of the ~14 us a group-stage tile takes
            // were exposed load time (profiles/r04k: tensor pipe 42 % active, L2 28 %).  The next tile is therefore requested into L2 one tile ahead.
            if (SMW_TC_PREFETCH && it + 1 < my_pairs)
                bulk_prefetch_l2(g.in + (blockIdx.x + (it + 1) * gridDim.x) * PAIR * g.in_tile_stride, (uint32_t)g.kw * KSTEP_BYTES);
            if (!mbar_wait(&S.a_empty, (uint32_t)(it & 1) ^ 1)) { atomicExch(g.err, 1); break; }
            mbar_expect_tx(&S.a_full, (uint32_t)(nt * g.kw) * KSTEP_BYTES);
            for (int tl = 0; tl < nt; tl++) bulk_g2s(S.a[tl], g.in + (tile0 + tl) * g.in_tile_stride, (uint32_t)g.kw * KSTEP_BYTES, &S.a_full);
            bool bad = false;
            for (int c = 0; c < n_chunks && !bad; c++) {
                const int words = min(CHUNK_WORDS, g.nw - c * CHUNK_WORDS);
                const uint32_t bytes = (uint32_t)words * WCOLS * 32u;
                const uint8_t* src = g.bmat + (size_t)c * g.kw * B_BLOCK_MAX;
                for (int s = 0; s < g.kw; s += KSUB, bit++) {
                    const int st = bit % NSTAGE, ks = min(KSUB, g.kw - s);
                    if (!mbar_wait(&S.b_empty[st], ((bit / NSTAGE) & 1) ^ 1)) { atomicExch(g.err, 2); bad = true; break; }
                    mbar_expect_tx(&S.b_full[st], bytes * ks);
                    bulk_g2s(S.b[st], src + (size_t)s * bytes, bytes * ks, &S.b_full[st]);
                }
            }
            if (bad) break;
        }
    } else if (warp == 1) {
        // ---------------- MMA issuer: the whole warp runs the loop converged (descriptors stay in uniform registers: no per-MMA
        // election loop, measured 2.4x on the issue rate), one elected lane issues.  One accumulator stage (256 TMEM columns:
        // [tile][word][64]) per chunk, double-buffered against the epilogue.
        uint32_t bit = 0, accit = 0;
        const bool leader = elect_one();
        const uint32_t hi = (128u >> 4) | (1u << 14), a_lbo = (uint32_t)((TILE_M * 16) >> 4) << 16;
        for (uint64_t it = 0; it < my_pairs; it++) {
            const uint64_t tile0 = (blockIdx.x + it * gridDim.x) * PAIR;
            const int nt = (int)min((uint64_t)PAIR, g.n_tiles - tile0);
            if (!mbar_wait(&S.a_full, (uint32_t)(it & 1))) { if (leader) atomicExch(g.err, 3); break; }
            tc_fence_after();
            bool bad = false;
            for (int c = 0; c < n_chunks && !bad; c++, accit++) {
                const int as = accit & 1, n = min(CHUNK_WORDS, g.nw - c * CHUNK_WORDS) * WCOLS;
                if (!mbar_wait(&S.acc_empty[as], ((accit >> 1) & 1) ^ 1)) { if (leader) atomicExch(g.err, 4); bad = true; break; }
                tc_fence_after();
                const uint32_t idesc = umma_idesc(TILE_M, n), b_lbo = (uint32_t)((n * 16) >> 4) << 16;
                for (int s = 0; s < g.kw; s += KSUB, bit++) {
                    const int st = bit % NSTAGE, ks = min(KSUB, g.kw - s);
                    if (!mbar_wait(&S.b_full[st], (bit / NSTAGE) & 1)) { if (leader) atomicExch(g.err, 5); bad = true; break; }
                    tc_fence_after();
                    const uint32_t b_lo = (smem_u32(S.b[st]) & 0x3ffffu) >> 4;
                    if (leader) {
#pragma unroll
                        for (int k = 0; k < KSUB; k++) {
                            if (k < ks) {
                                const uint64_t bd = (uint64_t)((b_lo + (uint32_t)k * (uint32_t)(n * 2)) | b_lbo) | ((uint64_t)hi << 32);
#pragma unroll
                                for (int tl = 0; tl < PAIR; tl++) {
                                    if (tl < nt) {
                                        const uint32_t a_lo = ((smem_u32(S.a[tl]) & 0x3ffffu) >> 4) + (uint32_t)(s + k) * (KSTEP_BYTES >> 4);
                                        umma_i8(tmem + (uint32_t)as * ACC_COLS + (uint32_t)tl * (CHUNK_WORDS * WCOLS), (uint64_t)(a_lo | a_lbo) | ((uint64_t)hi << 32), bd, idesc, (s + k) > 0 ? 1u : 0u);
                                    }
                                }
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
            if (leader) umma_commit(&S.a_empty);
            __syncwarp();
        }
    } else if (warp >= 2) {
        // ---------------- epilogue: warp (quarter q, slot) owns words slot, slot + EPI_SLOTS, .. of every chunk for the 32 permutations of its TMEM lane quarter
        const int q = warp & 3, slot = (warp - 2) >> 2, row = 32 * q + lane;
        uint32_t accit = 0; bool bad = false;
        for (uint64_t it = 0; it < my_pairs && !bad; it++) {
            const uint64_t tile = blockIdx.x + it * gridDim.x;
            const uint64_t perm = tile * TILE_M + row;
            uint8_t* out_tile = g.out + tile * g.out_tile_stride;
            NodeSeg ns; uint64_t nq = 0; bool valid = true;
            if (MODE == EPI_ABSORB_SBOX || MODE == EPI_EMIT) valid = resolve_node(g.B, g.node0 + perm, ns, nq);
            for (int c = 0; c < n_chunks; c++, accit++) {
                const int as = accit & 1;
                bool ok = true;
                if (lane == 0) ok = mbar_wait(&S.acc_full[as], (accit >> 1) & 1);      // one lane waits, the warp follows
                ok = __shfl_sync(0xffffffffu, ok ? 1 : 0, 0) != 0;
                if (!ok) { if (lane == 0) atomicExch(g.err, 6); bad = true; break; }
                tc_fence_after();
                const int w0 = c * CHUNK_WORDS + slot;
                const uint32_t taddr = tmem + ((uint32_t)(32 * q) << 16) + (uint32_t)as * ACC_COLS + (uint32_t)slot * WCOLS;
                uint64_t acc[WPS][6];
#pragma unroll
                for (int h = 0; h < WPS; h++) {
#pragma unroll
                    for (int i = 0; i < 6; i++) acc[h][i] = 0;
                    if (w0 + h * EPI_SLOTS < g.nw) {
                        uint32_t col[32];
                        tmem_ld16(taddr + h * EPI_SLOTS * WCOLS, col); tmem_ld16(taddr + h * EPI_SLOTS * WCOLS + 16, col + 16); tmem_ld_wait();
                        columns_to_limbs16<0>(col, acc[h]); columns_to_limbs16<16>(col + 16, acc[h]);
                    }
                }
                // the accumulator stage may be reused as soon as every warp has read its columns
                tc_fence_before();
                __syncwarp();
                if (lane == 0) mbar_arrive(&S.acc_empty[as]);
#pragma unroll
                for (int h = 0; h < WPS; h++) {
                    const int word = w0 + h * EPI_SLOTS;
                    if (word >= g.nw) break;
                    F52 rc = f52_zero(), add = f52_zero();
                    if (g.rc_idx >= 0) rc = tbl52_load(g.tbl, (uint32_t)(g.rc_idx + word));
                    if (MODE == EPI_ABSORB_SBOX && valid && word < 16) {
                        const int e = g.absorb_first + word - 4;                       // input index of this lane (4 DS fields precede the children)
                        if (e < (int)ns.arity) add = ext_times16(ns.children + (nq * ns.arity + (uint64_t)e) * 8);
                        else if (e == (int)ns.arity) add = f52_const(C_KONE52);        // the padding one: 2^260 mod p
                    }
                    const F52 r = fold_top_add2(acc[h], rc, add);
                    if (MODE == EPI_WORDS) {
                        if (word < g.nbytes) store_word_bytes(out_tile, g.out_koff + word, row, r);
                        else {
                            double* d = g.dots + (size_t)(word - g.nbytes) * 5 * g.dots_stride + perm;
#pragma unroll
                            for (int i = 0; i < 5; i++) d[(size_t)i * g.dots_stride] = r.v[i];
                        }
                    } else if (MODE == EPI_SBOX || MODE == EPI_ABSORB_SBOX) {
                        store_word_bytes(out_tile, g.out_koff + word, row, sbox52(r));
                    } else if (MODE == EPI_EMIT) {
                        if (valid && word == 0) {
                            Acc52 Q; acc52_mul(Q, r, f52_const(C_KOUT52));
                            emit_ext52(acc52_reduce_add2(Q, 1, f52_zero(), f52_zero()), ns.parents + nq * 8);
                        }
                    } else if (MODE == EPI_EMIT_LEAF) {
                        LeafSeg ls; uint64_t li = 0;
                        if (word == 0 && resolve_leaf(g.leaves, g.node0 + perm, ls, li)) {
                            Acc52 Q; acc52_mul(Q, r, f52_const(C_KOUT52));
                            emit_ext52(acc52_reduce_add2(Q, 1, f52_zero(), f52_zero()), ls.out + li * 8);
                        }
                    }
                }
            }
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 1) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" :: "r"(tmem), "r"(512u) : "memory");
}

// ------------------------------------------------------------------ head: the first rate block of the node sponge, ARK and S-boxes of round 0
// lane l of node q = element l of [arity, level, q + pos_offset, label, child_0 ..] (merkle/src/lib.rs:164-179, poseidon/src/lib.rs:288-312);
// x_l = (lane + rc_full[0][l])^5 goes to k-step l of the node's byte tile.  blockDim = (128 rows, 4): threadIdx.y strides the lanes.
struct HeadArgs { LevelBatch B; uint64_t node0, n_nodes; uint8_t* out; size_t out_tile_stride; const uint32_t* tbl; int rc_idx; };
template <int T> __global__ void __launch_bounds__(512) k_tc_head(const HeadArgs g) {
    const int row = threadIdx.x;
    const uint64_t tile = blockIdx.x, perm = tile * TILE_M + row;
    NodeSeg ns; uint64_t q = 0;
    const bool valid = perm < g.n_nodes && resolve_node(g.B, g.node0 + perm, ns, q);
    uint8_t* out_tile = g.out + tile * g.out_tile_stride;
    const F52 k_rr = f52_const(C_KRR52);
#pragma unroll 1
    for (int l = threadIdx.y; l < T; l += blockDim.y) {
        const F52 rc = tbl52_load(g.tbl, (uint32_t)(g.rc_idx + l));
        F52 s = rc;
        if (valid && l < T - 1) {
            Acc52 A;
            if (l < 4) {                       // a DS integer d -> d * 2^520 / 2^260
                acc52_mul(A, f52_from_u64(l == 0 ? (uint64_t)ns.arity : l == 1 ? (uint64_t)ns.level : l == 2 ? q + ns.pos_offset : ns.label), k_rr);
                s = acc52_reduce_add2(A, 1, rc, f52_zero());
            } else {
                const int e = l - 4;
                F52 add = f52_zero();
                if (e < (int)ns.arity) add = ext_times16(ns.children + (q * ns.arity + (uint64_t)e) * 8);
                else if (e == (int)ns.arity) add = f52_const(C_KONE52);
                acc52_zero(A);
                s = acc52_reduce_add2(A, 0, rc, add);          // rc + add with the carries normalised (limbs < 2^52)
            }
        }
        store_word_bytes(out_tile, l, row, sbox52(s));
    }
}

// head of a one-permutation leaf hash.  kind 0: FRI hashed leaf (deep_ali/src/fri.rs:38-44): a fixed 17-lane pre-image template (tags and
// label words as 256-bit little-endian integers) with f_i and s_{i / cp_div} dropped into lanes lane_f, lane_s.  kind 1: pair leaf
// (merkle/src/lib.rs:382-400): [arity, u32::MAX, i + pos_offset, label, f_i, cp_{i / cp_div}, 1, 0 ..].  x_l = (lane + rc_full[0][l])^5.
struct HeadLeafArgs { LeafRuns R; uint64_t item0, n_items; uint8_t* out; size_t out_tile_stride; const uint32_t* tbl; int rc_idx;
                      int kind; const uint32_t* tmpl; int lane_f, lane_s; };
template <int T> __global__ void __launch_bounds__(512) k_tc_head_leaf(const HeadLeafArgs g) {
    const int row = threadIdx.x;
    const uint64_t tile = blockIdx.x, perm = tile * TILE_M + row;
    LeafSeg ls; uint64_t i = 0;
    const bool valid = perm < g.n_items && resolve_leaf(g.R, g.item0 + perm, ls, i);
    uint8_t* out_tile = g.out + tile * g.out_tile_stride;
    const F52 k_rr = f52_const(C_KRR52);
    const int lane_f = g.kind == 0 ? g.lane_f : 4, lane_s = g.kind == 0 ? g.lane_s : 5;
#pragma unroll 1
    for (int l = threadIdx.y; l < T; l += blockDim.y) {
        const F52 rc = tbl52_load(g.tbl, (uint32_t)(g.rc_idx + l));
        F52 s = rc;
        if (valid) {
            Acc52 A;
            if (l == lane_f || l == lane_s) {
                F52 add = f52_zero();
                if (l == lane_f) add = ext_times16(ls.f + i * 8);
                else if (ls.cp) add = ext_times16(ls.cp + (i / ls.cp_div) * 8);
                acc52_zero(A);
                s = acc52_reduce_add2(A, 0, rc, add);
            } else if (g.kind == 0) {                  // a template lane: integer * 2^520 / 2^260
                acc52_mul(A, raw_ext52(g.tmpl + (size_t)l * 8), k_rr);
                s = acc52_reduce_add2(A, 1, rc, f52_zero());
            } else if (l < 4) {
                acc52_mul(A, f52_from_u64(l == 0 ? (uint64_t)ls.arity : l == 1 ? 0xffffffffull : l == 2 ? i + ls.pos_offset : ls.label), k_rr);
                s = acc52_reduce_add2(A, 1, rc, f52_zero());
            } else if (l == 6) {
                acc52_zero(A);
                s = acc52_reduce_add2(A, 0, rc, f52_const(C_KONE52));
            }
        }
        store_word_bytes(out_tile, l, row, sbox52(s));
    }
}

// ------------------------------------------------------------------ chain: the G dependent S-boxes of one group of partial rounds
// s_k = dot_k + sum_{m<k} gamma[k-1-m] x_m + rc_part[round0 + k],  x_k = s_k^5;  x_k -> k-step koff + k of the permutation's tile.
// One thread per permutation; the group's x values stay in shared memory ([k][limb][thread], conflict-free).
#ifndef SMW_TC_CHAIN_THREADS
#define SMW_TC_CHAIN_THREADS 128
#endif
#ifndef SMW_TC_CHAIN_MINB
#define SMW_TC_CHAIN_MINB 6      // 78 registers instead of 90, no spills: cfg2 755.4 -> 748.2 ms (profiles/r04h_tc_streams.txt)
#endif
constexpr int CHAIN_THREADS = SMW_TC_CHAIN_THREADS, MAX_G = 8;
struct ChainArgs { const double* dots; size_t dots_stride; uint64_t n_perm; uint8_t* out; size_t out_tile_stride; int koff; int G;
                   const double* gamma; const uint32_t* tbl; int rc_idx; };
__global__ void __launch_bounds__(CHAIN_THREADS, SMW_TC_CHAIN_MINB) k_tc_chain(const ChainArgs g) {
    __shared__ double xs[MAX_G * 5 * CHAIN_THREADS];
    const uint64_t perm = (uint64_t)blockIdx.x * CHAIN_THREADS + threadIdx.x;
    if (perm >= g.n_perm) return;
    const uint64_t tile = perm / TILE_M; const int row = (int)(perm % TILE_M);
    uint8_t* out_tile = g.out + tile * g.out_tile_stride;
#pragma unroll 1
    for (int k = 0; k < g.G; k++) {
        F52 d;
#pragma unroll
        for (int i = 0; i < 5; i++) d.v[i] = g.dots[((size_t)k * 5 + i) * g.dots_stride + perm];
        const F52 rc = tbl52_load(g.tbl, (uint32_t)(g.rc_idx + k));
        Acc52 A; acc52_zero(A);
#pragma unroll 1
        for (int m = 0; m < k; m++) {
            F52 x, gm;
#pragma unroll
            for (int i = 0; i < 5; i++) { x.v[i] = xs[(m * 5 + i) * CHAIN_THREADS + threadIdx.x]; gm.v[i] = __ldg(g.gamma + (size_t)(k - 1 - m) * 5 + i); }
            acc52_mac(A, x, gm);
        }
        const F52 x = sbox52(acc52_reduce_add2(A, (uint32_t)k, d, rc));
#pragma unroll
        for (int i = 0; i < 5; i++) xs[(k * 5 + i) * CHAIN_THREADS + threadIdx.x] = x.v[i];
        store_word_bytes(out_tile, g.koff + k, row, x);
    }
}

} // namespace tc
} // namespace smw
