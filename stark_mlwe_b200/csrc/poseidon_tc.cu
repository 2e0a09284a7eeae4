// poseidon_tc.cu — host side of the tensor-core node hash (poseidon_tc.cuh): the word-level matrices of every stage, derived on
// the host from the parameter set's external-form tables (small modular arithmetic, once per parameter set), their Toeplitz
// byte expansion in the UMMA block layout, and the launch sequence of one Merkle level.
#include "poseidon_tc.cuh"
#include <algorithm>
#include <memory>
#include <cstdlib>
#include <cstring>
#include <vector>

namespace smw {
namespace tc {

// ------------------------------------------------------------------ host arithmetic mod p (Pallas Fr), Montgomery radix 2^256 = the external form
namespace {
struct U256 { uint64_t v[4]; };
const U256 PMOD = {{0x8c46eb2100000001ull, 0x224698fc0994a8ddull, 0x0ull, 0x4000000000000000ull}};
const uint64_t PINV64 = 0x8c46eb20ffffffffull;          // -p^-1 mod 2^64
bool geq(const U256& a, const U256& b) { for (int i = 3; i >= 0; i--) if (a.v[i] != b.v[i]) return a.v[i] > b.v[i]; return true; }
U256 sub_raw(const U256& a, const U256& b) { U256 r; unsigned __int128 br = 0; for (int i = 0; i < 4; i++) { unsigned __int128 d = (unsigned __int128)a.v[i] - b.v[i] - (uint64_t)br; r.v[i] = (uint64_t)d; br = (d >> 64) & 1; } return r; }
U256 addm(const U256& a, const U256& b) {
    U256 r; unsigned __int128 c = 0;
    for (int i = 0; i < 4; i++) { c += (unsigned __int128)a.v[i] + b.v[i]; r.v[i] = (uint64_t)c; c >>= 64; }
    if (c || geq(r, PMOD)) r = sub_raw(r, PMOD);
    return r;
}
// a * b / 2^256 mod p, canonical (CIOS over 64-bit digits)
U256 mont_mul(const U256& a, const U256& b) {
    uint64_t t[6] = {0, 0, 0, 0, 0, 0};
    for (int i = 0; i < 4; i++) {
        unsigned __int128 c = 0;
        for (int j = 0; j < 4; j++) { c += (unsigned __int128)a.v[j] * b.v[i] + t[j]; t[j] = (uint64_t)c; c >>= 64; }
        c += t[4]; t[4] = (uint64_t)c; t[5] = (uint64_t)(c >> 64);
        const uint64_t m = t[0] * PINV64;
        c = (unsigned __int128)m * PMOD.v[0] + t[0]; c >>= 64;
        for (int j = 1; j < 4; j++) { c += (unsigned __int128)m * PMOD.v[j] + t[j]; t[j - 1] = (uint64_t)c; c >>= 64; }
        c += t[4]; t[3] = (uint64_t)c; t[4] = t[5] + (uint64_t)(c >> 64);
    }
    U256 r = {{t[0], t[1], t[2], t[3]}};
    if (t[4] || geq(r, PMOD)) r = sub_raw(r, PMOD);
    return r;
}
U256 pow2_mod(int e) { U256 x = {{1, 0, 0, 0}}; for (int i = 0; i < e; i++) x = addm(x, x); return x; }
const U256 ZERO = {{0, 0, 0, 0}};
typedef std::vector<std::vector<U256>> Mat;           // entries in the external (2^256 Montgomery) form
Mat mat(int r, int c) { return Mat(r, std::vector<U256>(c, ZERO)); }
Mat matmul(const Mat& A, const Mat& B) {
    Mat C = mat((int)A.size(), (int)B[0].size());
    for (size_t i = 0; i < A.size(); i++) for (size_t j = 0; j < B[0].size(); j++) { U256 s = ZERO; for (size_t k = 0; k < B.size(); k++) s = addm(s, mont_mul(A[i][k], B[k][j])); C[i][j] = s; }
    return C;
}

size_t core_pos(int rows, int row, int k) { return (size_t)(k / 16) * rows * 16 + (size_t)(row / 8) * 128 + (size_t)(row % 8) * 16 + (size_t)(k % 16); }
// Byte blocks [chunk][k-step] of W (nw x kw field elements, external form): row (w, c) of k-step j holds byte_c((W[w][j] * 2^(8a)) mod p)
// at k = a — the reduced constants of poseidon_tc.cuh, in the UMMA K-major core-matrix layout
std::vector<uint8_t> reduced_blocks(const Mat& W) {
    const int nw = (int)W.size(), kw = (int)W[0].size(), n_chunks = (nw + CHUNK_WORDS - 1) / CHUNK_WORDS;
    std::vector<uint8_t> out((size_t)n_chunks * kw * B_BLOCK_MAX, 0);
    U256 shift[32];
    for (int a = 0; a < 32; a++) { shift[a] = ZERO; shift[a].v[a / 8] = 1ull << (8 * (a % 8)); }       // the plain integer 2^(8a) < p
    for (int c = 0; c < n_chunks; c++) {
        const int words = std::min(CHUNK_WORDS, nw - c * CHUNK_WORDS), n_rows = words * WCOLS;
        for (int j = 0; j < kw; j++) {
            uint8_t* blk = out.data() + (size_t)c * kw * B_BLOCK_MAX + (size_t)j * n_rows * 32;
            for (int w = 0; w < words; w++)
                for (int a = 0; a < 32; a++) {
                    const U256 e = mont_mul(W[c * CHUNK_WORDS + w][j], shift[a]);      // (W 2^256) (2^(8a)) / 2^256 = W 2^(8a) mod p, canonical
                    uint8_t byte[32]; memcpy(byte, e.v, 32);
                    for (int col = 0; col < WCOLS; col++) blk[core_pos(n_rows, w * WCOLS + col, a)] = byte[col];
                }
        }
    }
    return out;
}
} // namespace

struct Tables {
    int G = 0;
    uint8_t *b_full = nullptr, *b_r3 = nullptr, *b_grp = nullptr, *b_last = nullptr, *b_final = nullptr;
    double* gamma = nullptr; int* d_err = nullptr;
    uint8_t *buf_p = nullptr, *buf_q = nullptr; double* dots = nullptr; uint64_t cap_nodes = 0;
    int n_sm = 0;
};
void free_tables(void* q) {
    Tables* t = (Tables*)q; if (!t) return;
    cudaFree(t->b_full); cudaFree(t->b_r3); cudaFree(t->b_grp); cudaFree(t->b_last); cudaFree(t->b_final); cudaFree(t->gamma); cudaFree(t->d_err);
    cudaFree(t->buf_p); cudaFree(t->buf_q); cudaFree(t->dots);
    delete t;
}
static int upload(smw_ctx* c, const std::vector<uint8_t>& h, uint8_t** d) {
    SMW_CUDA(c, cudaMalloc(d, h.size()));
    SMW_CUDA(c, cudaMemcpyAsync(*d, h.data(), h.size(), cudaMemcpyHostToDevice, c->stream));
    SMW_CUDA(c, cudaStreamSynchronize(c->stream));
    return SMW_OK;
}
static int build_tables(smw_ctx* c, smw_params* p, int G, Tables** out) {
    const int t = p->t, rp = p->rp;
    if ((t != 17 && t != 9) || G < 1 || G > MAX_G || rp % G || t + G > MAX_KW) return set_err(c, SMW_ERR_INVALID_ARG, "tensor-core hash: t = 9 or 17 and a group size dividing rp only");
    std::vector<uint64_t> ext((size_t)t * t * 4);
    SMW_CUDA(c, cudaMemcpyAsync(ext.data(), p->d_ext, ext.size() * 8, cudaMemcpyDeviceToHost, c->stream));
    SMW_CUDA(c, cudaStreamSynchronize(c->stream));
    Mat M = mat(t, t), Mh = mat(t, t);
    for (int i = 0; i < t; i++) for (int j = 0; j < t; j++) { memcpy(M[i][j].v, &ext[((size_t)i * t + j) * 4], 32); Mh[i][j] = j == 0 ? ZERO : M[i][j]; }
    const U256 one = pow2_mod(256), c260 = pow2_mod(260);
    // rows[d] = e0^T Mh^d, cols[d] = Mh^d m0, d < G;  MhG = Mh^G
    std::vector<std::vector<U256>> rows(G, std::vector<U256>(t, ZERO)), cols(G, std::vector<U256>(t, ZERO));
    rows[0][0] = one;
    for (int i = 0; i < t; i++) cols[0][i] = M[i][0];
    for (int d = 1; d < G; d++)
        for (int j = 0; j < t; j++) {
            U256 s = ZERO, u = ZERO;
            for (int i = 0; i < t; i++) { s = addm(s, mont_mul(rows[d - 1][i], Mh[i][j])); u = addm(u, mont_mul(Mh[j][i], cols[d - 1][i])); }
            rows[d][j] = s; cols[d][j] = u;
        }
    Mat MhG = Mh; for (int d = 1; d < G; d++) MhG = matmul(Mh, MhG);
    Mat Wv = mat(t, t + G);
    for (int i = 0; i < t; i++) { for (int j = 0; j < t; j++) Wv[i][j] = MhG[i][j]; for (int m = 0; m < G; m++) Wv[i][t + m] = cols[G - 1 - m][i]; }
    Mat R = mat(G, t); for (int k = 0; k < G; k++) R[k] = rows[k];
    Mat Wd = matmul(R, Wv), D0 = matmul(R, M);
    Mat Br3 = M; for (int k = 0; k < G; k++) Br3.push_back(D0[k]);
    Mat Bgrp = Wv; for (int k = 0; k < G; k++) Bgrp.push_back(Wd[k]);
    Mat Bfin(1, M[0]);
    Tables* T = new Tables; T->G = G;
    int rc = upload(c, reduced_blocks(M), &T->b_full);
    if (!rc) rc = upload(c, reduced_blocks(Br3), &T->b_r3);
    if (!rc) rc = upload(c, reduced_blocks(Bgrp), &T->b_grp);
    if (!rc) rc = upload(c, reduced_blocks(Wv), &T->b_last);
    if (!rc) rc = upload(c, reduced_blocks(Bfin), &T->b_final);
    if (!rc) {
        std::vector<double> gm((size_t)G * 5);
        for (int d = 0; d < G; d++) {
            const U256 e = mont_mul(cols[d][0], c260);
            const uint64_t l[5] = {e.v[0] & M52, ((e.v[0] >> 52) | (e.v[1] << 12)) & M52, ((e.v[1] >> 40) | (e.v[2] << 24)) & M52, ((e.v[2] >> 28) | (e.v[3] << 36)) & M52, e.v[3] >> 16};
            for (int i = 0; i < 5; i++) gm[(size_t)d * 5 + i] = (double)l[i];
        }
        std::vector<uint8_t> raw((uint8_t*)gm.data(), (uint8_t*)gm.data() + gm.size() * 8);
        uint8_t* q = nullptr; rc = upload(c, raw, &q); T->gamma = (double*)q;
    }
    if (!rc && cudaMalloc(&T->d_err, 4) != cudaSuccess) rc = set_err(c, SMW_ERR_CUDA, "cudaMalloc");
    if (!rc) { cudaMemsetAsync(T->d_err, 0, 4, c->stream); cudaDeviceGetAttribute(&T->n_sm, cudaDevAttrMultiProcessorCount, c->device); }
    if (rc) { free_tables(T); return rc; }
    *out = T;
    return SMW_OK;
}

// ------------------------------------------------------------------ one Merkle level
// partial rounds per group: t = 17 (rp = 64): 8; t = 9 (rp = 60): 6, so that a group stage has 15 input and output words
static int group_size(int t) {
    static int g17 = -1, g9 = -1;
    if (g17 < 0) { const char* e = getenv("SMW_TC_GROUP"); g17 = e ? atoi(e) : 8; if (g17 < 1 || g17 > MAX_G) g17 = 8; }
    if (g9 < 0) { const char* e = getenv("SMW_TC_GROUP9"); g9 = e ? atoi(e) : 6; if (g9 < 1 || g9 > MAX_G) g9 = 6; }
    return t == 9 ? g9 : g17;
}
static uint64_t batch_nodes() { static long v = -1; if (v < 0) { const char* e = getenv("SMW_TC_BATCH"); v = e ? atol(e) : (1L << 20); if (v < TILE_M) v = 1L << 20; } return (uint64_t)v; }

template <int MODE> static int launch_stage(smw_ctx* c, const Tables* T, StageArgs a, int kclass) {
    const size_t sm = sizeof(StageSmem) + 1024;
    static bool attr_set = false;
    if (!attr_set) { SMW_CUDA(c, cudaFuncSetAttribute(k_tc_stage<MODE>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm)); attr_set = true; }
    a.err = T->d_err;
    const unsigned grid = (unsigned)std::min<uint64_t>(a.n_tiles, (uint64_t)T->n_sm);
    { LaunchScope ls(c, kclass); k_tc_stage<MODE><<<grid, STAGE_THREADS, sm, c->stream>>>(a); }
    SMW_CUDA(c, cudaGetLastError());
    return SMW_OK;
}

// What one arity-16 node (two permutations) costs on this path, counted from the launch sequence below:
// out[0] = FP64-pipe instructions (DFMA.RZ + DADD: 75 per general limb product block, 45 per squaring, 85 per Montgomery
// reduction, 15 per stage output (limb conversions of the fold), 5 per byte packing), out[1] = int8 multiply-accumulates the tensor
// cores execute, out[2] = word products (row x word, 32 x 32 byte MACs each), out[3] = kernel launches per batch, out[4] = group size.
void node_counts(int t, int rp, uint64_t out[5]) {
    const uint64_t G = (uint64_t)group_size(t), NG = (uint64_t)rp / G, T = (uint64_t)t;
    const int n_perms = t == 17 ? 2 : 1;              // an arity-16 node absorbs two rate blocks; a t = 9 node or any pair / hashed leaf one
    const uint64_t RED = 85, MUL = 75, SQR = 45, PACK = 5, FOLD = 15, SBOX = 2 * (SQR + RED) + (MUL + RED);
    uint64_t fp = 0, wp = 0, launches = 1;
    fp += T * (SBOX + RED + PACK) + 4 * MUL;                                          // head: 17 lanes (4 DS products), ARK, S-box
    for (int perm = 0; perm < n_perms; perm++) {
        fp += 6 * T * (FOLD + SBOX + PACK); wp += 6 * T * T; launches += 6;           // full rounds 1-3, 5-7: stage + S-box epilogue
        fp += (T + G) * FOLD + T * PACK; wp += (T + G) * T; launches += 1;            // round 3 -> state and first dots
        fp += NG * (G * (RED + SBOX + PACK) + (G * (G - 1) / 2) * MUL); launches += NG;            // chains
        fp += (NG - 1) * ((T + G) * FOLD + T * PACK); wp += (NG - 1) * (T + G) * (T + G); launches += NG - 1;   // group stages
        fp += T * (FOLD + SBOX + PACK); wp += T * (T + G); launches += 1;             // last group stage -> round 4 S-boxes
        if (perm + 1 < n_perms) { fp += T * (FOLD + SBOX + PACK); wp += T * T; }      // round 7 + second rate block + round 0 S-boxes
        else { fp += FOLD + MUL + RED; wp += T; }                                     // round 7, lane 0 only, external form
        launches += 1;
    }
    out[0] = fp; out[1] = wp * 32 * WCOLS; out[2] = wp; out[3] = launches; out[4] = G;
}

// What a run of the pipeline hashes: Merkle nodes of full arity 16 (two permutations: the sponge absorbs 4 DS fields + 16 children + the
// padding one in two rate blocks) or one-permutation leaves (FRI hashed leaves / pair leaves).
struct Job {
    int kind;                      // 0: nodes of B (t = 17: full arity 16, two permutations; t = 9: full arity 2, one); 1: FRI hashed leaves (template); 2: pair leaves
    LevelBatch B; LeafRuns R; const uint32_t* tmpl; int lane_f, lane_s;
    uint64_t total; int kclass;
};
static int ensure(smw_ctx* c, smw_params* p, Tables** out) {
    if (!p->tc) { Tables* T = nullptr; int rc = build_tables(c, p, group_size(p->t), &T); if (rc) return rc; p->tc = T; p->tc_free = free_tables; }
    *out = (Tables*)p->tc;
    return SMW_OK;
}
// Two halves of a batch on two streams (the default; SMW_TC_STREAMS=1 keeps one): each half runs the whole launch sequence in order on its own stream, so a chain
// kernel of one half (FP64 / ALU pipes, small CTAs) can share the SMs with a group stage of the other half (tensor pipe, paced by the
// matrix stream from L2) whenever the two meet, and the tail of every persistent stage kernel is covered by the other half.  The halves use
// disjoint tile ranges of the same work buffers.  Measured on cfg2 (profiles/r04g, r04h): 755.4 -> 743.9 ms, with the chain kernel held to
// 80 registers (six CTAs per SM) 748.2 -> 725.9 ms; a three-stage ring that would leave shared memory for a resident chain CTA: slower.
static int n_streams() { static int v = -1; if (v < 0) { const char* e = getenv("SMW_TC_STREAMS"); v = e ? atoi(e) : 2; if (v < 1 || v > 2) v = 2; } return v; }
struct SideStream {
    smw_ctx* c; cudaStream_t main; bool used = false;
    explicit SideStream(smw_ctx* ctx) : c(ctx), main(ctx->stream) {
        if (c->side.empty()) {
            cudaStream_t s; cudaStreamCreateWithFlags(&s, cudaStreamNonBlocking); c->side.push_back(s);
            cudaEvent_t e; cudaEventCreateWithFlags(&e, cudaEventDisableTiming); c->side_ev.push_back(e);
        }
        if (!c->fan_ev) cudaEventCreateWithFlags(&c->fan_ev, cudaEventDisableTiming);
    }
    void enter() { cudaEventRecord(c->fan_ev, main); cudaStreamWaitEvent(c->side[0], c->fan_ev, 0); c->stream = c->side[0]; used = true; }
    void leave() { c->stream = main; cudaEventRecord(c->side_ev[0], c->side[0]); cudaStreamWaitEvent(main, c->side_ev[0], 0); }
    ~SideStream() { c->stream = main; }
};

// the launch sequence of `n` items (first one: item `node0` of the job) whose byte tiles start at P / Q and whose dots start at `dots`
static int run_sequence(smw_ctx* c, smw_params* p, Tables* T, const Job& J, uint64_t node0, uint64_t n, uint8_t* P, uint8_t* Q, double* dots) {
    const int t = p->t, G = T->G, NG = p->rp / G, n_perms = (J.kind == 0 && t == 17) ? 2 : 1;
    const TableLayout& L = p->layout;
    const size_t tile_stride = (size_t)MAX_KW * KSTEP_BYTES;
    const uint64_t n_tiles = (n + TILE_M - 1) / TILE_M;
    StageArgs base{}; base.in_tile_stride = tile_stride; base.out_tile_stride = tile_stride; base.tbl = p->d_tables52; base.rc_idx = -1; base.n_tiles = n_tiles;
    base.B = J.B; base.leaves = J.R; base.node0 = node0; base.dots = dots; base.dots_stride = T->cap_nodes; base.nbytes = t;
    auto full_round = [&](int round_next) -> int {          // P: x of a full round -> Q: x of the next one
        StageArgs a = base; a.in = P; a.kw = t; a.bmat = T->b_full; a.nw = t; a.out = Q; a.rc_idx = (int)(L.rc_full + (uint32_t)round_next * t);
        int rc = launch_stage<EPI_SBOX>(c, T, a, J.kclass); std::swap(P, Q); return rc;
    };
    if (J.kind == 0) {
        HeadArgs h{J.B, node0, n, P, tile_stride, p->d_tables52, (int)L.rc_full};
        LaunchScope ls(c, J.kclass);
        if (t == 17) k_tc_head<17><<<(unsigned)n_tiles, dim3(TILE_M, 4), 0, c->stream>>>(h);
        else k_tc_head<9><<<(unsigned)n_tiles, dim3(TILE_M, 3), 0, c->stream>>>(h);
    } else {
        HeadLeafArgs h{J.R, node0, n, P, tile_stride, p->d_tables52, (int)L.rc_full, J.kind == 1 ? 0 : 1, J.tmpl, J.lane_f, J.lane_s};
        LaunchScope ls(c, J.kclass);
        if (t == 17) k_tc_head_leaf<17><<<(unsigned)n_tiles, dim3(TILE_M, 4), 0, c->stream>>>(h);
        else k_tc_head_leaf<9><<<(unsigned)n_tiles, dim3(TILE_M, 3), 0, c->stream>>>(h);
    }
    SMW_CUDA(c, cudaGetLastError());
    for (int perm = 0; perm < n_perms; perm++) {
        int rc = 0;
        for (int r = 1; r <= 3 && !rc; r++) rc = full_round(r);
        if (rc) return rc;
        { StageArgs a = base; a.in = P; a.kw = t; a.bmat = T->b_r3; a.nw = t + G; a.out = Q; rc = launch_stage<EPI_WORDS>(c, T, a, J.kclass); std::swap(P, Q); if (rc) return rc; }
        for (int g = 0; g < NG; g++) {
            { ChainArgs ch{dots, (size_t)T->cap_nodes, n, P, tile_stride, t, G, T->gamma, p->d_tables52, (int)(L.rc_part + (uint32_t)g * G)};
              LaunchScope ls(c, J.kclass); k_tc_chain<<<(unsigned)((n + CHAIN_THREADS - 1) / CHAIN_THREADS), CHAIN_THREADS, 0, c->stream>>>(ch); }
            SMW_CUDA(c, cudaGetLastError());
            StageArgs a = base; a.in = P; a.kw = t + G; a.out = Q;
            if (g + 1 < NG) { a.bmat = T->b_grp; a.nw = t + G; rc = launch_stage<EPI_WORDS>(c, T, a, J.kclass); }
            else { a.bmat = T->b_last; a.nw = t; a.rc_idx = (int)(L.rc_full + 4u * t); rc = launch_stage<EPI_SBOX>(c, T, a, J.kclass); }
            std::swap(P, Q);
            if (rc) return rc;
        }
        for (int r = 5; r <= 7 && !rc; r++) rc = full_round(r);
        if (rc) return rc;
        StageArgs a = base; a.in = P; a.kw = t; a.out = Q;
        if (perm + 1 < n_perms) { a.bmat = T->b_full; a.nw = t; a.rc_idx = (int)L.rc_full; a.absorb_first = t - 1; rc = launch_stage<EPI_ABSORB_SBOX>(c, T, a, J.kclass); std::swap(P, Q); }
        else if (J.kind == 0) { a.bmat = T->b_final; a.nw = 1; rc = launch_stage<EPI_EMIT>(c, T, a, J.kclass); }
        else { a.bmat = T->b_final; a.nw = 1; rc = launch_stage<EPI_EMIT_LEAF>(c, T, a, J.kclass); }
        if (rc) return rc;
    }
    return SMW_OK;
}
static int run_job(smw_ctx* c, smw_params* p, const Job& J) {
    Tables* T; int rc0 = ensure(c, p, &T); if (rc0) return rc0;
    const uint64_t batch = std::min<uint64_t>(batch_nodes(), (J.total + TILE_M - 1) / TILE_M * TILE_M);
    if (T->cap_nodes < batch) {
        cudaStreamSynchronize(c->stream);
        cudaFree(T->buf_p); cudaFree(T->buf_q); cudaFree(T->dots); T->buf_p = T->buf_q = nullptr; T->dots = nullptr; T->cap_nodes = 0;
        const size_t tiles = (size_t)(batch / TILE_M);
        SMW_CUDA(c, cudaMalloc(&T->buf_p, tiles * MAX_KW * KSTEP_BYTES)); SMW_CUDA(c, cudaMalloc(&T->buf_q, tiles * MAX_KW * KSTEP_BYTES));
        SMW_CUDA(c, cudaMalloc(&T->dots, (size_t)MAX_G * 5 * batch * sizeof(double)));
        T->cap_nodes = batch;
    }
    const size_t tile_stride = (size_t)MAX_KW * KSTEP_BYTES;
    SideStream side(c);
    for (uint64_t node0 = 0; node0 < J.total; node0 += batch) {
        const uint64_t n = std::min<uint64_t>(batch, J.total - node0), n_tiles = (n + TILE_M - 1) / TILE_M;
        const bool split = n_streams() == 2 && n_tiles >= 8ull * (uint64_t)T->n_sm && std::find(c->side.begin(), c->side.end(), c->stream) == c->side.end();   // not inside another fan-out
        const uint64_t tiles0 = split ? (n_tiles + 1) / 2 : n_tiles, n0 = std::min<uint64_t>(n, tiles0 * TILE_M);
        int rc = SMW_OK;
        std::unique_ptr<ProfSpan> span;   // kernels of the two halves overlap: the profile gets ONE record for the batch (both events on the main stream)
        if (split) span.reset(new ProfSpan(c, J.kclass, c->stream));
        if (split) {                      // the second half first: it is enqueued on the side stream while the main stream is still empty
            side.enter();
            rc = run_sequence(c, p, T, J, node0 + n0, n - n0, T->buf_p + tiles0 * tile_stride, T->buf_q + tiles0 * tile_stride, T->dots + n0);
            c->stream = side.main;
        }
        if (!rc) rc = run_sequence(c, p, T, J, node0, n0, T->buf_p, T->buf_q, T->dots);
        if (split) side.leave();
        span.reset();
        if (rc) return rc;
    }
    int err = 0;
    SMW_CUDA(c, cudaMemcpyAsync(&err, T->d_err, 4, cudaMemcpyDeviceToHost, c->stream));
    SMW_CUDA(c, cudaStreamSynchronize(c->stream));
    if (err) { cudaMemsetAsync(T->d_err, 0, 4, c->stream); return set_err(c, SMW_ERR_CUDA, "tensor-core stage: mbarrier protocol timeout"); }
    return SMW_OK;
}

// Whether this parameter set can take the pipeline at all: FP64-pipe tables present, width 9 or 17, and the partial rounds divisible into
// groups (custom parameter sets with another rp — smw_params_from_seed / from_tables — stay on the interpreter instead of failing)
bool usable(const smw_params* p) {
    if (!p || !p->d_tables52 || (p->t != 17 && p->t != 9)) return false;
    const int G = group_size(p->t);
    return p->rp > 0 && p->rp % G == 0 && p->t + G <= MAX_KW;
}
// Hashes the nodes of `B` with the tensor-core pipeline: t = 17, every node with exactly 16 children (two rate blocks), or t = 9, every
// node with exactly 2 children (4 DS fields + 2 children + the padding one = 7 of the 8 rate lanes: one permutation).  The caller has checked the shape.
int merkle_level_tc(smw_ctx* c, smw_params* p, const LevelBatch& B) {
    Job J{}; J.kind = 0; J.B = B; J.total = B.total; J.kclass = KC_MERKLE_LEVEL;
    return run_job(c, p, J);
}
// FRI hashed leaves (deep_ali/src/fri.rs:38-44), one permutation each: d_segs == nullptr takes the single run `u`
int fri_leaves_tc(smw_ctx* c, smw_params* p, const uint32_t* d_tmpl, int lane_f, int lane_s, const LeafSeg* d_segs, int n_segs, const LeafSeg& u, uint64_t total) {
    Job J{}; J.kind = 1; J.R = LeafRuns{d_segs, n_segs, u, total}; J.tmpl = d_tmpl; J.lane_f = lane_f; J.lane_s = lane_s; J.total = total; J.kclass = KC_FRI_LEAF;
    return run_job(c, p, J);
}
// pair leaves of arity-16 trees (merkle/src/lib.rs:382-400), one permutation each
int pair_leaves_tc(smw_ctx* c, smw_params* p, const LeafSeg* d_segs, int n_segs, const LeafSeg& u, uint64_t total) {
    Job J{}; J.kind = 2; J.R = LeafRuns{d_segs, n_segs, u, total}; J.total = total; J.kclass = KC_PAIR_LEAVES;
    return run_job(c, p, J);
}

} // namespace tc
} // namespace smw
