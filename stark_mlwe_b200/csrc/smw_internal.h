// smw_internal.h — host-side structures behind the opaque C-ABI handles, and the launchers the ABI
// layer calls.  Everything here is product code: no oracle, no CPU arithmetic on field elements.
#pragma once
#include <cuda_runtime.h>
#include <cstdint>
#include <cstddef>
#include <string>
#include <vector>
#include <map>
#include <mutex>
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <nvtx3/nvToolsExt.h>
#include <nccl.h>                   // types and prototypes only: the functions are bound at run time (comm.cu)
#include "../../include/smw.h"

namespace smw {
struct PoseidonDev;
// element indices (EL-word elements) of the tables inside smw_params::d_tables
struct TableLayout { uint32_t lu_full, rc_full, rc_part, sp_row, sp_col, lu_post, mds, k_in, k_out, k_rr, k_one, post, total; };
struct Segs { int n; int first[8]; int count[8]; };   // (first, count) ranges of the op array, run in order
// One run of sibling groups of a Merkle level: parents[q] = H(DS(arity, level, q + pos_offset, label), children[q * arity ..]),
// q < n_parents (merkle/src/lib.rs:164-179).  `first` = index of the run's first node among all nodes of the launch.
struct NodeSeg { const uint32_t* children; uint32_t* parents; uint64_t n_children, n_parents, label, pos_offset; uint32_t level, arity; uint64_t first; };
// What one level launch covers: explicit runs (device array `segs`: levels of different trees hashed together, e.g. level j of
// every layer tree of a FRI commit phase) or, with segs == nullptr, `n_cols` equal columns derived from `u`.
struct LevelBatch { const NodeSeg* segs; int n_segs; NodeSeg u; int n_cols; uint64_t label_stride; size_t child_col_stride, parent_col_stride; uint64_t total; };
// Leaves of one FRI layer: out[i] = pair-leaf digest H(DS(arity, u32::MAX, i + pos_offset, label), [f_i, cp_{i / cp_div}]) (merkle/src/lib.rs:382-400)
// or hash_leaf_pair(f_i, cp_{i / cp_div}) (deep_ali/src/fri.rs:38-44); cp == nullptr stands for zeros.
struct LeafSeg { const uint32_t* f; const uint32_t* cp; uint32_t* out; uint64_t n, label, pos_offset, cp_div; uint32_t arity, pad; uint64_t first; };
}

struct smw_params {
    smw_ctx* ctx = nullptr;
    int t = 0, rp = 0;
    uint32_t* d_tables = nullptr;     // one allocation holding every internal-form table
    uint32_t* d_tables52 = nullptr;   // the FP64-pipe form of the same tables (widths 9 and 17 only; csrc/poseidon52.cuh)
    uint32_t* d_ext = nullptr;        // external-form copies (mds, rc_full, rc_partial) for export
    // device views (pointers into d_tables)
    const uint32_t *lu_full = nullptr, *rc_full = nullptr, *rc_part = nullptr, *sp_row = nullptr, *sp_col = nullptr, *lu_post = nullptr;
    smw::TableLayout layout{};
    // row-operation programs run by the interpreter in poseidon.cuh (built on the host by poseidon_prog.cu)
    std::vector<uint4> h_prog; uint4* d_prog = nullptr; size_t d_prog_cap = 0;
    int perm_first = 0, perm_count = 0, dense_first = 0, dense_count = 0;
    std::map<std::pair<int, int>, smw::Segs> sponge_cache;     // (n_ds, max_inputs) -> segments
    bool owned_by_ctx = false;
    void* tc = nullptr; void (*tc_free)(void*) = nullptr;     // tensor-core node-hash tables and work buffers (csrc/poseidon_tc.cu), built on first use
};

struct smw_ctx {
    int device = 0;
    cudaStream_t stream = nullptr;
    bool own_stream = true;
    std::string last_error;
    smw_params* commitment_default = nullptr;    // "POSEIDON-T17-X5-SEED"
    smw_params* transcript_default = nullptr;    // "POSEIDON-T17-X5-TRANSCRIPT"
    std::map<int, smw_params*> by_width;         // poseidon_params_for_width cache
    uint32_t* d_twiddles = nullptr;              // w_N^i, i < N/2, external Montgomery form
    int twiddle_log_n = 0;
    // grow-only device scratch slots reused across calls (cudaMalloc/cudaFree of GB-sized buffers cost ~100 ms per chunk)
    enum { SLOT_LDE_COEF = 0, SLOT_LDE_G, SLOT_TREE_A, SLOT_TREE_B, SLOT_EXT, SLOT_IN0, SLOT_IN1, SLOT_NTT_SCALE, N_SLOTS };
    void* slot_ptr[N_SLOTS] = {}; size_t slot_bytes[N_SLOTS] = {};
    void* h_pinned = nullptr; size_t pinned_bytes = 0;
    // side streams for independent launch chains (the per-layer trees of a FRI commit phase), joined back into `stream`
    std::vector<cudaStream_t> side; std::vector<cudaEvent_t> side_ev; cudaEvent_t fan_ev = nullptr;
    bool use_f52 = true;                         // widths 9/17 batch kernels on the FP64-pipe multiplier (SMW_REP=29 in the environment selects the integer one)
    uint64_t tc_min_nodes = 16384;               // full arity-16 levels of at least this many nodes take the tensor-core pipeline (poseidon_tc.cu); SMW_TC_MIN / smw_ctx_set_tc_min_nodes
    uint64_t launches = 0;                       // kernels launched through this context (bench.py's gpu_launches)
    uint64_t collectives = 0;                    // NCCL collectives / grouped exchanges issued through this context
    // optional per-kernel-class device timing (CUDA events on this context's stream), for bench.py's roofline
    // Arena pool: the large temporaries of the fold-and-recommit entry points (FRI layers, leaf digests, tree levels, exchange buffers) are
    // handed back here instead of to cudaFree and reused by the next call of a similar size.  cudaMalloc / cudaFree of GB-sized blocks cost
    // 10-200 ms each and varied from call to call (cfg5 FRI at 2^26: 1.51-1.77 s on one GPU, 649 ms on four where the kernels need ~400).
    // Blocks are only returned after the stream has drained (every owner synchronises before freeing), so reuse is stream-safe.
    std::map<void*, size_t> pool_owned;                       // every live block that came from pool_malloc: capacity in bytes
    std::vector<std::pair<void*, size_t>> pool_freelist;      // returned blocks, oldest first
    size_t pool_free_bytes = 0;
    bool prof_on = false;
    // prof_span: launches inside a span are counted into it instead of being timed one by one (kernels of two streams overlap: their
    // individual durations would add up to more than the wall time); n = launches the record stands for
    int prof_span = -1;
    struct ProfRec { int cls; cudaEvent_t a, b; uint64_t n = 1; };
    std::vector<ProfRec> prof;
    std::vector<cudaEvent_t> prof_pool;
};

struct smw_tree {
    smw_ctx* ctx = nullptr;
    smw_params* params = nullptr;
    int arity = 0;
    uint64_t tree_label = 0;
    std::vector<size_t> len;          // level 0 = leaves
    std::vector<uint32_t*> lv;        // device pointers, external form
    std::vector<bool> owned;
    uint32_t* arena = nullptr;        // the levels above level 0, when they share one allocation
};

struct smw_comm {
    smw_ctx* ctx = nullptr;
    ncclComm_t comm = nullptr;
    int rank = 0, world = 1;
};

struct smw_transcript {
    smw_ctx* ctx = nullptr;
    smw_params* params = nullptr;
    uint32_t* d_state = nullptr;      // 17 elements, external form; followed in the same allocation by 64 immediate words and one output slot
    uint32_t* d_imm = nullptr;        // scratch for the 256-bit immediates of one call (tags, label words): no allocation per absorb / challenge
    uint32_t* d_out = nullptr;
    int pos = 0;                      // rate cursor: a function of the call sequence only, so tracked on the host
};

struct smw_fri {
    smw_ctx* ctx = nullptr;
    int L = 0;
    std::vector<size_t> n, m;
    std::vector<int> arity, hashed;
    std::vector<uint32_t*> f;         // device, external form
    uint32_t* d_z = nullptr;          // L elements
    std::vector<smw_tree*> tree;
    uint32_t *arena_f = nullptr, *arena_leaves = nullptr, *arena_levels = nullptr;   // f layers, leaf digests and tree levels: one allocation each
};

namespace smw {

// A named host-side range: an NVTX range (visible in nsys / ncu --nvtx; a no-op without a tool attached) and, with SMW_TRACE=1 in
// the environment, a line on stderr with the host time spent inside — what located the per-call overheads of the small builds.
inline bool trace_on() { static int v = -1; if (v < 0) { const char* e = getenv("SMW_TRACE"); v = (e && *e && *e != '0') ? 1 : 0; } return v == 1; }
struct Range {
    const char* name; std::chrono::steady_clock::time_point t0;
    explicit Range(const char* n) : name(n) { nvtxRangePushA(n); if (trace_on()) t0 = std::chrono::steady_clock::now(); }
    ~Range() {
        nvtxRangePop();
        if (trace_on()) fprintf(stderr, "[smw] %-28s %9.3f ms (host)\n", name, std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count());
    }
};
static const char* const KC_NAMES[] = {"merkle_level", "pair_leaves", "fri_leaf", "permute", "ntt_pass", "lde_aux", "fold", "misc"};

// kernel classes for the optional device-side timing (smw_ctx_profile_*)
enum { KC_MERKLE_LEVEL = 0, KC_PAIR_LEAVES, KC_FRI_LEAF, KC_PERMUTE, KC_NTT_PASS, KC_LDE_AUX, KC_FOLD, KC_MISC, KC_COUNT };
// counts one launch; when profiling is on, brackets it with two events on the context's stream
struct LaunchScope {
    smw_ctx* c; cudaEvent_t b = nullptr;
    LaunchScope(smw_ctx* ctx, int cls) : c(ctx) {
        c->launches++;
        nvtxRangePushA(KC_NAMES[cls < 8 ? cls : 7]);
        if (!c->prof_on) return;
        if (c->prof_span >= 0) { c->prof[(size_t)c->prof_span].n++; return; }
        cudaEvent_t a;
        if (c->prof_pool.size() >= 2) { a = c->prof_pool.back(); c->prof_pool.pop_back(); b = c->prof_pool.back(); c->prof_pool.pop_back(); }
        else { cudaEventCreate(&a); cudaEventCreate(&b); }
        cudaEventRecord(a, c->stream);
        c->prof.push_back({cls, a, b});
    }
    ~LaunchScope() { if (b) cudaEventRecord(b, c->stream); nvtxRangePop(); }
};
// One timed record for a whole group of launches that run on several streams: both events go to `main` (the first before the streams
// fork, the second after they have been joined), the launches in between are only counted.  No-op unless profiling is on.
struct ProfSpan {
    smw_ctx* c; cudaStream_t main; bool on = false;
    ProfSpan(smw_ctx* ctx, int cls, cudaStream_t main_stream) : c(ctx), main(main_stream) {
        if (!c->prof_on || c->prof_span >= 0) return;
        cudaEvent_t a, b;
        if (c->prof_pool.size() >= 2) { a = c->prof_pool.back(); c->prof_pool.pop_back(); b = c->prof_pool.back(); c->prof_pool.pop_back(); }
        else { cudaEventCreate(&a); cudaEventCreate(&b); }
        cudaEventRecord(a, main);
        smw_ctx::ProfRec r; r.cls = cls; r.a = a; r.b = b; r.n = 0;
        c->prof.push_back(r); c->prof_span = (int)c->prof.size() - 1; on = true;
    }
    ~ProfSpan() { if (on) { cudaEventRecord(c->prof[(size_t)c->prof_span].b, main); c->prof_span = -1; } }
};

// returns a device buffer of at least `bytes` for `slot` (valid until the next request for the same slot on this context)
int ctx_scratch(smw_ctx* c, int slot, size_t bytes, void** out);
cudaError_t pool_malloc(smw_ctx* c, void** out, size_t bytes);   // blocks >= 1 MiB are pooled, smaller ones are plain cudaMalloc
void pool_free(smw_ctx* c, void* p);                              // the caller has synchronised the stream(s) that used p; unknown pointers go to cudaFree
void pool_release(smw_ctx* c);                                     // cudaFree every returned block
cudaError_t pool_malloc(smw_ctx* c, void** out, size_t bytes);   // blocks >= 1 MiB are pooled, smaller ones are plain cudaMalloc
void pool_free(smw_ctx* c, void* p);                              // the caller has synchronised the stream(s) that used p; unknown pointers go to cudaFree
void pool_release(smw_ctx* c);                                     // cudaFree every returned block

// NCCL entry points, bound by dlopen (comm.cu)
struct NcclApi {
    decltype(&ncclGetUniqueId) GetUniqueId; decltype(&ncclCommInitRank) CommInitRank; decltype(&ncclCommDestroy) CommDestroy;
    decltype(&ncclAllGather) AllGather; decltype(&ncclSend) Send; decltype(&ncclRecv) Recv;
    decltype(&ncclGroupStart) GroupStart; decltype(&ncclGroupEnd) GroupEnd; decltype(&ncclGetErrorString) GetErrorString;
};
const NcclApi* nccl_api(std::string* err);
int comm_all_gather_bytes(smw_ctx* c, smw_comm* m, const void* send, void* recv, size_t bytes);        // m == nullptr or world 1: a device copy
int comm_all_to_all_bytes(smw_ctx* c, smw_comm* m, const void* send, void* recv, size_t piece_bytes);  // grouped ncclSend / ncclRecv

// status helpers
int set_err(smw_ctx* c, int code, const char* what);
int cuda_err(smw_ctx* c, cudaError_t e, const char* what);
#define SMW_CUDA(ctx, call) do { cudaError_t e__ = (call); if (e__ != cudaSuccess) return smw::cuda_err((ctx), e__, #call); } while (0)

// programs (poseidon_prog.cu)
TableLayout table_layout(int t, int rp);
void build_perm_program(std::vector<uint4>& ops, int t, int rp, bool sparse);
void mark_deps(std::vector<uint4>& ops, size_t from);                             // sets OPF_DEP (poseidon.cuh) on ops[from..]
uint64_t count_macs(const std::vector<uint4>& ops, int first, int count);         // executed 32x32+64 multiply-accumulates
Segs build_sponge_program(std::vector<uint4>& ops, int t, int rp, int n_ds, int max_inputs, int perm_first, int perm_count);
int upload_programs(smw_ctx* c, smw_params* p);                                   // h_prog -> d_prog (re-allocates when it grew)
int get_sponge_segs(smw_ctx* c, smw_params* p, int n_ds, int max_inputs, Segs* out); // cached per parameter set

// launchers (poseidon_kernels.cu)
int launch_derive_params(smw_ctx* c, smw_params* p, const unsigned char* seed, int seed_len);            // BLAKE3 -> tables
int launch_upload_params(smw_ctx* c, smw_params* p, const uint32_t* d_mds, const uint32_t* d_rcf, const uint32_t* d_rcp); // device ext tables -> internal
int launch_permute_batch(smw_ctx* c, const smw_params* p, uint32_t* d_states, size_t batch, bool sparse);
int launch_merkle_level(smw_ctx* c, const smw_params* p, const uint32_t* d_children, size_t n_children, uint32_t* d_parents, size_t n_parents,
                        int arity, uint32_t level, uint64_t label0, uint64_t label_stride, uint64_t pos_offset,
                        int n_cols, size_t child_col_stride, size_t parent_col_stride);
int launch_pair_leaves(smw_ctx* c, const smw_params* p, const uint32_t* d_f, const uint32_t* d_cp, size_t n, int arity, uint64_t label, uint64_t pos_offset,
                       uint32_t* d_out, size_t cp_div);
int launch_hash_ds_single(smw_ctx* c, const smw_params* p, const uint32_t* d_ds, int nds, const uint32_t* d_in, size_t n, uint32_t* d_out);
// out[g] = H(DS(arity, level, pos[g], label), pool[child_idx[g*row + c]] for c < cnt[g] <= row) — recomputed parents / pair-leaf digests of a verifier
int launch_hash_groups(smw_ctx* c, const smw_params* p, const uint32_t* d_pool, const uint64_t* d_child_idx, const uint8_t* d_cnt, const uint64_t* d_pos,
                       size_t n_groups, int arity, int row, uint32_t level, uint64_t label, uint32_t* d_out);
// legacy-sponge nodes, one warp each: mode 0 chunks of 16 of d_src, 1 pairs (d_src[g], d_cp[g]), 2 gathered pool entries (d_idx rows of 16, d_cnt)
int launch_legacy_nodes(smw_ctx* c, const smw_params* p, int mode, const uint32_t* d_src, const uint32_t* d_cp, size_t n_src, const uint64_t* d_idx, const uint8_t* d_cnt,
                        size_t n_out, const uint32_t* d_tag, uint32_t* d_out);
// count one-shot hashes sharing a 17-lane template (LE256 words), item i's k elements in lanes lane0.. : one warp and one permutation each
int launch_tr_hash_batch(smw_ctx* c, const smw_params* trp, const uint32_t* d_tmpl, int lane0, int k, const uint32_t* d_xs, size_t count, uint32_t* d_out);
int launch_hash_legacy_single(smw_ctx* c, const smw_params* p, const uint32_t* d_in, size_t n, const uint32_t* d_tag, uint32_t* d_out);
// several runs of one level in one launch (explicit device segment array); all runs share the parameter set and the arity
int launch_merkle_level_multi(smw_ctx* c, const smw_params* p, const NodeSeg* d_segs, int n_segs, uint64_t total_nodes, int arity, bool all_full = false);   // all_full: every node of every run has exactly `arity` children
int launch_pair_leaves_multi(smw_ctx* c, const smw_params* p, const LeafSeg* d_segs, int n_segs, uint64_t total);
int launch_fri_leaf_hashed_multi(smw_ctx* c, const smw_params* tr_params, const uint32_t* d_template_le256, int lane_f, int lane_s, const LeafSeg* d_segs, int n_segs, uint64_t total);
// transcript program: state' = run(state, pos; immediates A, ext elements B, immediates C, optional final permute)
int launch_transcript_run(smw_ctx* c, const smw_params* p, uint32_t* d_state, int fresh, int pos,
                          const uint32_t* d_immA, int nA, const uint32_t* d_extB, size_t nB, const uint32_t* d_immC, int nC, int final_permute,
                          uint32_t* d_out /* optional: state[0] after the run */);
// the B phase of launch_transcript_run for long inputs: one warp, literal dense rounds with one state lane per thread; the state must live in d_state
int launch_transcript_absorb_coop(smw_ctx* c, const smw_params* p, uint32_t* d_state, int pos, const uint32_t* d_ext, size_t n);
int launch_fri_leaf_hashed(smw_ctx* c, const smw_params* tr_params, const uint32_t* d_template_le256 /* 17 x 8 words */, int lane_f, int lane_s,
                           const uint32_t* d_f, const uint32_t* d_s, size_t s_div, size_t n, uint32_t* d_out);
int launch_fri_sample_z(smw_ctx* c, const smw_params* tr_params, const uint32_t* d_templates_le256 /* L x 17 x 8 */, const uint64_t* d_level_size /* L x 2 */, int L,
                        uint64_t seed_z, uint32_t* d_z);

// per-width launchers (poseidon_T.cu, one translation unit per width)
template <int T> int permute_batch_T(smw_ctx* c, const smw_params* p, uint32_t* d_states, size_t batch, bool sparse);
template <int T> int merkle_level_T(smw_ctx* c, const smw_params* p, const uint32_t* d_children, size_t n_children, uint32_t* d_parents, size_t n_parents,
                                    int arity, uint32_t level, uint64_t label0, uint64_t label_stride, uint64_t pos_offset,
                                    int n_cols, size_t child_col_stride, size_t parent_col_stride);
template <int T> int pair_leaves_T(smw_ctx* c, const smw_params* p, const uint32_t* d_f, const uint32_t* d_cp, size_t n, int arity, uint64_t label, uint64_t pos_offset,
                                   uint32_t* d_out, size_t cp_div);
template <int T> int hash_groups_T(smw_ctx* c, const smw_params* p, const uint32_t* d_pool, const uint64_t* d_child_idx, const uint8_t* d_cnt, const uint64_t* d_pos,
                                   size_t n_groups, int arity, int row, uint32_t level, uint64_t label, uint32_t* d_out);
namespace tc {
int merkle_level_tc(smw_ctx* c, smw_params* p, const LevelBatch& B); void node_counts(int t, int rp, uint64_t out[5]); bool usable(const smw_params* p);
int fri_leaves_tc(smw_ctx* c, smw_params* p, const uint32_t* d_tmpl, int lane_f, int lane_s, const LeafSeg* d_segs, int n_segs, const LeafSeg& u, uint64_t total);
int pair_leaves_tc(smw_ctx* c, smw_params* p, const LeafSeg* d_segs, int n_segs, const LeafSeg& u, uint64_t total);
}     // poseidon_tc.cu
template <int T> int merkle_level_multi_T(smw_ctx* c, const smw_params* p, const NodeSeg* d_segs, int n_segs, uint64_t total_nodes, int arity, bool all_full);
template <int T> int pair_leaves_multi_T(smw_ctx* c, const smw_params* p, const LeafSeg* d_segs, int n_segs, uint64_t total);
template <int T> int hash_ds_single_T(smw_ctx* c, const smw_params* p, const uint32_t* d_ds, int nds, const uint32_t* d_in, size_t n, uint32_t* d_out);

// field_kernels.cu
int launch_fri_fold(smw_ctx* c, const uint32_t* d_f, size_t n, const uint32_t* d_z, size_t m, uint32_t* d_out);
int launch_broadcast_div(smw_ctx* c, const uint32_t* d_src, size_t n, size_t m, uint32_t* d_out);   // out[i] = src[i / m]
int launch_to_mont(smw_ctx* c, const uint32_t* d_in, uint32_t* d_out, size_t n, int inverse);
int launch_mf_coeffs(smw_ctx* c, const uint32_t* d_layer, size_t n, uint32_t* d_c0c1);
int launch_mf_fold(smw_ctx* c, const uint32_t* d_cur, size_t n, const uint32_t* d_r, uint32_t* d_next);
int launch_gather(smw_ctx* c, const uint32_t* d_src, const uint64_t* d_idx, size_t n, uint32_t* d_out);   // out[i] = src[idx[i]]
int launch_deep_ali_merge(smw_ctx* c, const uint32_t* a, const uint32_t* s, const uint32_t* e, const uint32_t* t, size_t n, const uint32_t* d_omega, const uint32_t* d_z, uint32_t* d_f0);

int launch_scale_powers(smw_ctx* c, uint32_t* d_data, size_t n, const uint32_t* d_g);                     // data[i] *= g^i
int launch_dft_strided(smw_ctx* c, uint32_t* d_data, int log_g, size_t stride, size_t count, const uint32_t* d_wpow);
int launch_fr_pow(smw_ctx* c, const uint32_t* d_base, uint64_t e, uint32_t* d_out);                       // out = base^e (one thread)
int launch_index_from_seed(smw_ctx* c, const uint32_t* d_seeds, size_t n, uint64_t n_pow2, uint64_t* d_out);
int launch_ali_sample(smw_ctx* c, const uint32_t* d_fused, const uint32_t* d_roots_seed, uint64_t n0, uint32_t* d_z, uint32_t* d_beta);
int launch_f52_dot(smw_ctx* c, const uint32_t* d_a, const uint32_t* d_b, size_t n_out, int k, uint32_t* d_out);
int launch_diag_mac_rate(smw_ctx* c, double* mac_per_s, double* warp_instr_per_clk_per_sm);
int launch_diag_rate(smw_ctx* c, int fp64, double* per_s, double* warp_instr_per_clk_per_sm);       // fp64 = 1: DFMA.RZ issue rate
int launch_diag_rate_ex(smw_ctx* c, int fp64, double* per_s, double* warp_instr_per_clk_per_sm, double* sm_hz_during, double* n_sm, double* sm_max_hz);
uint64_t count_fp64_ops(const std::vector<uint4>& ops, int first, int count);                          // FP64-pipe instructions of the fr52 path

// ntt_kernels.cu
int ensure_twiddles(smw_ctx* c, int log_n);
int launch_ntt(smw_ctx* c, uint32_t* d_buf, int log_n, int inverse, int n_cols, size_t col_stride);
int launch_lde(smw_ctx* c, const uint32_t* d_in, int log_n, int log_blowup, const uint32_t* d_shift, uint32_t* d_out, int n_cols);
int launch_root_of_unity(smw_ctx* c, int log_n, uint32_t* d_out);

} // namespace smw
