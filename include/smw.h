/*
 * smw.h — C ABI of libsmw.so: the B200 (sm_100a) implementation of stark-mlwe's commitment hot path.
 *
 * One entry point per reference function on the path; each declaration cites the
 * reference interface it replaces (paths relative to the reference repository root).  A Rust shim
 * binds these with `extern "C"` (INTEGRATION.md shows the stubs); no torch / C++ types cross the ABI.
 *
 * Conventions
 *  - Field elements ("Fr") are 32 bytes: four little-endian u64 limbs in Montgomery form, R = 2^256,
 *    canonical (< p) — byte-identical to ark-ff 0.5 `Fp256<MontBackend<ark_pallas::FrConfig,4>>`
 *    (field/src/lib.rs:13), so a Rust `&[F]` can be passed without conversion.
 *    smw_fr_from_canonical / smw_fr_to_canonical convert to/from canonical little-endian integers
 *    (the ark-serialize wire form, field/src/lib.rs:206-215).
 *  - `mem` says where a caller buffer lives: SMW_HOST (copied through the context's stream) or
 *    SMW_DEVICE (a CUDA device pointer on the context's device, used in place).
 *  - Every function returns an int32 status (SMW_OK = 0).  The reference panics on the same
 *    conditions (file:line given per code); a shim turns non-zero into that panic.
 *  - A context is bound to one device and one stream and is not thread-safe; use one per GPU.
 *  - All arithmetic runs on the GPU.  There is no CPU fallback: without a CUDA device
 *    smw_ctx_create fails with SMW_ERR_CUDA.
 */
#ifndef SMW_H
#define SMW_H
#include <stdint.h>
#include <stddef.h>
#ifdef __cplusplus
extern "C" {
#endif
#if defined(__GNUC__)
#pragma GCC visibility push(default)
#endif

typedef struct smw_ctx smw_ctx;
typedef struct smw_params smw_params;          /* PoseidonParamsDynamic (poseidon/src/lib.rs:104-114), device resident */
typedef struct smw_tree smw_tree;              /* MerkleTree (merkle/src/lib.rs:114-128) */
typedef struct smw_proof smw_proof;            /* MerkleProof (merkle/src/lib.rs:131-143) */
typedef struct smw_transcript smw_transcript;  /* Transcript (transcript/src/lib.rs:48-52) */
typedef struct smw_fri smw_fri;                /* FriProverState (deep_ali/src/fri.rs:210-216) */

enum {
    SMW_OK = 0,
    SMW_ERR_CUDA = 1,              /* CUDA runtime failure; see smw_last_error */
    SMW_ERR_EMPTY = 2,             /* "no leaves"                     merkle/src/lib.rs:148,394 */
    SMW_ERR_ARITY_WIDTH = 3,       /* arity incompatible with width   merkle/src/lib.rs:161,413 */
    SMW_ERR_LENGTH_MISMATCH = 4,   /* "f and cp length mismatch"      merkle/src/lib.rs:393 */
    SMW_ERR_NOT_DIVISIBLE = 5,     /* layer size not divisible by m   deep_ali/src/fri.rs:87,124 */
    SMW_ERR_UNSUPPORTED_WIDTH = 6, /* unsupported Poseidon width      poseidon/src/lib.rs:127 */
    SMW_ERR_UNSUPPORTED_ARITY = 7, /* unsupported Merkle arity        poseidon/src/lib.rs:163 */
    SMW_ERR_INVALID_ARG = 8,
    SMW_ERR_PARAMS_SINGULAR = 9,   /* matrix has a zero pivot: the exact sparse/LU factorisation does not exist */
    SMW_ERR_NOMEM = 10,
    SMW_ERR_EMPTY_INDICES = 11,    /* "open_many: empty indices"      merkle/src/lib.rs:247 */
    SMW_ERR_Z_IN_DOMAIN = 12,      /* "z must be outside H"           deep_ali/src/lib.rs:77,97 */
    SMW_ERR_COMM = 13              /* NCCL failure or no communicator; see smw_last_error */
};
enum { SMW_HOST = 0, SMW_DEVICE = 1 };

/* ---- context ------------------------------------------------------------------------------ */
int32_t smw_ctx_create(int32_t device, smw_ctx** out);
int32_t smw_ctx_create_on_stream(int32_t device, void* cuda_stream, smw_ctx** out);
void smw_ctx_destroy(smw_ctx* ctx);
int32_t smw_ctx_sync(smw_ctx* ctx);
/* The context keeps the large device temporaries of the fold-and-recommit entry points (FRI layers, leaf digests, tree levels, exchange
 * buffers: up to 16 blocks / 24 GiB) for the next call instead of returning them to the driver after every call.  This gives all idle
 * blocks back now; smw_ctx_destroy does the same.  Not in the reference (it has no device). */
int32_t smw_ctx_release_cached(smw_ctx* ctx);
const char* smw_last_error(const smw_ctx* ctx);
uint64_t smw_ctx_launch_count(const smw_ctx* ctx);   /* kernels launched so far through this context */
void* smw_ctx_stream(const smw_ctx* ctx);
/* measurement aids (no reference counterpart): per-kernel-class device time from CUDA events on the context's
 * stream — classes: 0 merkle_level, 1 pair_leaves, 2 fri_leaf, 3 permute, 4 ntt_pass, 5 lde_aux, 6 fold, 7 misc —
 * and the measured 32x32+64->64 multiply-accumulate (IMAD.WIDE.U32) issue rate that the roofline fractions use */
int32_t smw_ctx_profile_enable(smw_ctx* ctx, int32_t on);
int32_t smw_ctx_profile_read(smw_ctx* ctx, double* ms_out, uint64_t* launches_out, int32_t n_classes);   /* syncs, sums, resets */
int32_t smw_diag_mac_rate(smw_ctx* ctx, double* mac_per_s, double* warp_instr_per_clk_per_sm);
/* 32x32+64 multiply-accumulates one production permutation of this parameter set executes (counted from its program) */
/* the FP64-pipe counterparts: measured DFMA.RZ issue rate, and the FP64 instructions (DFMA + DADD) one production permutation of
 * this parameter set executes when it runs on the fr52 multiplier (0 otherwise) */
int32_t smw_diag_fp64_rate(smw_ctx* ctx, double* dfma_per_s, double* warp_instr_per_clk_per_sm);
/* the same measurements with what is needed to re-derive them: out5 = {lanes per s, warp instructions per clock per SM, SM clock (Hz) held
 * during the measurement (in-kernel clock64 ticks / event time), SM count, nominal maximum SM clock (Hz)}; fp64 = 0: IMAD.WIDE.U32, 1: DFMA.RZ */
int32_t smw_diag_pipe_rate(smw_ctx* ctx, int32_t fp64, double* out5);
int32_t smw_diag_perm_fp64_ops(const smw_params* params, uint64_t* ops_out);
/* self-test of the FP64-pipe field arithmetic (csrc/fr52.cuh): out[i] = sum_{j<k} a[i k + j] * b[i k + j] */
int32_t smw_diag_f52_dot(smw_ctx* ctx, const void* a, const void* b, size_t n_out, int32_t k, void* out, int32_t mem);
int32_t smw_diag_perm_macs(const smw_params* params, uint64_t* macs_out);

/* ---- field (field/src/lib.rs:13,43-53,206-215; ark-ff) ------------------------------------- */
int32_t smw_fr_from_canonical(smw_ctx* ctx, const void* canon_le32, void* out_fr, size_t n, int32_t mem);
int32_t smw_fr_to_canonical(smw_ctx* ctx, const void* fr, void* out_canon_le32, size_t n, int32_t mem);
int32_t smw_root_of_unity(smw_ctx* ctx, int32_t log_n, void* out_fr_host);   /* F::get_root_of_unity(2^log_n), field/src/lib.rs:46 */

/* ---- Poseidon parameters (poseidon/src/lib.rs:120-216, 318-356; utils/src/lib.rs:7-22) ---- */
int32_t smw_params_from_seed(smw_ctx* ctx, const uint8_t* seed, size_t seed_len, int32_t t, int32_t rp, smw_params** out); /* params::generate_params_t17_x5(seed) for t=17, rp=64 */
int32_t smw_params_for_width(smw_ctx* ctx, int32_t t, smw_params** out);          /* poseidon_params_for_width; cached, owned by ctx */
int32_t smw_params_for_arity(smw_ctx* ctx, int32_t arity, smw_params** out);      /* poseidon_params_for_arity; cached, owned by ctx */
int32_t smw_params_commitment_default(smw_ctx* ctx, smw_params** out);            /* commitment::default_params (commitment/src/lib.rs:48-51) */
int32_t smw_params_transcript_default(smw_ctx* ctx, smw_params** out);            /* transcript::default_params (transcript/src/lib.rs:44-46) */
int32_t smw_params_from_tables(smw_ctx* ctx, int32_t t, int32_t rp, const void* mds, const void* rc_full, const void* rc_partial, int32_t mem, smw_params** out); /* caller-supplied PoseidonParamsDynamic */
int32_t smw_params_export(smw_params* p, void* mds_out, void* rc_full_out, void* rc_partial_out);   /* host buffers: t*t, 8*t, rp Fr */
int32_t smw_params_width(const smw_params* p);
int32_t smw_params_partial_rounds(const smw_params* p);
void smw_params_free(smw_params* p);   /* no-op for ctx-owned (cached) sets */

/* ---- permutation and sponges (poseidon/src/lib.rs:31-68, 85-100, 219-312) ------------------ */
int32_t smw_poseidon_permute(smw_ctx* ctx, const smw_params* p, void* states, size_t batch, int32_t mem);              /* permute / permute_dynamic, in place, [batch][t] Fr */
int32_t smw_poseidon_permute_dense(smw_ctx* ctx, const smw_params* p, void* states, size_t batch, int32_t mem);        /* same result, literal round structure (test aid) */
int32_t smw_hash_with_ds_dynamic(smw_ctx* ctx, const smw_params* p, const void* ds_fields, size_t n_ds, const void* inputs, size_t n, void* out_fr, int32_t mem);
int32_t smw_hash_with_ds(smw_ctx* ctx, const smw_params* p, const void* inputs, size_t n, const void* ds_tag, void* out_fr, int32_t mem);   /* legacy, t = 17 */

/* ---- Merkle (merkle/src/lib.rs:58-112, 147-193, 380-445) ----------------------------------- */
int32_t smw_merkle_new(smw_ctx* ctx, smw_params* p, int32_t arity, uint64_t tree_label, const void* leaves, size_t n, int32_t mem, smw_tree** out);               /* MerkleTree::new */
int32_t smw_merkle_new_pairs(smw_ctx* ctx, smw_params* p, int32_t arity, uint64_t tree_label, const void* f, const void* cp, size_t n_f, size_t n_cp, int32_t mem, smw_tree** out); /* MerkleTree::new_pairs */
int32_t smw_tree_root(smw_tree* t, void* out_fr_host);          /* MerkleTree::root */
int32_t smw_tree_num_levels(const smw_tree* t);                 /* levels.len(); level 0 = leaves */
size_t smw_tree_level_len(const smw_tree* t, int32_t level);
int32_t smw_tree_level(smw_tree* t, int32_t level, void* out, int32_t mem);
int32_t smw_tree_arity(const smw_tree* t);
void smw_tree_free(smw_tree* t);
/* many equally sized columns at once: one tree per column, tree_label = label0 + c*label_stride; roots only
 * (the device-side column-batched form of calling MerkleTree::new once per column) */
int32_t smw_merkle_commit_columns(smw_ctx* ctx, smw_params* p, int32_t arity, uint64_t label0, uint64_t label_stride,
                                  const void* cols, size_t n, int32_t n_cols, int32_t mem, void* roots_out_host);

/* ---- commitment (commitment/src/lib.rs:13-27, 60-113; channel/src/lib.rs:172-179) ---------- */
int32_t smw_commitment_commit(smw_ctx* ctx, uint64_t ds_tag, smw_params* params_or_null, const void* leaves, size_t n, int32_t mem, smw_tree** out); /* MerkleCommitment::commit: arity 16, tree_label = ds_tag */

/* ---- NTT / coset LDE (fft/src/lib.rs:6-32) ---------------------------------- */
int32_t smw_ntt(smw_ctx* ctx, void* buf, int32_t log_n, int32_t inverse, int32_t n_cols, int32_t mem);   /* fft_in_place / ifft_in_place per column, natural order in and out */
/* building blocks of the cross-GPU four-step NTT (no reference counterpart): buf[i] *= g^i, and an in-place
 * DFT of length 2^log_g (<= 16) across stride-spaced elements for every offset j < count */
int32_t smw_scale_powers(smw_ctx* ctx, void* buf, size_t n, const void* g_fr_host, int32_t mem);
int32_t smw_dft_strided(smw_ctx* ctx, void* buf, int32_t log_g, size_t stride, size_t count, int32_t mem);
int32_t smw_lde(smw_ctx* ctx, const void* in, int32_t log_n, int32_t log_blowup, const void* shift_fr_host, void* out, int32_t n_cols, int32_t mem);
/* the benchmark pipeline: LDE of n_cols columns, then one arity-16 commitment-default tree per extended column */
int32_t smw_lde_commit(smw_ctx* ctx, const void* cols, int32_t log_n, int32_t log_blowup, const void* shift_fr_host, int32_t n_cols,
                       uint64_t label0, int32_t mem, void* roots_out_host);

/* ---- transcript (transcript/src/lib.rs:55-117; deep_ali/src/fri.rs:28-44) ------------------ */
int32_t smw_transcript_new(smw_ctx* ctx, const uint8_t* label, size_t len, smw_params* params_or_null, smw_transcript** out);
int32_t smw_transcript_absorb_bytes(smw_transcript* t, const uint8_t* bytes, size_t len);
int32_t smw_transcript_absorb_field(smw_transcript* t, const void* fr_host);
int32_t smw_transcript_absorb_fields(smw_transcript* t, const void* frs, size_t n, int32_t mem);
int32_t smw_transcript_challenge(smw_transcript* t, const uint8_t* label, size_t len, void* out_fr_host);
int32_t smw_transcript_state(smw_transcript* t, void* out_17_fr_host, int32_t* pos_out);
void smw_transcript_free(smw_transcript* t);
int32_t smw_tr_hash_fields_tagged(smw_ctx* ctx, const uint8_t* tag, size_t tag_len, const void* xs, size_t n, int32_t mem, void* out_fr_host);
/* `count` independent tr_hash_fields_tagged calls whose serial sponges overlap on side streams (the four ALI sponges of DeepAliRealBuilder::build_f0, deep_ali/src/fri.rs:541-548) */
/* `count` one-shot hashes sharing `tag`, k field elements each (xs: count x k), one launch, one permutation per hash — the
 * index seeds of fri_prove_queries (deep_ali/src/fri.rs:184-191, 388-400).  The pre-image must fit one rate block (k <= ~8). */
int32_t smw_tr_hash_fields_tagged_batch(smw_ctx* ctx, const uint8_t* tag, size_t tag_len, const void* xs, int32_t k, size_t count, int32_t mem, void* out);
int32_t smw_tr_hash_fields_tagged_many(smw_ctx* ctx, int32_t count, const uint8_t* const* tags, const size_t* tag_lens, const void* const* xs, const size_t* n,
                                       int32_t mem, void* out_fr_host);
int32_t smw_hash_leaf_pair(smw_ctx* ctx, const void* f, const void* s, size_t n, void* out, int32_t mem);   /* hash_leaf_pair, element-wise */

/* ---- FRI-like commit phase (deep_ali/src/fri.rs:59-143, 220-312) --------------------------- */
int32_t smw_fri_sample_z_ell(smw_ctx* ctx, uint64_t seed_z, uint64_t level, uint64_t domain_size, void* out_fr_host);
int32_t smw_fri_fold_layer(smw_ctx* ctx, const void* f, size_t n, const void* z_fr_host, size_t m, void* out, int32_t mem);
int32_t smw_compute_s_layer(smw_ctx* ctx, const void* f, size_t n, const void* z_fr_host, size_t m, void* out, int32_t mem);
int32_t smw_pick_arity_for_layer(size_t n, size_t requested_m);
int32_t smw_fri_build_transcript(smw_ctx* ctx, const void* f0, size_t n0, const uint64_t* schedule, int32_t L, uint64_t seed_z, int32_t mem, smw_fri** out);
int32_t smw_fri_num_layers(const smw_fri* s);                                          /* L + 1 */
int32_t smw_fri_layer_info(const smw_fri* s, int32_t l, size_t* n, size_t* m, int32_t* arity, int32_t* hashed_leaves);
int32_t smw_fri_layer_f(smw_fri* s, int32_t l, void* out, int32_t mem);
int32_t smw_fri_layer_s(smw_fri* s, int32_t l, void* out, int32_t mem);                /* s_l[i] = f_{l+1}[i / m]; zeros for the last layer */
int32_t smw_fri_roots(smw_fri* s, void* out_host);                                     /* L + 1 Fr */
int32_t smw_fri_z(smw_fri* s, void* out_host);                                         /* L Fr */
smw_tree* smw_fri_layer_tree(smw_fri* s, int32_t l);                                   /* borrowed */
void smw_fri_free(smw_fri* s);
/* layers level0.. of a larger instance: tree_label and the z_l level index start at level0 (fri.rs:250,284-289) */
int32_t smw_fri_build_transcript_from(smw_ctx* ctx, const void* f, size_t n, const uint64_t* schedule, int32_t L, uint64_t seed_z, uint64_t level0, int32_t mem, smw_fri** out);

/* ---- range-sharded commit phase of ONE large polynomial (no reference counterpart: the reference is
 * single-process).  Rank g passes f0[first_index, first_index + n_local) with n_local = n0 / G; folds (fri.rs:95-100),
 * leaf hashes (fri.rs:38-44, merkle/src/lib.rs:382-400) and subtrees (merkle/src/lib.rs:164-181, DS positions offset by the
 * slice start) are local; the host all-gathers each layer's cap nodes and finishes the roots. */
typedef struct smw_fri_shard smw_fri_shard;
int32_t smw_fri_shard_build(smw_ctx* ctx, const void* f0_slice, size_t n_local, size_t n0, size_t first_index, const uint64_t* schedule, int32_t L,
                            uint64_t seed_z, int32_t mem, smw_fri_shard** out);
int32_t smw_fri_shard_num_local_layers(const smw_fri_shard* s);      /* K: layers [0, K) are committed from slices */
int32_t smw_fri_shard_cap_info(const smw_fri_shard* s, int32_t l, size_t* n_nodes_local, int32_t* cap_level, int32_t* arity);
int32_t smw_fri_shard_cap(smw_fri_shard* s, int32_t l, void* out, int32_t mem);   /* this rank's nodes of the cap level of layer l */
size_t smw_fri_shard_tail_len(const smw_fri_shard* s);               /* length of this rank's slice of f_K (0 if every layer was local) */
int32_t smw_fri_shard_tail(smw_fri_shard* s, void* out, int32_t mem);
void smw_fri_shard_free(smw_fri_shard* s);
/* root of the tree whose level-`level` nodes are given (level 0 = leaves / leaf digests): finishes a gathered cap */
int32_t smw_merkle_root_from_level(smw_ctx* ctx, smw_params* p, int32_t arity, uint64_t tree_label, int32_t level, const void* nodes, size_t n, int32_t mem, void* root_out_fr_host);

/* ---- Merkle-folded sum-check commit loop (channel/src/lib.rs:406-416, 592-598, 600-657) ---- */
int32_t smw_sumcheck_round_coeffs(smw_ctx* ctx, const void* layer, size_t n, void* c0_host, void* c1_host, int32_t mem);
/* NO COUNTERPART IN THE REFERENCE.  smw_sumcheck_round_coeffs for callers that may
 * not allocate (CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc): `layer_dev` (n elements, n even,
 * 0 < n <= 2^26) and `c0c1_out_dev` (64 bytes: c0 then c1) are DEVICE pointers; the per-block partial sums live in a fixed 64 KiB
 * device array inside the library (held by every GPU that loads it until the process exits).  Calls no cudaMalloc / cudaFree;
 * enqueues two kernels on the context's stream and returns without synchronising.  Larger n is rejected with SMW_ERR_INVALID_ARG.
 * Every call on a device shares the array: calls that overlap on the GPU race.  The first call per device looks the array's address
 * up (cudaGetSymbolAddress): make it once before capturing a graph. */
int32_t smw_sumcheck_round_coeffs_noalloc(smw_ctx* ctx, const void* layer_dev, size_t n, void* c0c1_out_dev);
int32_t smw_mf_round_challenge_from_root(smw_ctx* ctx, uint64_t round_idx, const void* prev_root_host, smw_params* tr_params_or_null, void* out_fr_host);
int32_t smw_mf_fold(smw_ctx* ctx, const void* cur, size_t n, const void* r_fr_host, void* next, int32_t mem);
int32_t smw_mf_commit_phase(smw_ctx* ctx, uint64_t ds_tag, const void* table, size_t n, int32_t mem, void* roots_out_host /* k+1 */, void* r_out_host /* k */, void* final_out_host);

/* ---- legacy trees (merkle/src/lib.rs:196-223 new_legacy, :448-470 new_pairs_legacy, :487-584 verify_many, :704-721 verify_pairs_legacy):
 * fixed t = 17, chunks of poseidon::RATE = 16, every node / pair leaf = hash_with_ds(children, ds_tag) (poseidon/src/lib.rs:85-100).
 * params NULL = merkle::default_params (:364-370).  The handle is an ordinary tree of arity 16: smw_tree_root / _level / _open_many apply. */
int32_t smw_merkle_new_legacy(smw_ctx* ctx, smw_params* p_or_null, const void* ds_tag_fr_host, const void* leaves, size_t n, int32_t mem, smw_tree** out);
int32_t smw_merkle_new_pairs_legacy(smw_ctx* ctx, smw_params* p_or_null, const void* ds_tag_fr_host, const void* f, const void* cp, size_t n_f, size_t n_cp, int32_t mem, smw_tree** out);
int32_t smw_verify_many_legacy(smw_ctx* ctx, smw_params* p_or_null, int32_t proof_arity, const void* ds_tag_fr_host, const void* root_fr_host, const uint64_t* indices, const void* values_host, size_t n,
                               const uint64_t* proof_indices, size_t n_proof_indices, int32_t n_levels, const size_t* n_siblings_per_level, const void* siblings_host,
                               const size_t* n_groups_per_level, const uint8_t* group_sizes, int32_t* ok_out);
int32_t smw_verify_pairs_legacy(smw_ctx* ctx, smw_params* p_or_null, int32_t proof_arity, const void* ds_tag_fr_host, const void* root_fr_host, const uint64_t* indices, const void* pairs_host, size_t n,
                                const uint64_t* proof_indices, size_t n_proof_indices, int32_t n_levels, const size_t* n_siblings_per_level, const void* siblings_host,
                                const size_t* n_groups_per_level, const uint8_t* group_sizes, int32_t* ok_out);

/* ---- multi-GPU data plane: one process per GPU, NCCL bound at run time -----------------------------------
 * The reference has no distributed path; these entry points shard the SAME computations — results are
 * bit-identical to the single-GPU calls above.  Bootstrap: rank 0 obtains a 128-byte id, the host carries it to every rank
 * (any transport), every rank calls smw_comm_init with the context of its GPU.  All exchanges run on the context's stream
 * on device buffers; a NULL communicator means one rank. */
typedef struct smw_comm smw_comm;
int32_t smw_comm_unique_id(uint8_t* out128);
int32_t smw_comm_init(smw_ctx* ctx, const uint8_t* id128, int32_t world, int32_t rank, smw_comm** out);
int32_t smw_comm_rank(const smw_comm* m);
int32_t smw_comm_world(const smw_comm* m);
void smw_comm_free(smw_comm* m);
int32_t smw_comm_all_gather(smw_comm* m, const void* send_dev, void* recv_dev, size_t n_fr);   /* e.g. the per-column roots of a column-sharded commit */
/* fri_build_transcript's layer roots (deep_ali/src/fri.rs:231-312) for ONE polynomial of n0 evaluations held as world * n_chunks
 * equal contiguous slices, n_chunks per rank (chunk c of this rank = f0[first_indices[c], first_indices[c] + n_chunk)).  Every rank
 * folds, hashes and builds the subtrees of its own slices (DS positions offset by the slice start, merkle/src/lib.rs:168-173), ONE
 * all-gather carries every layer's cap and the first layer too small to split, every rank finishes the tops.  roots: (L + 1) Fr,
 * identical on every rank and to smw_fri_build_transcript + smw_fri_roots. */
int32_t smw_fri_roots_sharded(smw_ctx* ctx, smw_comm* comm_or_null, const void* chunks, size_t n_chunk, int32_t n_chunks, const uint64_t* first_indices,
                              size_t n0, const uint64_t* schedule, int32_t L, uint64_t seed_z, int32_t mem, void* roots_out_host);
/* Forward NTT (fft/src/lib.rs:6-12 convention) of ONE vector of 2^log_n elements spread CYCLICALLY over the ranks
 * (in[j] = x[rank + world * j], j < M = 2^log_n / world; device buffers; `in` is overwritten): local NTT, twiddle, ONE all-to-all,
 * a length-world DFT across the received pieces.  out[k1 * (M / world) + j] = X[M * k1 + rank * (M / world) + j]: `world` contiguous chunks,
 * the layout smw_fri_roots_sharded takes. */
int32_t smw_ntt_cyclic_to_chunks(smw_ctx* ctx, smw_comm* comm_or_null, void* in_dev, void* out_dev, int32_t log_n);
/* smw_mf_commit_phase for a table held as `world` equal contiguous slices (this rank: table[rank * n_local, (rank + 1) * n_local)):
 * every round commits the local slice up to its cap, ONE all-gather of the caps per round (r_i needs the root of round i,
 * channel/src/lib.rs:643-646; the root is then on every rank, so r_i needs no broadcast), local fold; once a slice is a single
 * element the layer is gathered and the remaining rounds run replicated.  Outputs as smw_mf_commit_phase, identical on every rank. */
int32_t smw_mf_commit_phase_sharded(smw_ctx* ctx, smw_comm* comm_or_null, uint64_t ds_tag, const void* table_slice, size_t n_local, int32_t mem,
                                    void* roots_out_host, void* r_out_host, void* final_out_host);
uint64_t smw_ctx_collective_count(const smw_ctx* ctx);   /* NCCL collectives issued through this context */
/* Tuning: launches of at least `n` hashes of one shape -- full arity-16 nodes (t = 17), full arity-2 nodes (t = 9), pair leaves
 * (t = 9, 17) or FRI hashed leaves (t = 17) -- run their linear layers on the INT8 tensor cores (csrc/poseidon_tc.cuh; same
 * digests).  0 = never.  Default 16384, or SMW_TC_MIN from the environment. */
int32_t smw_ctx_set_tc_min_nodes(smw_ctx* ctx, uint64_t n);
/* What one node (t = 17: arity 16, two permutations; t = 9: arity 2, one) costs on that path: out6 = {FP64-pipe instructions, int8 MACs on the tensor cores, word products,
 * kernel launches per batch, partial-round group size, the context's current threshold}; zeros 0..4 when the path is unavailable. */
int32_t smw_diag_tc_node_counts(const smw_ctx* ctx, const smw_params* params, uint64_t* out6);

/* ---- "next" tier: openings, verification, DEEP-ALI merge ---------------------------------- */
int32_t smw_index_from_seed(smw_ctx* ctx, const void* seeds_fr, size_t n, uint64_t n_pow2, uint64_t* out_idx_host, int32_t mem);   /* index_from_seed, deep_ali/src/fri.rs:178-183, element-wise */
int32_t smw_ali_sample_z_beta(smw_ctx* ctx, const void* fused_fr_host, const void* roots_seed_fr_host, uint64_t n0, void* z_out_fr_host, void* beta_out_fr_host); /* the ChaCha12 part of ali_sample_z_beta_fs, fri.rs:515-533 */
int32_t smw_tree_open_many(smw_tree* t, const uint64_t* indices, size_t n, smw_proof** out);   /* open_union_of_paths, merkle/src/lib.rs:246-315 */
/* verify_many_ds / verify_pairs_ds (merkle/src/lib.rs:587-701, 725-773).  The MerkleProof (merkle/src/lib.rs:131-143) is passed by its fields:
 * proof_indices[n_pidx]; for level l < n_levels, n_sib[l] siblings and n_grp[l] group sizes, levels consecutive in `siblings_host` / `group_sizes`;
 * arity = proof.arity.  values_host: n Fr; pairs_host: n x (f, cp).  *ok_out = 1 / 0 where the reference returns true / false (malformed proofs
 * are `false`, not errors); the hashes of every recomputed parent (and pair-leaf digest) run on the device, one launch per level. */
int32_t smw_verify_many_ds(smw_ctx* ctx, smw_params* p, int32_t arity, uint64_t tree_label, const void* root_fr_host, const uint64_t* indices, const void* values_host, size_t n,
                           const uint64_t* proof_indices, size_t n_pidx, int32_t n_levels, const size_t* n_sib, const void* siblings_host,
                           const size_t* n_grp, const uint8_t* group_sizes, int32_t* ok_out);
int32_t smw_verify_pairs_ds(smw_ctx* ctx, smw_params* p, int32_t arity, uint64_t tree_label, const void* root_fr_host, const uint64_t* indices, const void* pairs_host, size_t n,
                            const uint64_t* proof_indices, size_t n_pidx, int32_t n_levels, const size_t* n_sib, const void* siblings_host,
                            const size_t* n_grp, const uint8_t* group_sizes, int32_t* ok_out);
int32_t smw_proof_num_levels(const smw_proof* p);
size_t smw_proof_num_indices(const smw_proof* p);
int32_t smw_proof_indices(const smw_proof* p, uint64_t* out);
size_t smw_proof_level_siblings(const smw_proof* p, int32_t level);
size_t smw_proof_level_groups(const smw_proof* p, int32_t level);
int32_t smw_proof_level(const smw_proof* p, int32_t level, void* siblings_out_host, uint8_t* group_sizes_out);
void smw_proof_free(smw_proof* p);
int32_t smw_deep_ali_merge_evals(smw_ctx* ctx, const void* a, const void* s, const void* e, const void* t, size_t n, const void* omega_fr_host, const void* z_fr_host, void* f0_out, int32_t mem); /* deep_ali/src/lib.rs:60-105 (f0 only; c_star is discarded by build_f0) */

#if defined(__GNUC__)
#pragma GCC visibility pop
#endif
#ifdef __cplusplus
}
#endif
#endif
