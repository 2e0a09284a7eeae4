// poseidon_T.cu — compiled once per Poseidon width (-DSMW_T=9|17|33|65|129) so the widths build in
// parallel.  Holds the launchers for the width-templated kernels; the t = 17 unit also holds the
// transcript / FRI-leaf kernels, which are fixed at t = 17 (transcript/src/lib.rs:3).
#include "poseidon_kernels.cuh"
#include "chacha.cuh"
#if SMW_T == 9 || SMW_T == 17
#include "poseidon_coop_kernels.cuh"
#endif
#include <cstdlib>
#ifndef SMW_T
#error "compile with -DSMW_T=<width>"
#endif

namespace smw {

constexpr int T = SMW_T;

// threads per CTA actually launched: the kernels' __launch_bounds__ give the maximum (and the register budget); small CTAs
// retire and refill independently, which measurably shortens the levels (profiles/r01r_cta_size_sweep.txt), so the
// defaults are one or two warps per CTA.  SMW_NT52 / SMW_NT29 override them for tuning runs.
static int launch_threads(int deflt, int max_nt, const char* env) {
    if (const char* e = getenv(env)) { int v = atoi(e); if (v >= 32 && v <= max_nt && v % 32 == 0) return v; }
    return deflt;
}

// Launches of at most this many nodes take the cooperative kernel (one lane group per node, poseidon_coop52.cuh): below
// one wave the thread-per-node kernel costs a lone warp's latency (3.2 ms per t = 17 node, 148 x 32 nodes or 1), the
// cooperative one ~0.4 ms per 592 x (32 / G) nodes.  SMW_COOP_MAX overrides (0 = never).
static uint64_t coop_max_nodes() {
    static long v = -1;
    if (v < 0) { const char* e = getenv("SMW_COOP_MAX"); v = e ? atol(e) : (T == 17 ? 4096 : 2048); }
    return (uint64_t)v;
}
#if SMW_T == 9 || SMW_T == 17
static int launch_nodes_coop(smw_ctx* c, const smw_params* p, const LevelBatch& B) {
    constexpr int NG = 32 / CoopCfg<T>::G;
    { LaunchScope ls(c, KC_MERKLE_LEVEL);
      k_nodes_coop52<T><<<(unsigned)((B.total + NG - 1) / NG), 32, 0, c->stream>>>(B, p->d_tables52, coop_tab(p)); }
    SMW_CUDA(c, cudaGetLastError());
    return SMW_OK;
}
#endif

// segment list [load lanes][permutation][lanes -> external scale] of the batched permutation
static int io_segs(smw_ctx* c, smw_params* p, bool sparse, Segs* sg) {
    Segs io; int rc = get_sponge_segs(c, p, -1, 0, &io); if (rc) return rc;      // key (-1, 0): first[0] = load ops, first[1] = output ops
    sg->n = 3;
    sg->first[0] = io.first[0]; sg->count[0] = p->t;
    sg->first[1] = sparse ? p->perm_first : p->dense_first; sg->count[1] = sparse ? p->perm_count : p->dense_count;
    sg->first[2] = io.first[1]; sg->count[2] = p->t;
    return SMW_OK;
}
static int out_first_op(smw_ctx* c, smw_params* p, int* out_first) {
    Segs io; int rc = get_sponge_segs(c, p, -1, 0, &io); if (rc) return rc;
    *out_first = io.first[1]; return SMW_OK;
}

template <> int permute_batch_T<T>(smw_ctx* c, const smw_params* cp, uint32_t* d_states, size_t batch, bool sparse) {
    smw_params* p = const_cast<smw_params*>(cp);
#if SMW_T == 9 || SMW_T == 17
    if (sparse && c->use_f52 && p->d_tables52 && (uint64_t)batch <= coop_max_nodes() * (T == 17 ? 2 : 1)) {   // small batch: the latency form (a node is two permutations at t = 17)
        constexpr int NG = 32 / CoopCfg<T>::G;
        { LaunchScope ls(c, KC_PERMUTE); k_permute_coop52<T><<<(unsigned)((batch + NG - 1) / NG), 32, 0, c->stream>>>(d_states, batch, p->d_tables52, coop_tab(p)); }
        SMW_CUDA(c, cudaGetLastError());
        return SMW_OK;
    }
#endif
    Segs sg; int rc = io_segs(c, p, sparse, &sg); if (rc) return rc;
    PoseidonDev P = dev_view(p);
#if SMW_T == 9 || SMW_T == 17
    if (c->use_f52 && p->d_tables52) {
        const int nt = launch_threads(Cfg52<T>::RUN, Cfg52<T>::NT, "SMW_NT52"); size_t sm = state_smem_bytes52<T>(nt); unsigned grid = (unsigned)((batch + nt - 1) / nt);
        rc = set_smem(c, k_permute_batch52<T>, sm); if (rc) return rc;
        { LaunchScope ls(c, KC_PERMUTE); k_permute_batch52<T><<<grid, nt, sm, c->stream>>>(d_states, batch, sg, P); }
        SMW_CUDA(c, cudaGetLastError());
        return SMW_OK;
    }
#endif
    const int nt = launch_threads(Cfg<T>::RUN, Cfg<T>::NT, "SMW_NT29"); size_t sm = state_smem_bytes<T>(nt); unsigned grid = (unsigned)((batch + nt - 1) / nt);
    rc = set_smem(c, k_permute_batch<T>, sm); if (rc) return rc;
    { LaunchScope ls(c, KC_PERMUTE); k_permute_batch<T><<<grid, nt, sm, c->stream>>>(d_states, batch, sg, P); }
    SMW_CUDA(c, cudaGetLastError());
    return SMW_OK;
}
template <> int merkle_level_T<T>(smw_ctx* c, const smw_params* cp, const uint32_t* d_children, size_t n_children, uint32_t* d_parents, size_t n_parents,
                                  int arity, uint32_t level, uint64_t label0, uint64_t label_stride, uint64_t pos_offset,
                                  int n_cols, size_t child_col_stride, size_t parent_col_stride) {
    smw_params* p = const_cast<smw_params*>(cp);
    Segs sg; int rc = get_sponge_segs(c, p, 4, arity, &sg); if (rc) return rc;
    int out_first; rc = out_first_op(c, p, &out_first); if (rc) return rc;
    PoseidonDev P = dev_view(p);
#if SMW_T == 17 || SMW_T == 9
    // full arity-16 (t = 17) / arity-2 (t = 9) levels above the threshold: linear layers on the INT8 tensor cores (poseidon_tc.cuh)
    if (c->use_f52 && tc::usable(p) && arity == (T == 17 ? 16 : 2) && n_children == n_parents * (size_t)arity && (uint64_t)n_parents * (uint64_t)n_cols >= c->tc_min_nodes) {
        LevelBatch B{}; B.segs = nullptr; B.n_segs = 0; B.n_cols = n_cols; B.label_stride = label_stride; B.child_col_stride = child_col_stride; B.parent_col_stride = parent_col_stride;
        B.u = NodeSeg{d_children, d_parents, (uint64_t)n_children, (uint64_t)n_parents, label0, pos_offset, level, (uint32_t)arity, 0};
        B.total = (uint64_t)n_parents * (uint64_t)n_cols;
        return tc::merkle_level_tc(c, p, B);
    }
#endif
#if SMW_T == 9 || SMW_T == 17
    if (c->use_f52 && p->d_tables52 && (uint64_t)n_parents * (uint64_t)n_cols <= coop_max_nodes()) {
        LevelBatch B{}; B.segs = nullptr; B.n_segs = 0; B.n_cols = n_cols; B.label_stride = label_stride; B.child_col_stride = child_col_stride; B.parent_col_stride = parent_col_stride;
        B.u = NodeSeg{d_children, d_parents, (uint64_t)n_children, (uint64_t)n_parents, label0, pos_offset, level, (uint32_t)arity, 0};
        B.total = (uint64_t)n_parents * (uint64_t)n_cols;
        return launch_nodes_coop(c, p, B);
    }
    if (c->use_f52 && p->d_tables52) {
        const int nt = launch_threads(Cfg52<T>::RUN, Cfg52<T>::NT, "SMW_NT52");
        size_t sm = state_smem_bytes52<T>(nt);
        rc = set_smem(c, k_merkle_level52<T>, sm); if (rc) return rc;
        dim3 grid((unsigned)((n_parents + nt - 1) / nt), (unsigned)n_cols);
        { LaunchScope ls(c, KC_MERKLE_LEVEL);
          k_merkle_level52<T><<<grid, nt, sm, c->stream>>>(d_children, n_children, d_parents, n_parents, arity, level, label0, label_stride, pos_offset,
                                                           child_col_stride, parent_col_stride, sg, out_first, P); }
        SMW_CUDA(c, cudaGetLastError());
        return SMW_OK;
    }
#endif
    const int nt = launch_threads(Cfg<T>::RUN, Cfg<T>::NT, "SMW_NT29");
    size_t sm = state_smem_bytes<T>(nt);
    rc = set_smem(c, k_merkle_level<T>, sm); if (rc) return rc;
    dim3 grid((unsigned)((n_parents + nt - 1) / nt), (unsigned)n_cols);
    { LaunchScope ls(c, KC_MERKLE_LEVEL);
      k_merkle_level<T><<<grid, nt, sm, c->stream>>>(d_children, n_children, d_parents, n_parents, arity, level, label0, label_stride, pos_offset,
                                                     child_col_stride, parent_col_stride, sg, out_first, P); }
    SMW_CUDA(c, cudaGetLastError());
    return SMW_OK;
}
// several runs (explicit device segment array) of one level in one launch; widths without FP64-pipe tables take one launch per run
template <> int merkle_level_multi_T<T>(smw_ctx* c, const smw_params* cp, const NodeSeg* d_segs, int n_segs, uint64_t total, int arity, bool all_full) {
    smw_params* p = const_cast<smw_params*>(cp);
    if (n_segs <= 0 || total == 0) return SMW_OK;
#if SMW_T == 17 || SMW_T == 9
    if (c->use_f52 && tc::usable(p) && arity == (T == 17 ? 16 : 2) && all_full && total >= c->tc_min_nodes) {
        LevelBatch B{}; B.segs = d_segs; B.n_segs = n_segs; B.total = total;
        return tc::merkle_level_tc(c, p, B);
    }
#endif
#if SMW_T == 9 || SMW_T == 17
    if (c->use_f52 && p->d_tables52) {
        if (total <= coop_max_nodes()) {
            LevelBatch B{}; B.segs = d_segs; B.n_segs = n_segs; B.total = total;
            return launch_nodes_coop(c, p, B);
        }
        Segs sg; int rc = get_sponge_segs(c, p, 4, arity, &sg); if (rc) return rc;
        int out_first; rc = out_first_op(c, p, &out_first); if (rc) return rc;
        const int nt = launch_threads(Cfg52<T>::RUN, Cfg52<T>::NT, "SMW_NT52"); size_t sm = state_smem_bytes52<T>(nt);
        rc = set_smem(c, k_merkle_level52_multi<T>, sm); if (rc) return rc;
        { LaunchScope ls(c, KC_MERKLE_LEVEL);
          k_merkle_level52_multi<T><<<(unsigned)((total + nt - 1) / nt), nt, sm, c->stream>>>(d_segs, n_segs, total, sg, out_first, dev_view(p)); }
        SMW_CUDA(c, cudaGetLastError());
        return SMW_OK;
    }
#endif
    return set_err(c, SMW_ERR_UNSUPPORTED_WIDTH, "multi-run level launch needs the FP64-pipe tables (t = 9, 17)");
}
template <> int pair_leaves_multi_T<T>(smw_ctx* c, const smw_params* cp, const LeafSeg* d_segs, int n_segs, uint64_t total) {
    smw_params* p = const_cast<smw_params*>(cp);
    if (n_segs <= 0 || total == 0) return SMW_OK;
#if SMW_T == 17 || SMW_T == 9
    if (c->use_f52 && tc::usable(p) && total >= c->tc_min_nodes) return tc::pair_leaves_tc(c, p, d_segs, n_segs, LeafSeg{}, total);
#endif
#if SMW_T == 9 || SMW_T == 17
    if (c->use_f52 && p->d_tables52) {
        Segs sg; int rc = get_sponge_segs(c, p, 4, 2, &sg); if (rc) return rc;
        int out_first; rc = out_first_op(c, p, &out_first); if (rc) return rc;
        const int nt = launch_threads(Cfg52<T>::RUN, Cfg52<T>::NT, "SMW_NT52"); size_t sm = state_smem_bytes52<T>(nt);
        rc = set_smem(c, k_pair_leaves52_multi<T>, sm); if (rc) return rc;
        { LaunchScope ls(c, KC_PAIR_LEAVES);
          k_pair_leaves52_multi<T><<<(unsigned)((total + nt - 1) / nt), nt, sm, c->stream>>>(d_segs, n_segs, total, sg, out_first, dev_view(p)); }
        SMW_CUDA(c, cudaGetLastError());
        return SMW_OK;
    }
#endif
    return set_err(c, SMW_ERR_UNSUPPORTED_WIDTH, "multi-run leaf launch needs the FP64-pipe tables (t = 9, 17)");
}
template <> int hash_groups_T<T>(smw_ctx* c, const smw_params* cp, const uint32_t* d_pool, const uint64_t* d_child_idx, const uint8_t* d_cnt, const uint64_t* d_pos,
                                 size_t n_groups, int arity, int row, uint32_t level, uint64_t label, uint32_t* d_out) {
    smw_params* p = const_cast<smw_params*>(cp);
    Segs sg; int rc = get_sponge_segs(c, p, 4, row, &sg); if (rc) return rc;
    int out_first; rc = out_first_op(c, p, &out_first); if (rc) return rc;
    PoseidonDev P = dev_view(p);
    const int nt = launch_threads(Cfg<T>::RUN, Cfg<T>::NT, "SMW_NT29"); size_t sm = state_smem_bytes<T>(nt);
    rc = set_smem(c, k_hash_groups<T>, sm); if (rc) return rc;
    { LaunchScope ls(c, KC_MERKLE_LEVEL);
      k_hash_groups<T><<<(unsigned)((n_groups + nt - 1) / nt), nt, sm, c->stream>>>(d_pool, d_child_idx, d_cnt, d_pos, n_groups, arity, row, level, label, d_out, sg, out_first, P); }
    SMW_CUDA(c, cudaGetLastError());
    return SMW_OK;
}
template <> int pair_leaves_T<T>(smw_ctx* c, const smw_params* cp, const uint32_t* d_f, const uint32_t* d_cp, size_t n, int arity, uint64_t label, uint64_t pos_offset,
                                 uint32_t* d_out, size_t cp_div) {
    smw_params* p = const_cast<smw_params*>(cp);
#if SMW_T == 17 || SMW_T == 9
    if (c->use_f52 && tc::usable(p) && (uint64_t)n >= c->tc_min_nodes) {      // linear layers on the tensor cores (poseidon_tc.cuh)
        LeafSeg u{d_f, d_cp, d_out, (uint64_t)n, label, pos_offset, (uint64_t)(cp_div ? cp_div : 1), (uint32_t)arity, 0, 0};
        return tc::pair_leaves_tc(c, p, nullptr, 0, u, (uint64_t)n);
    }
#endif
    Segs sg; int rc = get_sponge_segs(c, p, 4, 2, &sg); if (rc) return rc;
    int out_first; rc = out_first_op(c, p, &out_first); if (rc) return rc;
    PoseidonDev P = dev_view(p);
#if SMW_T == 9 || SMW_T == 17
    if (c->use_f52 && p->d_tables52) {
        const int nt = launch_threads(Cfg52<T>::RUN, Cfg52<T>::NT, "SMW_NT52"); size_t sm = state_smem_bytes52<T>(nt);
        rc = set_smem(c, k_pair_leaves52<T>, sm); if (rc) return rc;
        { LaunchScope ls(c, KC_PAIR_LEAVES);
          k_pair_leaves52<T><<<(unsigned)((n + nt - 1) / nt), nt, sm, c->stream>>>(d_f, d_cp, n, arity, label, pos_offset, d_out, cp_div, sg, out_first, P); }
        SMW_CUDA(c, cudaGetLastError());
        return SMW_OK;
    }
#endif
    const int nt = launch_threads(Cfg<T>::RUN, Cfg<T>::NT, "SMW_NT29"); size_t sm = state_smem_bytes<T>(nt);
    rc = set_smem(c, k_pair_leaves<T>, sm); if (rc) return rc;
    { LaunchScope ls(c, KC_PAIR_LEAVES);
      k_pair_leaves<T><<<(unsigned)((n + nt - 1) / nt), nt, sm, c->stream>>>(d_f, d_cp, n, arity, label, pos_offset, d_out, cp_div, sg, out_first, P); }
    SMW_CUDA(c, cudaGetLastError());
    return SMW_OK;
}
template <> int hash_ds_single_T<T>(smw_ctx* c, const smw_params* p, const uint32_t* d_ds, int nds, const uint32_t* d_in, size_t n, uint32_t* d_out) {
    size_t sm = state_smem_bytes<T>(1) + 64;
    int rc = set_smem(c, k_hash_ds_single<T>, sm); if (rc) return rc;
    k_hash_ds_single<T><<<1, 32, sm, c->stream>>>(d_ds, nds, d_in, n, d_out, dev_view(p));
    c->launches++;
    SMW_CUDA(c, cudaGetLastError());
    return SMW_OK;
}

#if SMW_T == 17
// legacy sponge (poseidon/src/lib.rs:85-100): state[16] = ds_tag, 16-chunks, no padding; t = 17
__global__ void k_hash_legacy_single(const uint32_t* in, size_t n, const uint32_t* tag, uint32_t* out, PoseidonDev P) {
    extern __shared__ __align__(16) unsigned char smem[];
    if (threadIdx.x != 0 || blockIdx.x != 0) return;
    StateRef S = state_ref<17>(smem, 1, 0);
    uint32_t w[8];
    for (int i = 0; i < 16; i++) S.store(i, f29_zero());
    load_ext(tag, w); S.store(16, f29_from_ext(w));
    for (size_t off = 0; off < n; off += 16) {
        size_t c = n - off < 16 ? n - off : 16;
        for (size_t i = 0; i < c; i++) { load_ext(in + (off + i) * 8, w); S.store((int)i, f29_add(S.load((int)i), f29_from_ext(w))); }
        run_perm<17>(S.nt, P);
    }
    f29_to_ext(S.load(0), w); store_ext(out, w);
}

// ------------------------------------------------------------------ transcript (transcript/src/lib.rs:55-101), t = 17
// One launch runs a whole absorb program on the device-resident state:
//   [fresh: state = 0, lane 16 = INIT tag (immA[0])] ; absorb immA[fresh..nA) ; absorb extB[0..nB) ; absorb immC[0..nC) ; [final permute]
// Immediates are 256-bit little-endian integers (tags, 31-byte words, small integers), reduced mod p here.
SMW_DEV void tr_absorb(const StateRef& S, const PoseidonDev& P, int& pos, const F29& x) {
    if (pos == 16) { run_perm<17>(S.nt, P); pos = 0; }
    S.store(pos, f29_add(S.load(pos), x)); pos++;
}
__global__ void k_transcript_run(uint32_t* state, int fresh, int pos, const uint32_t* immA, int nA, const uint32_t* extB, size_t nB,
                                 const uint32_t* immC, int nC, int final_permute, uint32_t* out, PoseidonDev P) {
    extern __shared__ __align__(16) unsigned char smem[];
    if (threadIdx.x != 0 || blockIdx.x != 0) return;
    StateRef S = state_ref<17>(smem, 1, 0);
    uint32_t w[8];
    int a0 = 0;
    if (fresh) {
        for (int i = 0; i < 16; i++) S.store(i, f29_zero());
        load_ext(immA, w); S.store(16, f29_from_le256(w)); a0 = 1; pos = 0;
    } else {
        for (int i = 0; i < 17; i++) { load_ext(state + i * 8, w); S.store(i, f29_from_ext(w)); }
    }
    for (int i = a0; i < nA; i++) { load_ext(immA + i * 8, w); tr_absorb(S, P, pos, f29_from_le256(w)); }
    for (size_t i = 0; i < nB; i++) { load_ext(extB + i * 8, w); tr_absorb(S, P, pos, f29_from_ext(w)); }
    for (int i = 0; i < nC; i++) { load_ext(immC + i * 8, w); tr_absorb(S, P, pos, f29_from_le256(w)); }
    if (final_permute) run_perm<17>(S.nt, P);
    if (state) for (int i = 0; i < 17; i++) { f29_to_ext(S.load(i), w); store_ext(state + i * 8, w); }
    if (out) { f29_to_ext(S.load(0), w); store_ext(out, w); }
}

// ------------------------------------------------------------------ bulk absorb, one warp per sponge
// The transcript sponge over n elements (tr_hash_fields_tagged over a whole evaluation vector, deep_ali/src/fri.rs:28-36, 535-560)
// is strictly serial: n/16 permutations one after the other.  On one thread a permutation takes ~2.5 ms; here the warp runs the
// LITERAL round structure (poseidon/src/lib.rs:219-258: ARK, S-box, dense 17x17 mat-vec) with one state lane / output row per
// thread: 17 lazily accumulated products + one reduction per lane and round, S-boxes lane-parallel.  Field results are exact, so
// the state after the absorb is bit-identical to the serial kernel's.
struct CoopTables { uint32_t mds, rc_full, rc_part, k_one; int rp; };
SMW_DEV void coop_matvec17(F29* st, const uint32_t* tbl, const CoopTables& L, int lane, bool add_rc, uint32_t rc_idx) {
    F29 y = f29_zero();
    if (lane < 17) {
        Acc A; acc_zero(A);
        F29 b = tbl_load(tbl, L.mds + (uint32_t)lane * 17);
#pragma unroll 1
        for (int j = 0; j < 17; j++) {
            const F29 bb = b;
            if (j + 1 < 17) b = tbl_load(tbl, L.mds + (uint32_t)lane * 17 + j + 1);
            if (j == 6 || j == 12) acc_normalize(A);
            acc_mac(A, st[j], bb);
        }
        F29 rc = f29_zero();
        if (add_rc) rc = tbl_load(tbl, rc_idx);
        y = acc_reduce_add(A, rc);
    }
    __syncwarp();
    if (lane < 17) st[lane] = y;
    __syncwarp();
}
SMW_DEV void coop_perm17(F29* st, const uint32_t* tbl, const CoopTables& L, int lane) {
    if (lane < 17) {           // ARK of round 0 as a product with the Montgomery one (normalises the lazily added input lanes)
        Acc A; acc_zero(A); acc_mac(A, st[lane], tbl_load(tbl, L.k_one));
        st[lane] = acc_reduce_add(A, tbl_load(tbl, L.rc_full + lane));
    }
    __syncwarp();
    for (int r = 0; r < 4; r++) {
        if (lane < 17) st[lane] = f29_sbox5(st[lane]);
        __syncwarp();
        if (r < 3) coop_matvec17(st, tbl, L, lane, true, L.rc_full + (uint32_t)(r + 1) * 17 + lane);
        else coop_matvec17(st, tbl, L, lane, lane == 0, L.rc_part);
    }
    for (int k = 0; k < L.rp; k++) {
        if (lane == 0) st[0] = f29_sbox5(st[0]);
        __syncwarp();
        if (k + 1 < L.rp) coop_matvec17(st, tbl, L, lane, lane == 0, L.rc_part + (uint32_t)(k + 1));
        else coop_matvec17(st, tbl, L, lane, true, L.rc_full + 4u * 17 + lane);
    }
    for (int r = 4; r < 8; r++) {
        if (lane < 17) st[lane] = f29_sbox5(st[lane]);
        __syncwarp();
        coop_matvec17(st, tbl, L, lane, r < 7, L.rc_full + (uint32_t)(r + 1) * 17 + lane);
    }
}
__global__ void __launch_bounds__(32) k_transcript_absorb_coop(uint32_t* state, int pos, const uint32_t* __restrict__ ext, size_t n, const uint32_t* __restrict__ tbl, CoopTables L) {
    __shared__ F29 st[17];
    const int lane = threadIdx.x;
    uint32_t w[8];
    if (lane < 17) { load_ext(state + lane * 8, w); st[lane] = f29_from_ext(w); }
    __syncwarp();
    size_t i = 0;
    while (i < n) {
        if (pos == 16) { coop_perm17(st, tbl, L, lane); pos = 0; }
        const int cnt = (int)((n - i < (size_t)(16 - pos)) ? (n - i) : (size_t)(16 - pos));
        if (lane >= pos && lane < pos + cnt) { load_ext(ext + (i + (size_t)(lane - pos)) * 8, w); st[lane] = f29_add(st[lane], f29_from_ext(w)); }
        __syncwarp();
        pos += cnt; i += (size_t)cnt;
    }
    if (lane < 17) { f29_to_ext(st[lane], w); store_ext(state + lane * 8, w); }
}

// FRI hashed leaves (deep_ali/src/fri.rs:38-44 hash_leaf_pair): the pre-image is a fixed 17-lane template
// (tags and label words; transcript/src/lib.rs absorbs 9 lanes, no intermediate permutation) with f and s
// dropped into two lanes: exactly one permutation per leaf.  s_i = s[i / s_div] (or 0 when s == nullptr).
struct InFriLeaf {
    const uint32_t* f; const uint32_t* s; const uint32_t* tmpl;
    SMW_DEV F29 operator()(uint32_t idx) const {
        if (idx == 0) return raw_ext(f);
        if (idx == 1) return s ? raw_ext(s) : f29_zero();
        return raw_ext(tmpl + (size_t)(idx - 2) * 8);          // template lane idx-2: a 256-bit little-endian integer (tag / label word)
    }
    SMW_DEV F52 f52(uint32_t idx) const {
        if (idx == 0) return raw_ext52(f);
        if (idx == 1) return s ? raw_ext52(s) : f52_zero();
        return raw_ext52(tmpl + (size_t)(idx - 2) * 8);
    }
};
__global__ void __launch_bounds__(Cfg52<17>::NT, Cfg52<17>::MINB)
k_fri_leaf_hashed52(const uint32_t* __restrict__ tmpl_le256, const uint32_t* __restrict__ f, const uint32_t* __restrict__ s,
                    size_t s_div, size_t n, uint32_t* __restrict__ out, Segs sg, int out_first, PoseidonDev P) {
    extern __shared__ __align__(16) unsigned char smem[];
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    StateRef52 S = state_ref52<17>(smem, blockDim.x, threadIdx.x);
    InFriLeaf in{f + i * 8, s ? s + (i / s_div) * 8 : nullptr, tmpl_le256};
    run_segs52<17>(blockDim.x, P, sg, in);
    run_ops52<17>(blockDim.x, P.tbl52, P.prog, out_first, 1, in);
    emit_ext52(S.load(0), out + i * 8);
}
__global__ void __launch_bounds__(Cfg52<17>::NT, Cfg52<17>::MINB)
k_fri_leaf_hashed52_multi(const uint32_t* __restrict__ tmpl_le256, const LeafSeg* __restrict__ segs, int n_segs, uint64_t total, Segs sg, int out_first, PoseidonDev P) {
    extern __shared__ __align__(16) unsigned char smem[];
    const uint64_t g = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (g >= total) return;
    const LeafSeg ls = segs[find_seg(g, &segs[0].first, sizeof(LeafSeg) / 8, n_segs)];
    const uint64_t i = g - ls.first;
    StateRef52 S = state_ref52<17>(smem, blockDim.x, threadIdx.x);
    InFriLeaf in{ls.f + i * 8, ls.cp ? ls.cp + (i / ls.cp_div) * 8 : nullptr, tmpl_le256};
    run_segs52<17>(blockDim.x, P, sg, in);
    run_ops52<17>(blockDim.x, P.tbl52, P.prog, out_first, 1, in);
    emit_ext52(S.load(0), ls.out + i * 8);
}
__global__ void __launch_bounds__(Cfg<17>::NT, Cfg<17>::MINB)
k_fri_leaf_hashed(const uint32_t* __restrict__ tmpl_le256, const uint32_t* __restrict__ f, const uint32_t* __restrict__ s,
                  size_t s_div, size_t n, uint32_t* __restrict__ out, Segs sg, int out_first, PoseidonDev P) {
    extern __shared__ __align__(16) unsigned char smem[];
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    StateRef S = state_ref<17>(smem, blockDim.x, threadIdx.x);
    InFriLeaf in{f + i * 8, s ? s + (i / s_div) * 8 : nullptr, tmpl_le256};
    run_segs<17>(blockDim.x, P, sg, in);         // [17 lane loads][permutation]
    run_ops<17>(blockDim.x, P.tbl, P.prog, out_first, 1, in);
    emit_ext(S.load(0), out + i * 8);
}

// fri_sample_z_ell (deep_ali/src/fri.rs:59-82) for all L levels at once: thread l hashes its template
// (one permutation), seeds ChaCha12 with the canonical bytes of the digest, and takes the first u64 c with
// c != 0 and c^n != 1.
__global__ void k_fri_sample_z(const uint32_t* tmpl_le256, const uint64_t* level_size /* L x (level, size) */, int L, uint64_t seed_z, uint32_t* z_out, PoseidonDev P) {
    extern __shared__ __align__(16) unsigned char smem[];
    int l = blockIdx.x * blockDim.x + threadIdx.x;
    if (l >= L) return;
    StateRef S = state_ref<17>(smem, blockDim.x, threadIdx.x);
    uint32_t w[8];
    for (int j = 0; j < 17; j++) { load_ext(tmpl_le256 + ((size_t)l * 17 + j) * 8, w); S.store(j, f29_from_le256(w)); }
    run_perm<17>(S.nt, P);
    // canonical integer of the digest: x*2^261 -> x
    F29 one_int = f29_zero(); one_int.v[0] = 1;
    F29 d = f29_canonicalize(f29_mul(S.load(0), one_int));
    uint32_t key[8]; f29_repack_to32(d, key);
    const uint64_t level = level_size[2 * l], n = level_size[2 * l + 1];
    ChaChaRng rng; chacha_init(rng, key);
    F29 z = f29_zero(); bool found = false;
    for (int tries = 0; tries < 1000 && !found; tries++) {
        uint64_t c = chacha_next_u64(rng);
        if (c == 0) continue;
        F29 cand = f29_from_u64(c);
        F29 pw = f29_const(C_ONE29);                 // cand^n by square-and-multiply over the bits of n
        for (int b = 63; b >= 0; b--) { pw = f29_sqr(pw); if ((n >> b) & 1ull) pw = f29_mul(pw, cand); }
        F29 diff = f29_sub_canon(f29_full_reduce(pw), f29_const(C_ONE29));
        if (!f29_is_zero_canon(diff)) { z = cand; found = true; }
    }
    if (!found) {   // fallbacks of fri.rs:75-79
        F29 fb = f29_from_u64(seed_z + level + 7ull);
        F29 pw = f29_const(C_ONE29);
        for (int b = 63; b >= 0; b--) { pw = f29_sqr(pw); if ((n >> b) & 1ull) pw = f29_mul(pw, fb); }
        F29 diff = f29_sub_canon(f29_full_reduce(pw), f29_const(C_ONE29));
        z = f29_is_zero_canon(diff) ? f29_from_u64(11) : fb;
    }
    f29_to_ext(z, w); store_ext(z_out + (size_t)l * 8, w);
}


int launch_hash_legacy_single(smw_ctx* c, const smw_params* p, const uint32_t* d_in, size_t n, const uint32_t* d_tag, uint32_t* d_out) {
    if (p->t != 17) return set_err(c, SMW_ERR_UNSUPPORTED_WIDTH, "legacy hash_with_ds is fixed at t = 17");
    k_hash_legacy_single<<<1, 32, state_smem_bytes<17>(1) + 64, c->stream>>>(d_in, n, d_tag, d_out, dev_view(p));
    c->launches++;
    SMW_CUDA(c, cudaGetLastError());
    return SMW_OK;
}
int launch_transcript_run(smw_ctx* c, const smw_params* p, uint32_t* d_state, int fresh, int pos,
                          const uint32_t* d_immA, int nA, const uint32_t* d_extB, size_t nB, const uint32_t* d_immC, int nC, int final_permute, uint32_t* d_out) {
    if (p->t != 17) return set_err(c, SMW_ERR_UNSUPPORTED_WIDTH, "transcript is fixed at t = 17");
    if (c->use_f52 && p->d_tables52) {       // one warp per sponge (poseidon_coop52.cuh): ~0.2 ms per permutation instead of 2.5 ms on one thread
        k_transcript_coop52<<<1, 32, 0, c->stream>>>(d_state, fresh, pos, d_immA, nA, d_extB, nB, d_immC, nC, final_permute, d_out, p->d_tables52, coop_tab(p));
        c->launches++;
        SMW_CUDA(c, cudaGetLastError());
        return SMW_OK;
    }
    k_transcript_run<<<1, 32, state_smem_bytes<17>(1) + 64, c->stream>>>(d_state, fresh, pos, d_immA, nA, d_extB, nB, d_immC, nC, final_permute, d_out, dev_view(p));
    c->launches++;
    SMW_CUDA(c, cudaGetLastError());
    return SMW_OK;
}
int launch_transcript_absorb_coop(smw_ctx* c, const smw_params* p, uint32_t* d_state, int pos, const uint32_t* d_ext, size_t n) {
    if (p->t != 17) return set_err(c, SMW_ERR_UNSUPPORTED_WIDTH, "transcript is fixed at t = 17");
    if (n == 0) return SMW_OK;
    if (c->use_f52 && p->d_tables52) {
        k_transcript_coop52<<<1, 32, 0, c->stream>>>(d_state, 0, pos, nullptr, 0, d_ext, n, nullptr, 0, 0, nullptr, p->d_tables52, coop_tab(p));
        c->launches++;
        SMW_CUDA(c, cudaGetLastError());
        return SMW_OK;
    }
    CoopTables L{p->layout.mds, p->layout.rc_full, p->layout.rc_part, p->layout.k_one, p->rp};
    k_transcript_absorb_coop<<<1, 32, 0, c->stream>>>(d_state, pos, d_ext, n, p->d_tables, L);
    c->launches++;
    SMW_CUDA(c, cudaGetLastError());
    return SMW_OK;
}
int launch_fri_leaf_hashed(smw_ctx* c, const smw_params* ctrp, const uint32_t* d_tmpl, int lane_f, int lane_s,
                           const uint32_t* d_f, const uint32_t* d_s, size_t s_div, size_t n, uint32_t* d_out) {
    if (n == 0) return SMW_OK;
    smw_params* trp = const_cast<smw_params*>(ctrp);
    if (c->use_f52 && trp->t == 17 && tc::usable(trp) && (uint64_t)n >= c->tc_min_nodes) {      // linear layers on the tensor cores (poseidon_tc.cuh)
        LeafSeg u{d_f, d_s, d_out, (uint64_t)n, 0, 0, (uint64_t)(s_div ? s_div : 1), 0, 0, 0};
        return tc::fri_leaves_tc(c, trp, d_tmpl, lane_f, lane_s, nullptr, 0, u, (uint64_t)n);
    }
    Segs sg; int rc = get_sponge_segs(c, trp, -2, lane_f * 256 + lane_s, &sg); if (rc) return rc;
    int out_first; rc = out_first_op(c, trp, &out_first); if (rc) return rc;
    if (c->use_f52 && trp->d_tables52) {
        const int nt = launch_threads(Cfg52<17>::RUN, Cfg52<17>::NT, "SMW_NT52"); size_t sm = state_smem_bytes52<17>(nt);
        rc = set_smem(c, k_fri_leaf_hashed52, sm); if (rc) return rc;
        { LaunchScope ls(c, KC_FRI_LEAF);
          k_fri_leaf_hashed52<<<(unsigned)((n + nt - 1) / nt), nt, sm, c->stream>>>(d_tmpl, d_f, d_s, s_div, n, d_out, sg, out_first, dev_view(trp)); }
        SMW_CUDA(c, cudaGetLastError());
        return SMW_OK;
    }
    const int nt = launch_threads(Cfg<17>::RUN, Cfg<17>::NT, "SMW_NT29"); size_t sm = state_smem_bytes<17>(nt);
    rc = set_smem(c, k_fri_leaf_hashed, sm); if (rc) return rc;
    { LaunchScope ls(c, KC_FRI_LEAF);
      k_fri_leaf_hashed<<<(unsigned)((n + nt - 1) / nt), nt, sm, c->stream>>>(d_tmpl, d_f, d_s, s_div, n, d_out, sg, out_first, dev_view(trp)); }
    SMW_CUDA(c, cudaGetLastError());
    return SMW_OK;
}
int launch_fri_leaf_hashed_multi(smw_ctx* c, const smw_params* ctrp, const uint32_t* d_tmpl, int lane_f, int lane_s, const LeafSeg* d_segs, int n_segs, uint64_t total) {
    if (n_segs <= 0 || total == 0) return SMW_OK;
    smw_params* trp = const_cast<smw_params*>(ctrp);
    if (!(c->use_f52 && trp->d_tables52)) return set_err(c, SMW_ERR_UNSUPPORTED_WIDTH, "multi-run leaf launch needs the FP64-pipe tables");
    if (c->use_f52 && trp->t == 17 && tc::usable(trp) && total >= c->tc_min_nodes) return tc::fri_leaves_tc(c, trp, d_tmpl, lane_f, lane_s, d_segs, n_segs, LeafSeg{}, total);
    Segs sg; int rc = get_sponge_segs(c, trp, -2, lane_f * 256 + lane_s, &sg); if (rc) return rc;
    int out_first; rc = out_first_op(c, trp, &out_first); if (rc) return rc;
    const int nt = launch_threads(Cfg52<17>::RUN, Cfg52<17>::NT, "SMW_NT52"); size_t sm = state_smem_bytes52<17>(nt);
    rc = set_smem(c, k_fri_leaf_hashed52_multi, sm); if (rc) return rc;
    { LaunchScope ls(c, KC_FRI_LEAF);
      k_fri_leaf_hashed52_multi<<<(unsigned)((total + nt - 1) / nt), nt, sm, c->stream>>>(d_tmpl, d_segs, n_segs, total, sg, out_first, dev_view(trp)); }
    SMW_CUDA(c, cudaGetLastError());
    return SMW_OK;
}
int launch_tr_hash_batch(smw_ctx* c, const smw_params* trp, const uint32_t* d_tmpl, int lane0, int k, const uint32_t* d_xs, size_t count, uint32_t* d_out) {
    if (count == 0) return SMW_OK;
    if (trp->t != 17 || !trp->d_tables52) return set_err(c, SMW_ERR_UNSUPPORTED_WIDTH, "transcript is fixed at t = 17");
    k_tr_hash_batch_coop52<<<(unsigned)count, 32, 0, c->stream>>>(d_tmpl, lane0, k, d_xs, count, d_out, trp->d_tables52, coop_tab(trp));
    c->launches++;
    SMW_CUDA(c, cudaGetLastError());
    return SMW_OK;
}
int launch_legacy_nodes(smw_ctx* c, const smw_params* p, int mode, const uint32_t* d_src, const uint32_t* d_cp, size_t n_src, const uint64_t* d_idx, const uint8_t* d_cnt,
                        size_t n_out, const uint32_t* d_tag, uint32_t* d_out) {
    if (n_out == 0) return SMW_OK;
    if (p->t != 17 || !p->d_tables52) return set_err(c, SMW_ERR_UNSUPPORTED_WIDTH, "legacy trees are fixed at t = 17 (poseidon::RATE = 16)");
    { LaunchScope ls(c, KC_MERKLE_LEVEL);
      k_legacy_nodes_coop52<<<(unsigned)n_out, 32, 0, c->stream>>>(mode, d_src, d_cp, n_src, d_idx, d_cnt, n_out, d_tag, d_out, p->d_tables52, coop_tab(p)); }
    SMW_CUDA(c, cudaGetLastError());
    return SMW_OK;
}
int launch_fri_sample_z(smw_ctx* c, const smw_params* trp, const uint32_t* d_tmpl, const uint64_t* d_sizes, int L, uint64_t seed_z, uint32_t* d_z) {
    if (L <= 0) return SMW_OK;
    if (c->use_f52 && trp->d_tables52) {
        k_fri_sample_z_coop52<<<L, 32, 0, c->stream>>>(d_tmpl, d_sizes, L, seed_z, d_z, trp->d_tables52, coop_tab(trp));
        c->launches++;
        SMW_CUDA(c, cudaGetLastError());
        return SMW_OK;
    }
    const int nt = 32; size_t sm = state_smem_bytes<17>(nt);
    k_fri_sample_z<<<(L + nt - 1) / nt, nt, sm, c->stream>>>(d_tmpl, d_sizes, L, seed_z, d_z, dev_view(trp));
    c->launches++;
    SMW_CUDA(c, cudaGetLastError());
    return SMW_OK;
}
#endif

} // namespace smw
