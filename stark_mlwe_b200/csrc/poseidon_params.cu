// poseidon_params.cu — Poseidon constant derivation on the device (BLAKE3 -> Fr), the LU / sparse
// factorisation of the round matrix, and the width dispatch of the templated launchers.
#include "poseidon_kernels.cuh"
#include "blake3.cuh"
#include <cstring>

namespace smw {

// ------------------------------------------------------------------ constant derivation (BLAKE3)
struct SeedArg { unsigned char b[96]; int len; };
// utils/src/lib.rs:16-22 fr_from_hash + poseidon/src/lib.rs:176-216: one thread per constant
__global__ void k_derive_consts(int t, int rp, SeedArg seed, uint32_t* mds, uint32_t* rcf, uint32_t* rcp) {
    int id = blockIdx.x * blockDim.x + threadIdx.x;
    int n_mds = t * t, n_rcf = 8 * t;
    if (id >= n_mds + n_rcf + rp) return;
    unsigned char msg[160]; int len = 0;
    const char* tag; unsigned long long a, b = 0; int two = 1; uint32_t* dst; int di;
    if (id < n_mds) { tag = "POSEIDON-MDS"; a = id / t; b = id % t; dst = mds; di = id; }
    else if (id < n_mds + n_rcf) { int k = id - n_mds; tag = "POSEIDON-RC-FULL"; a = k / t; b = k % t; dst = rcf; di = k; }
    else { int k = id - n_mds - n_rcf; tag = "POSEIDON-RC-PART"; a = k; two = 0; dst = rcp; di = k; }
    for (const char* q = tag; *q; q++) msg[len++] = (unsigned char)*q;
    for (int i = 0; i < 8; i++) msg[len++] = (unsigned char)(a >> (8 * i));
    if (two) for (int i = 0; i < 8; i++) msg[len++] = (unsigned char)(b >> (8 * i));
    for (int i = 0; i < seed.len; i++) msg[len++] = seed.b[i];
    uint32_t h[8];
    blake3_chunk(msg, len, h);
    gs(dst, di, f29_from_le256(h));   // the 64-byte buffer's upper half is zero: value = LE(hash) mod p
}
// external-form tables supplied by the caller -> internal form
__global__ void k_ext_to_internal(const uint32_t* ext, size_t n, uint32_t* out) {
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    uint32_t w[8]; load_ext(ext + i * 8, w);
    gs(out, i, f29_full_reduce(f29_from_ext(w)));
}
__global__ void k_internal_to_ext(const uint32_t* tbl, size_t n, uint32_t* ext) {
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    uint32_t w[8]; f29_to_ext(gl(tbl, i), w); store_ext(ext + i * 8, w);
}

// ------------------------------------------------------------------ LU / sparse factorisation (one CTA)
// scratch layout (elements of EL words): see launch_finish_params
struct DeriveArgs {
    int t, rp;
    const uint32_t* mds;   // [t][t]
    uint32_t *lu_full, *sp_row, *sp_col, *lu_post, *post;
    uint32_t *mh, *mhi, *aug, *pw, *tmp, *vk, *wk, *vt;   // scratch
    int* status;
};
__device__ void lu_inplace(uint32_t* A, int n, int* status, F29* sh_inv) {
    for (int k = 0; k < n; k++) {
        if (threadIdx.x == 0) {
            F29 piv = gl(A, (size_t)k * n + k);
            if (f29_is_zero_canon(piv)) { *status = SMW_ERR_PARAMS_SINGULAR; *sh_inv = f29_zero(); }
            else *sh_inv = f29_inv_canon(piv);
        }
        __syncthreads();
        F29 inv = *sh_inv;
        for (int i = k + 1 + threadIdx.x; i < n; i += blockDim.x) gs(A, (size_t)i * n + k, mulc(gl(A, (size_t)i * n + k), inv));
        __syncthreads();
        int w = n - k - 1;
        for (int e = threadIdx.x; e < w * w; e += blockDim.x) {
            int i = k + 1 + e / w, j = k + 1 + e % w;
            F29 prod = mulc(gl(A, (size_t)i * n + k), gl(A, (size_t)k * n + j));
            gs(A, (size_t)i * n + j, f29_sub_canon(gl(A, (size_t)i * n + j), prod));
        }
        __syncthreads();
    }
}
// C = A * B (n x n), canonical
__device__ void matmul(uint32_t* C, const uint32_t* A, const uint32_t* B, int n) {
    for (int e = threadIdx.x; e < n * n; e += blockDim.x) {
        int i = e / n, j = e % n; Acc acc; acc_zero(acc); int c = 0;
        for (int k = 0; k < n; k++) { acc_mac(acc, gl(A, (size_t)i * n + k), gl(B, (size_t)k * n + j)); if (++c == 6) { acc_normalize(acc); c = 0; } }
        gs(C, e, f29_canonicalize(acc_reduce(acc)));   // <= 129 canonical terms: sum/2^261 + p < 2p
    }
    __syncthreads();
}
__device__ void matcopy(uint32_t* D, const uint32_t* S, int n) { for (int e = threadIdx.x; e < n * n * EL; e += blockDim.x) D[e] = S[e]; __syncthreads(); }

__global__ void __launch_bounds__(256) k_derive_factor(DeriveArgs a) {
    __shared__ F29 sh_inv; __shared__ int sh_piv;
    const int t = a.t, n = t - 1, rp = a.rp;
    // 1. LU of the full matrix
    for (int e = threadIdx.x; e < t * t * EL; e += blockDim.x) a.lu_full[e] = a.mds[e];
    __syncthreads();
    lu_inplace(a.lu_full, t, a.status, &sh_inv);
    // 2. Mh = M[1:,1:], Mh^-1 by Gauss-Jordan on [Mh | I] (row pivoting allowed: it does not change the inverse)
    for (int e = threadIdx.x; e < n * n; e += blockDim.x) {
        int i = e / n, j = e % n; F29 x = gl(a.mds, (size_t)(i + 1) * t + (j + 1));
        gs(a.mh, e, x); gs(a.aug, (size_t)i * 2 * n + j, x); gs(a.aug, (size_t)i * 2 * n + n + j, i == j ? f29_const(C_ONE29) : f29_zero());
    }
    __syncthreads();
    for (int c = 0; c < n; c++) {
        if (threadIdx.x == 0) {
            int r = c; while (r < n && f29_is_zero_canon(gl(a.aug, (size_t)r * 2 * n + c))) r++;
            sh_piv = r;
            if (r == n) { *a.status = SMW_ERR_PARAMS_SINGULAR; sh_inv = f29_zero(); }
            else sh_inv = f29_inv_canon(gl(a.aug, (size_t)r * 2 * n + c));
        }
        __syncthreads();
        int r = sh_piv; if (r == n) r = c;
        if (r != c) { for (int j = threadIdx.x; j < 2 * n; j += blockDim.x) { F29 x = gl(a.aug, (size_t)r * 2 * n + j), y = gl(a.aug, (size_t)c * 2 * n + j); gs(a.aug, (size_t)r * 2 * n + j, y); gs(a.aug, (size_t)c * 2 * n + j, x); } }
        __syncthreads();
        F29 inv = sh_inv;
        for (int j = threadIdx.x; j < 2 * n; j += blockDim.x) gs(a.aug, (size_t)c * 2 * n + j, mulc(gl(a.aug, (size_t)c * 2 * n + j), inv));
        for (int i = threadIdx.x; i < n; i += blockDim.x) gs(a.tmp, i, gl(a.aug, (size_t)i * 2 * n + c));   // column factors before they change
        __syncthreads();
        for (int e = threadIdx.x; e < n * 2 * n; e += blockDim.x) {
            int i = e / (2 * n), j = e % (2 * n);
            if (i == c) continue;
            F29 prod = mulc(gl(a.tmp, i), gl(a.aug, (size_t)c * 2 * n + j));
            gs(a.aug, (size_t)i * 2 * n + j, f29_sub_canon(gl(a.aug, (size_t)i * 2 * n + j), prod));
        }
        __syncthreads();
    }
    for (int e = threadIdx.x; e < n * n; e += blockDim.x) gs(a.mhi, e, gl(a.aug, (size_t)(e / n) * 2 * n + n + e % n));
    __syncthreads();
    // 3. sparse rounds: row_k = (M00, v^T Mh^k), col_k = Mh^-(k+1) w
    for (int j = threadIdx.x; j < n; j += blockDim.x) { gs(a.vk, j, gl(a.mds, (size_t)(j + 1))); gs(a.vt, j, gl(a.mds, (size_t)(j + 1) * t)); }
    __syncthreads();
    // wk = Mh^-1 * w
    for (int i = threadIdx.x; i < n; i += blockDim.x) {
        Acc acc; acc_zero(acc); int c = 0;
        for (int j = 0; j < n; j++) { acc_mac(acc, gl(a.mhi, (size_t)i * n + j), gl(a.vt, j)); if (++c == 6) { acc_normalize(acc); c = 0; } }
        gs(a.wk, i, f29_canonicalize(acc_reduce(acc)));
    }
    __syncthreads();
    for (int k = 0; k < rp; k++) {
        // layout (TableLayout, poseidon_prog.cu): sp_row has t + 1 slots per round — even k: [row_k[1..], M00], odd k:
        // [row_k[1..], G, M00] with G = sum_i row_k[i] * col_{k-1}[i-1] (two partial rounds run as one block over the
        // lanes u_1 .., x, y); sp_col interleaves the column vectors of rounds 2b and 2b + 1
        const int tp = t + 1;
        if (threadIdx.x == 0) gs(a.sp_row, (size_t)k * tp + n + (k & 1), gl(a.mds, 0));
        for (int j = threadIdx.x; j < n; j += blockDim.x) {
            gs(a.sp_row, (size_t)k * tp + j, gl(a.vk, j));
            gs(a.sp_col, (size_t)(k >> 1) * 2 * n + 2 * j + (k & 1), gl(a.wk, j));
        }
        __syncthreads();
        if ((k & 1) && threadIdx.x == 0) {
            Acc acc; acc_zero(acc); int c = 0;
            for (int j = 0; j < n; j++) { acc_mac(acc, gl(a.vk, j), gl(a.sp_col, (size_t)(k >> 1) * 2 * n + 2 * j)); if (++c == 6) { acc_normalize(acc); c = 0; } }
            gs(a.sp_row, (size_t)k * tp + n, f29_canonicalize(acc_reduce(acc)));
        }
        __syncthreads();
        // vk <- vk^T * Mh ; wk <- Mh^-1 * wk   (through tmp / vt)
        for (int j = threadIdx.x; j < n; j += blockDim.x) {
            Acc acc; acc_zero(acc); int c = 0;
            for (int i = 0; i < n; i++) { acc_mac(acc, gl(a.vk, i), gl(a.mh, (size_t)i * n + j)); if (++c == 6) { acc_normalize(acc); c = 0; } }
            gs(a.tmp, j, f29_canonicalize(acc_reduce(acc)));
            Acc ac2; acc_zero(ac2); c = 0;
            for (int i = 0; i < n; i++) { acc_mac(ac2, gl(a.mhi, (size_t)j * n + i), gl(a.wk, i)); if (++c == 6) { acc_normalize(ac2); c = 0; } }
            gs(a.vt, j, f29_canonicalize(acc_reduce(ac2)));
        }
        __syncthreads();
        for (int j = threadIdx.x; j < n; j += blockDim.x) { gs(a.vk, j, gl(a.tmp, j)); gs(a.wk, j, gl(a.vt, j)); }
        __syncthreads();
    }
    // 4. post = Mh^rp by square-and-multiply, then its LU
    for (int e = threadIdx.x; e < n * n; e += blockDim.x) gs(a.lu_post, e, (e / n == e % n) ? f29_const(C_ONE29) : f29_zero());
    matcopy(a.pw, a.mh, n);
    for (int bit = 0; (rp >> bit) != 0; bit++) {
        if ((rp >> bit) & 1) { matmul(a.aug, a.lu_post, a.pw, n); matcopy(a.lu_post, a.aug, n); }
        if ((rp >> (bit + 1)) != 0) { matmul(a.aug, a.pw, a.pw, n); matcopy(a.pw, a.aug, n); }
    }
    matcopy(a.post, a.lu_post, n);             // the dense Mh^rp for the cooperative kernels, before the in-place LU
    lu_inplace(a.lu_post, n, a.status, &sh_inv);
}

// internal 29-bit tables (x * 2^261, canonical) -> the FP64-pipe form (5 doubles of x * 2^260 mod p) in the same element slots
__device__ __constant__ const uint32_t C_2P260_29[9] = {0x1fffffc1u, 0x14611707u, 0x1a9c5e9eu, 0x00b3f348u, 0x1fff790au, 0x1fffffffu, 0x1fffffffu, 0x1fffffffu, 0x003fffffu};
SMW_DEV void gs52(uint32_t* base, size_t idx, const F52& x) { double* q = reinterpret_cast<double*>(base + idx * EL);
#pragma unroll
    for (int i = 0; i < 5; i++) q[i] = x.v[i];
    q[5] = 0.0; }
__global__ void k_tables_to_52(const uint32_t* tbl, size_t n, uint32_t* tbl52, TableLayout L) {
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    uint32_t w[8]; f29_repack_to32(f29_canonicalize(f29_mul(gl(tbl, i), f29_const(C_2P260_29))), w);     // x 2^261 * 2^260 / 2^261
    F52 x = f52_repack_from32(w);
    if (i == L.k_in) x = f52_const(C_KIN52); else if (i == L.k_out) x = f52_const(C_KOUT52);
    else if (i == L.k_rr) x = f52_const(C_KRR52); else if (i == L.k_one) x = f52_const(C_KONE52);
    gs52(tbl52, i, x);
}
__global__ void k_store_consts(uint32_t* tbl, TableLayout L) {
    if (threadIdx.x != 0 || blockIdx.x != 0) return;
    gs(tbl, L.k_in, f29_const(C_KIN29)); gs(tbl, L.k_out, f29_const(C_KOUT29)); gs(tbl, L.k_rr, f29_const(C_RR29)); gs(tbl, L.k_one, f29_const(C_ONE29));
}
static const uint4 PAD_OP = pad_op();

int upload_programs(smw_ctx* c, smw_params* p) {
    size_t n = p->h_prog.size();
    mark_deps(p->h_prog, 0);        // idempotent: an operation's flag depends only on itself and its array predecessor
    if (n > p->d_prog_cap) {
        SMW_CUDA(c, cudaStreamSynchronize(c->stream));          // kernels in flight still read the old copy
        if (p->d_prog) cudaFree(p->d_prog);
        p->d_prog_cap = n + 4096;
        SMW_CUDA(c, cudaMalloc(&p->d_prog, p->d_prog_cap * sizeof(uint4)));
    }
    // the host vector only ever grows at the end, so in-flight kernels read unchanged ops
    SMW_CUDA(c, cudaMemcpyAsync(p->d_prog, p->h_prog.data(), n * sizeof(uint4), cudaMemcpyHostToDevice, c->stream));
    SMW_CUDA(c, cudaStreamSynchronize(c->stream));               // h_prog may be re-allocated by the next append
    return SMW_OK;
}
// Cached segment lists.  n_ds >= 0: DS sponge with n_ds integer DS fields and up to max_inputs elements.
// n_ds == -1: first[0] = "load lane i from external element i" ops, first[1] = "lane i -> external scale" ops (t each).
// n_ds == -2: FRI hashed-leaf pre-image (max_inputs = lane_f * 256 + lane_s): [17 lane loads][permutation].
int get_sponge_segs(smw_ctx* c, smw_params* p, int n_ds, int max_inputs, Segs* out) {
    auto key = std::make_pair(n_ds, max_inputs);
    auto it = p->sponge_cache.find(key);
    if (it != p->sponge_cache.end()) { *out = it->second; return SMW_OK; }
    const int t = p->t; const TableLayout& L = p->layout;
    Segs sg; sg.n = 0;
    if (n_ds >= 0) {
        if (n_ds > t - 1 || (n_ds + max_inputs + 1 + (t - 2)) / (t - 1) > 4) return set_err(c, SMW_ERR_INVALID_ARG, "sponge program too long");
        sg = build_sponge_program(p->h_prog, t, p->rp, n_ds, max_inputs, p->perm_first, p->perm_count);
    } else if (n_ds == -1) {
        sg.n = 2;
        sg.first[0] = (int)p->h_prog.size(); sg.count[0] = t;
        for (int i = 0; i < t; i++) p->h_prog.push_back(make_op(L.k_in, 0, 1, i, OPF_AEXT, 0, (uint32_t)i));
        p->h_prog.push_back(PAD_OP);
        sg.first[1] = (int)p->h_prog.size(); sg.count[1] = t;
        for (int i = 0; i < t; i++) p->h_prog.push_back(make_op(L.k_out, i, 1, i, 0));
        p->h_prog.push_back(PAD_OP);
    } else {
        const int lane_f = max_inputs / 256, lane_s = max_inputs % 256;
        sg.n = 2;
        sg.first[0] = (int)p->h_prog.size(); sg.count[0] = t;
        for (int j = 0; j < t; j++) {
            if (j == lane_f) p->h_prog.push_back(make_op(L.k_in, 0, 1, j, OPF_AEXT, 0, 0));
            else if (j == lane_s) p->h_prog.push_back(make_op(L.k_in, 0, 1, j, OPF_AEXT, 0, 1));
            else p->h_prog.push_back(make_op(L.k_rr, 0, 1, j, OPF_AEXT, 0, (uint32_t)(2 + j)));
        }
        p->h_prog.push_back(PAD_OP);
        sg.first[1] = p->perm_first; sg.count[1] = p->perm_count;
    }
    int rc = upload_programs(c, p); if (rc) return rc;
    p->sponge_cache[key] = sg;
    *out = sg;
    return SMW_OK;
}

static size_t tables_words(int t, int rp) { return (size_t)EL * table_layout(t, rp).total; }

// allocate tables; returns pointer to the raw (un-factored) matrix slot
static int alloc_tables(smw_ctx* c, smw_params* p, uint32_t** mds_internal) {
    int t = p->t, rp = p->rp;
    p->layout = table_layout(t, rp);
    const TableLayout& L = p->layout;
    SMW_CUDA(c, cudaMalloc(&p->d_tables, tables_words(t, rp) * 4));
    SMW_CUDA(c, cudaMemsetAsync(p->d_tables, 0, tables_words(t, rp) * 4, c->stream));
    uint32_t* q = p->d_tables;
    p->lu_full = q + (size_t)EL * L.lu_full; p->rc_full = q + (size_t)EL * L.rc_full; p->rc_part = q + (size_t)EL * L.rc_part;
    p->sp_row = q + (size_t)EL * L.sp_row; p->sp_col = q + (size_t)EL * L.sp_col; p->lu_post = q + (size_t)EL * L.lu_post;
    *mds_internal = q + (size_t)EL * L.mds;
    k_store_consts<<<1, 32, 0, c->stream>>>(p->d_tables, L); c->launches++;
    SMW_CUDA(c, cudaMalloc(&p->d_ext, ((size_t)t * t + 8 * t + rp) * 32));
    // the permutation programs (indices only; the tables they point into are filled below)
    p->h_prog.clear();
    p->perm_first = 0; build_perm_program(p->h_prog, t, rp, true); p->perm_count = (int)p->h_prog.size(); p->h_prog.push_back(PAD_OP);
    p->dense_first = (int)p->h_prog.size(); build_perm_program(p->h_prog, t, rp, false); p->dense_count = (int)p->h_prog.size() - p->dense_first; p->h_prog.push_back(PAD_OP);
    return upload_programs(c, p);
}
static int finish_params(smw_ctx* c, smw_params* p, uint32_t* mds_internal) {
    int t = p->t, rp = p->rp, n = t - 1;
    // external-form copies for export
    size_t n_ext = (size_t)t * t;
    k_internal_to_ext<<<(unsigned)((n_ext + 127) / 128), 128, 0, c->stream>>>(mds_internal, n_ext, p->d_ext);
    k_internal_to_ext<<<(unsigned)((8 * t + 127) / 128), 128, 0, c->stream>>>(p->rc_full, 8 * t, p->d_ext + n_ext * 8);
    k_internal_to_ext<<<(unsigned)((rp + 127) / 128), 128, 0, c->stream>>>(p->rc_part, rp, p->d_ext + (n_ext + 8 * t) * 8);
    c->launches += 3;
    // scratch
    size_t nn = (size_t)n * n, words = EL * (nn + nn + 2 * nn + nn + 4 * (size_t)t);
    uint32_t* scratch; int* d_status;
    SMW_CUDA(c, cudaMalloc(&scratch, words * 4)); SMW_CUDA(c, cudaMalloc(&d_status, 4));
    SMW_CUDA(c, cudaMemsetAsync(d_status, 0, 4, c->stream));
    DeriveArgs a; a.t = t; a.rp = rp; a.mds = mds_internal;
    a.lu_full = (uint32_t*)p->lu_full; a.sp_row = (uint32_t*)p->sp_row; a.sp_col = (uint32_t*)p->sp_col; a.lu_post = (uint32_t*)p->lu_post; a.post = p->d_tables + (size_t)EL * p->layout.post;
    uint32_t* q = scratch; a.mh = q; q += EL * nn; a.mhi = q; q += EL * nn; a.aug = q; q += EL * 2 * nn; a.pw = q; q += EL * nn;
    a.tmp = q; q += EL * t; a.vk = q; q += EL * t; a.wk = q; q += EL * t; a.vt = q; q += EL * t; a.status = d_status;
    k_derive_factor<<<1, 256, 0, c->stream>>>(a); c->launches++;
    int st = 0;
    SMW_CUDA(c, cudaMemcpyAsync(&st, d_status, 4, cudaMemcpyDeviceToHost, c->stream));
    SMW_CUDA(c, cudaStreamSynchronize(c->stream));
    cudaFree(scratch); cudaFree(d_status);
    if (st != 0) return set_err(c, st, "Poseidon matrix has a zero pivot (no LU / sparse factorisation)");
    if (t == 9 || t == 17) {       // the throughput-bound widths also get the FP64-pipe tables (poseidon52.cuh)
        const size_t ne = p->layout.total;
        SMW_CUDA(c, cudaMalloc(&p->d_tables52, tables_words(t, rp) * 4));
        k_tables_to_52<<<(unsigned)((ne + 127) / 128), 128, 0, c->stream>>>(p->d_tables, ne, p->d_tables52, p->layout); c->launches++;
        SMW_CUDA(c, cudaStreamSynchronize(c->stream));
    }
    return SMW_OK;
}
int launch_derive_params(smw_ctx* c, smw_params* p, const unsigned char* seed, int seed_len) {
    if (seed_len > 96) return set_err(c, SMW_ERR_INVALID_ARG, "seed longer than 96 bytes");
    uint32_t* mds_i; int rc = alloc_tables(c, p, &mds_i); if (rc) return rc;
    SeedArg s; memset(&s, 0, sizeof s); memcpy(s.b, seed, seed_len); s.len = seed_len;
    int total = p->t * p->t + 8 * p->t + p->rp;
    k_derive_consts<<<(total + 63) / 64, 64, 0, c->stream>>>(p->t, p->rp, s, mds_i, (uint32_t*)p->rc_full, (uint32_t*)p->rc_part); c->launches++;
    return finish_params(c, p, mds_i);
}
int launch_upload_params(smw_ctx* c, smw_params* p, const uint32_t* d_mds, const uint32_t* d_rcf, const uint32_t* d_rcp) {
    uint32_t* mds_i; int rc = alloc_tables(c, p, &mds_i); if (rc) return rc;
    size_t n1 = (size_t)p->t * p->t, n2 = 8 * p->t, n3 = p->rp;
    k_ext_to_internal<<<(unsigned)((n1 + 127) / 128), 128, 0, c->stream>>>(d_mds, n1, mds_i);
    k_ext_to_internal<<<(unsigned)((n2 + 127) / 128), 128, 0, c->stream>>>(d_rcf, n2, (uint32_t*)p->rc_full);
    k_ext_to_internal<<<(unsigned)((n3 + 127) / 128), 128, 0, c->stream>>>(d_rcp, n3, (uint32_t*)p->rc_part);
    c->launches += 3;
    return finish_params(c, p, mds_i);
}


#define DISPATCH_T(t, ...)                                                                \
    switch (t) {                                                                           \
        case 9: { constexpr int T = 9; __VA_ARGS__; } break;                               \
        case 17: { constexpr int T = 17; __VA_ARGS__; } break;                             \
        case 33: { constexpr int T = 33; __VA_ARGS__; } break;                             \
        case 65: { constexpr int T = 65; __VA_ARGS__; } break;                             \
        case 129: { constexpr int T = 129; __VA_ARGS__; } break;                           \
        default: return set_err(c, SMW_ERR_UNSUPPORTED_WIDTH, "unsupported Poseidon width"); \
    }

int launch_permute_batch(smw_ctx* c, const smw_params* p, uint32_t* d_states, size_t batch, bool sparse) {
    if (batch == 0) return SMW_OK;
    DISPATCH_T(p->t, return permute_batch_T<T>(c, p, d_states, batch, sparse));
    return SMW_OK;
}
int launch_merkle_level(smw_ctx* c, const smw_params* p, const uint32_t* d_children, size_t n_children, uint32_t* d_parents, size_t n_parents,
                        int arity, uint32_t level, uint64_t label0, uint64_t label_stride, uint64_t pos_offset,
                        int n_cols, size_t child_col_stride, size_t parent_col_stride) {
    if (n_parents == 0 || n_cols == 0) return SMW_OK;
    DISPATCH_T(p->t, return merkle_level_T<T>(c, p, d_children, n_children, d_parents, n_parents, arity, level, label0, label_stride, pos_offset, n_cols, child_col_stride, parent_col_stride));
    return SMW_OK;
}
int launch_pair_leaves(smw_ctx* c, const smw_params* p, const uint32_t* d_f, const uint32_t* d_cp, size_t n, int arity, uint64_t label, uint64_t pos_offset,
                       uint32_t* d_out, size_t cp_div) {
    if (n == 0) return SMW_OK;
    DISPATCH_T(p->t, return pair_leaves_T<T>(c, p, d_f, d_cp, n, arity, label, pos_offset, d_out, cp_div));
    return SMW_OK;
}
int launch_merkle_level_multi(smw_ctx* c, const smw_params* p, const NodeSeg* d_segs, int n_segs, uint64_t total_nodes, int arity, bool all_full) {
    DISPATCH_T(p->t, return merkle_level_multi_T<T>(c, p, d_segs, n_segs, total_nodes, arity, all_full));
    return SMW_OK;
}
int launch_pair_leaves_multi(smw_ctx* c, const smw_params* p, const LeafSeg* d_segs, int n_segs, uint64_t total) {
    DISPATCH_T(p->t, return pair_leaves_multi_T<T>(c, p, d_segs, n_segs, total));
    return SMW_OK;
}
int launch_hash_groups(smw_ctx* c, const smw_params* p, const uint32_t* d_pool, const uint64_t* d_child_idx, const uint8_t* d_cnt, const uint64_t* d_pos,
                       size_t n_groups, int arity, int row, uint32_t level, uint64_t label, uint32_t* d_out) {
    if (n_groups == 0) return SMW_OK;
    DISPATCH_T(p->t, return hash_groups_T<T>(c, p, d_pool, d_child_idx, d_cnt, d_pos, n_groups, arity, row, level, label, d_out));
    return SMW_OK;
}
int launch_hash_ds_single(smw_ctx* c, const smw_params* p, const uint32_t* d_ds, int nds, const uint32_t* d_in, size_t n, uint32_t* d_out) {
    DISPATCH_T(p->t, return hash_ds_single_T<T>(c, p, d_ds, nds, d_in, n, d_out));
    return SMW_OK;
}

} // namespace smw
