// capi.cu — the C ABI (include/smw.h) over the CUDA kernels: handle management, host<->device staging,
// and the orchestration of level-by-level tree builds, the FRI-like commit phase and the
// Merkle-folded sum-check commit loop.  No field arithmetic happens on the host: byte strings are only
// packed into 256-bit little-endian words (transcript/src/lib.rs:13-41) and reduced on the device.
#include "smw_internal.h"
#include <cstring>
#include <cstdlib>
#include <cstdio>
#include <algorithm>
#include <memory>
#include <array>
#include <map>

using namespace smw;

namespace smw {
int set_err(smw_ctx* c, int code, const char* what) { if (c) c->last_error = what ? what : ""; return code; }
int cuda_err(smw_ctx* c, cudaError_t e, const char* what) {
    if (c) { c->last_error = std::string(what ? what : "") + ": " + cudaGetErrorString(e); }
    return SMW_ERR_CUDA;
}
int ctx_scratch(smw_ctx* c, int slot, size_t bytes, void** out) {
    if (bytes > c->slot_bytes[slot]) {
        SMW_CUDA(c, cudaStreamSynchronize(c->stream));
        if (c->slot_ptr[slot]) cudaFree(c->slot_ptr[slot]);
        c->slot_ptr[slot] = nullptr; c->slot_bytes[slot] = 0;
        SMW_CUDA(c, cudaMalloc(&c->slot_ptr[slot], bytes));
        c->slot_bytes[slot] = bytes;
    }
    *out = c->slot_ptr[slot];
    return SMW_OK;
}
// ---- arena pool (smw_internal.h) ----
static constexpr size_t POOL_MIN = (size_t)1 << 20, POOL_MAX_FREE_BYTES = (size_t)24 << 30, POOL_MAX_FREE_BLOCKS = 16;
void pool_release(smw_ctx* c) {
    for (auto& b : c->pool_freelist) { c->pool_owned.erase(b.first); cudaFree(b.first); }
    c->pool_freelist.clear(); c->pool_free_bytes = 0;
}
cudaError_t pool_malloc(smw_ctx* c, void** out, size_t bytes) {
    *out = nullptr;
    if (bytes < POOL_MIN) return cudaMalloc(out, bytes ? bytes : 1);
    int best = -1;                                             // the smallest returned block that fits without wasting more than half of itself
    for (int i = 0; i < (int)c->pool_freelist.size(); i++) {
        const size_t cap = c->pool_freelist[i].second;
        if (cap >= bytes && cap <= 2 * bytes && (best < 0 || cap < c->pool_freelist[best].second)) best = i;
    }
    if (best >= 0) {
        *out = c->pool_freelist[best].first; c->pool_free_bytes -= c->pool_freelist[best].second;
        c->pool_freelist.erase(c->pool_freelist.begin() + best);
        return cudaSuccess;
    }
    cudaError_t e = cudaMalloc(out, bytes);
    if (e != cudaSuccess) { cudaGetLastError(); pool_release(c); e = cudaMalloc(out, bytes); }      // give the pooled blocks back and try once more
    if (e == cudaSuccess) c->pool_owned[*out] = bytes;
    return e;
}
void pool_free(smw_ctx* c, void* p) {
    if (!p) return;
    auto it = c->pool_owned.find(p);
    if (it == c->pool_owned.end()) { cudaFree(p); return; }
    c->pool_freelist.push_back({p, it->second}); c->pool_free_bytes += it->second;
    while (c->pool_freelist.size() > POOL_MAX_FREE_BLOCKS || c->pool_free_bytes > POOL_MAX_FREE_BYTES) {      // oldest first
        auto b = c->pool_freelist.front(); c->pool_freelist.erase(c->pool_freelist.begin());
        c->pool_free_bytes -= b.second; c->pool_owned.erase(b.first); cudaFree(b.first);
    }
}
} // namespace smw

namespace {

constexpr size_t FR = 32;

// caller buffer -> device pointer (copy in when it lives on the host)
struct In {
    smw_ctx* c; const uint32_t* d = nullptr; uint32_t* owned = nullptr; int rc = SMW_OK;
    In(smw_ctx* ctx, const void* p, size_t n_fr, int mem) : c(ctx) {
        if (n_fr == 0 || p == nullptr) return;
        if (mem == SMW_DEVICE) { d = (const uint32_t*)p; return; }
        cudaError_t e = cudaMalloc(&owned, n_fr * FR);
        if (e != cudaSuccess) { rc = cuda_err(c, e, "cudaMalloc(input staging)"); return; }
        e = cudaMemcpyAsync(owned, p, n_fr * FR, cudaMemcpyHostToDevice, c->stream);
        if (e != cudaSuccess) { rc = cuda_err(c, e, "cudaMemcpyAsync(H2D)"); return; }
        d = owned;
    }
    uint32_t* release() { uint32_t* r = owned; owned = nullptr; return r; }
    ~In() { if (owned) { cudaStreamSynchronize(c->stream); cudaFree(owned); } }
};
// caller output buffer: device scratch when it lives on the host, copied back by finish()
struct Out {
    smw_ctx* c; uint32_t* d = nullptr; void* host = nullptr; size_t n; bool owned = false; int rc = SMW_OK;
    Out(smw_ctx* ctx, void* p, size_t n_fr, int mem) : c(ctx), n(n_fr) {
        if (n_fr == 0) return;
        if (mem == SMW_DEVICE) { d = (uint32_t*)p; return; }
        host = p; owned = true;
        cudaError_t e = cudaMalloc(&d, n_fr * FR);
        if (e != cudaSuccess) rc = cuda_err(c, e, "cudaMalloc(output staging)");
    }
    int finish() {
        if (!owned || n == 0) return SMW_OK;
        SMW_CUDA(c, cudaMemcpyAsync(host, d, n * FR, cudaMemcpyDeviceToHost, c->stream));
        SMW_CUDA(c, cudaStreamSynchronize(c->stream));
        return SMW_OK;
    }
    ~Out() { if (owned && d) { cudaStreamSynchronize(c->stream); cudaFree(d); } }
};
struct DevFree { std::vector<void*> v; cudaStream_t st; smw_ctx* pool = nullptr;      // pool != nullptr: blocks go back to the context's arena pool
                 ~DevFree() { cudaStreamSynchronize(st); for (void* p : v) if (p) { if (pool) smw::pool_free(pool, p); else cudaFree(p); } } };
int d2h(smw_ctx* c, void* host, const void* dev, size_t bytes) {
    SMW_CUDA(c, cudaMemcpyAsync(host, dev, bytes, cudaMemcpyDeviceToHost, c->stream));
    SMW_CUDA(c, cudaStreamSynchronize(c->stream));
    return SMW_OK;
}

int width_for_arity(int arity) {   // poseidon/src/lib.rs:155-166
    if (arity < 0) return -1;
    if (arity <= 8) return 9; if (arity <= 16) return 17; if (arity <= 32) return 33; if (arity <= 64) return 65; if (arity <= 128) return 129;
    return -1;
}
int rp_for_width(int t) { switch (t) { case 9: return 60; case 17: return 64; case 33: return 68; case 65: return 76; case 129: return 84; default: return -1; } }

// ---- transcript byte packing (transcript/src/lib.rs:13-41): 256-bit LE words, reduced on the device
using Word = std::array<uint32_t, 8>;
Word word_from_bytes(const uint8_t* b, size_t n) {   // n <= 32
    uint8_t le[32]; memset(le, 0, 32); memcpy(le, b, n);
    Word w; for (int i = 0; i < 8; i++) w[i] = (uint32_t)le[4 * i] | ((uint32_t)le[4 * i + 1] << 8) | ((uint32_t)le[4 * i + 2] << 16) | ((uint32_t)le[4 * i + 3] << 24);
    return w;
}
Word word_from_u64(uint64_t v) { Word w{}; w[0] = (uint32_t)v; w[1] = (uint32_t)(v >> 32); return w; }
Word tag_word(const char* s) { return word_from_bytes((const uint8_t*)s, strlen(s)); }
const char* TAG_INIT = "FSv1-TRANSCRIPT-INIT"; const char* TAG_ABSORB = "FSv1-ABSORB-BYTES"; const char* TAG_CHAL = "FSv1-CHALLENGE";
void push_absorb_bytes(std::vector<Word>& v, const uint8_t* b, size_t n) {   // marker, then 31-byte words
    v.push_back(tag_word(TAG_ABSORB));
    for (size_t off = 0; off < n; off += 31) v.push_back(word_from_bytes(b + off, std::min<size_t>(31, n - off)));
}
void push_challenge(std::vector<Word>& v, const uint8_t* label, size_t n) { v.push_back(tag_word(TAG_CHAL)); push_absorb_bytes(v, label, n); }
int advance_pos(int pos, size_t count) { for (size_t i = 0; i < count; i++) { if (pos == 16) pos = 0; pos++; } return pos; }

int upload_words(smw_ctx* c, const std::vector<Word>& w, uint32_t** d) {
    *d = nullptr; if (w.empty()) return SMW_OK;
    SMW_CUDA(c, cudaMalloc(d, w.size() * FR));
    SMW_CUDA(c, cudaMemcpyAsync(*d, w.data(), w.size() * FR, cudaMemcpyHostToDevice, c->stream));
    return SMW_OK;
}

int get_transcript_params(smw_ctx* c, smw_params* given, smw_params** out) {
    if (given) { if (given->t != 17) return set_err(c, SMW_ERR_UNSUPPORTED_WIDTH, "transcript params must have t = 17"); *out = given; return SMW_OK; }
    return smw_params_transcript_default(c, out);
}

constexpr size_t COOP_ABSORB_MIN = 64;      // elements from which the warp-cooperative absorb (k_transcript_absorb_coop) pays
// one-shot transcript hash on the device: new(label0) ; absorb_bytes(tag) ; fields ; challenge(chal)
// `defer` != nullptr: nothing is synchronised or freed here; the caller syncs the stream(s) and frees the collected pointers
int oneshot_hash(smw_ctx* c, smw_params* trp, const char* label0, const uint8_t* tag, size_t tag_len, const std::vector<Word>& imm_fields,
                 const uint32_t* d_ext, size_t n_ext, const char* chal, uint32_t* d_out, std::vector<void*>* defer = nullptr) {
    std::vector<Word> A, C;
    A.push_back(tag_word(TAG_INIT));
    push_absorb_bytes(A, (const uint8_t*)label0, strlen(label0));
    push_absorb_bytes(A, tag, tag_len);
    for (auto& w : imm_fields) A.push_back(w);
    push_challenge(C, (const uint8_t*)chal, strlen(chal));
    uint32_t *dA, *dC; int rc = upload_words(c, A, &dA); if (rc) return rc; rc = upload_words(c, C, &dC); if (rc) { cudaFree(dA); return rc; }
    if (n_ext >= COOP_ABSORB_MIN) {          // long input: the serial sponge runs on one warp instead of one thread
        uint32_t* d_st = nullptr;
        if (cudaMalloc(&d_st, 17 * FR) != cudaSuccess) { cudaFree(dA); cudaFree(dC); return set_err(c, SMW_ERR_NOMEM, "cudaMalloc(sponge state)"); }
        rc = launch_transcript_run(c, trp, d_st, 1, 0, dA, (int)A.size(), nullptr, 0, nullptr, 0, 0, nullptr);
        int pos = advance_pos(0, A.size() - 1);
        if (!rc) rc = launch_transcript_absorb_coop(c, trp, d_st, pos, d_ext, n_ext);
        pos = advance_pos(pos, n_ext);
        if (!rc) rc = launch_transcript_run(c, trp, d_st, 0, pos, nullptr, 0, nullptr, 0, dC, (int)C.size(), 1, d_out);
        if (defer) { defer->push_back(d_st); defer->push_back(dA); defer->push_back(dC); return rc; }
        cudaStreamSynchronize(c->stream);
        cudaFree(d_st); cudaFree(dA); cudaFree(dC);
        return rc;
    }
    rc = launch_transcript_run(c, trp, nullptr, 1, 0, dA, (int)A.size(), d_ext, n_ext, dC, (int)C.size(), 1, d_out);
    if (defer) { defer->push_back(dA); defer->push_back(dC); return rc; }
    cudaStreamSynchronize(c->stream);
    cudaFree(dA); cudaFree(dC);
    return rc;
}

// Runs independent chains of launches on side streams of the context and joins them back into its main stream.  The layers of
// a FRI commit phase are independent trees whose upper levels (< 148 CTAs) are latency-bound single-wave launches (~0.3-1 ms
// each, ~16 per layer): on one stream cfg3 spent a third of its time there; on eight streams they overlap the next layers' wide levels.
struct StreamFan {
    smw_ctx* c; cudaStream_t main; int K; std::vector<char> used;
    StreamFan(smw_ctx* ctx, int k) : c(ctx), main(ctx->stream), K(k), used(k, 0) {
        while ((int)c->side.size() < K) {
            cudaStream_t s; cudaStreamCreateWithFlags(&s, cudaStreamNonBlocking); c->side.push_back(s);
            cudaEvent_t e; cudaEventCreateWithFlags(&e, cudaEventDisableTiming); c->side_ev.push_back(e);
        }
        if (!c->fan_ev) cudaEventCreateWithFlags(&c->fan_ev, cudaEventDisableTiming);
        cudaEventRecord(c->fan_ev, main);
    }
    void enter(int i) { int k = i % K; if (!used[k]) { cudaStreamWaitEvent(c->side[k], c->fan_ev, 0); used[k] = 1; } c->stream = c->side[k]; }
    void leave() { c->stream = main; }
    ~StreamFan() {
        c->stream = main;
        for (int k = 0; k < K; k++) if (used[k]) { cudaEventRecord(c->side_ev[k], c->side[k]); cudaStreamWaitEvent(main, c->side_ev[k], 0); }
    }
};
constexpr int FAN_STREAMS = 8;

// wpow[i] = w_G^i, i < G = 2^log_g: a vector of ones scaled by the powers of the order-G root (freed by `fr`)
int make_wpow(smw_ctx* c, int log_g, uint32_t** d_w_out, DevFree& fr) {
    const size_t G = (size_t)1 << log_g;
    uint32_t *d_w = nullptr, *d_root = nullptr, *d_one = nullptr;
    SMW_CUDA(c, cudaMalloc(&d_w, G * FR)); fr.v.push_back(d_w);
    SMW_CUDA(c, cudaMalloc(&d_root, FR)); fr.v.push_back(d_root);
    SMW_CUDA(c, cudaMalloc(&d_one, FR)); fr.v.push_back(d_one);
    int rc = launch_root_of_unity(c, log_g, d_root); if (rc) return rc;
    rc = launch_root_of_unity(c, 0, d_one); if (rc) return rc;
    for (size_t i = 0; i < G; i++) SMW_CUDA(c, cudaMemcpyAsync(d_w + i * 8, d_one, FR, cudaMemcpyDeviceToDevice, c->stream));
    rc = launch_scale_powers(c, d_w, G, d_root); if (rc) return rc;
    *d_w_out = d_w;
    return SMW_OK;
}

// ---- forests: the levels of many independent trees hashed round by round, one launch per round and (parameter set, arity)
// level sizes of a tree over n leaves (level 0 = leaves); whole_only: stop at the last level made of complete sibling groups
// (the "cap" of a slice of a range-sharded tree: the ragged top belongs to whoever gathers the caps)
std::vector<size_t> level_sizes(size_t n, int arity, bool whole_only) {
    std::vector<size_t> len{n};
    if (arity < 2) return len;
    while (len.back() > 1) {
        const size_t cn = len.back();
        if (whole_only && (cn < (size_t)arity || cn % (size_t)arity)) break;
        len.push_back((cn + arity - 1) / arity);
    }
    return len;
}
struct ForestTree {
    smw_params* p; int arity; uint64_t label; uint64_t first_leaf;     // first_leaf: global index of leaf 0 (DS positions of a slice, merkle/src/lib.rs:168-173)
    uint32_t first_ds_level; bool whole_only;
    uint32_t* leaves; size_t n;
    std::vector<size_t> len; std::vector<uint32_t*> lv;                // filled by build_forest: level 0 = leaves
};
bool has_multi(const smw_ctx* c, const smw_params* p) { return c->use_f52 && p->d_tables52 != nullptr; }
// Hashes levels 1.. of every tree (merkle/src/lib.rs:164-181) on the context's stream.  All levels >= 1 of all trees live in ONE
// allocation (*arena_out); the segment table goes to `defer` (freed by the caller after the stream has drained).
int build_forest(smw_ctx* c, std::vector<ForestTree>& trees, uint32_t** arena_out, std::vector<void*>& defer) {
    Range rg("build_forest");
    *arena_out = nullptr;
    size_t total = 0, max_levels = 0;
    for (auto& t : trees) {
        t.len = level_sizes(t.n, t.arity, t.whole_only);
        for (size_t j = 1; j < t.len.size(); j++) total += t.len[j];
        max_levels = std::max(max_levels, t.len.size());
    }
    uint32_t* arena = nullptr;
    if (total) SMW_CUDA(c, pool_malloc(c, (void**)&arena, total * FR));
    *arena_out = arena;
    { size_t off = 0; for (auto& t : trees) { t.lv.assign(1, t.leaves); for (size_t j = 1; j < t.len.size(); j++) { t.lv.push_back(arena + off * 8); off += t.len[j]; } } }
    struct Launch { smw_params* p; int arity; size_t seg0; int n; uint64_t total; };
    std::vector<NodeSeg> segs; std::vector<Launch> launches;
    for (size_t j = 0; j + 1 < max_levels; j++) {
        std::vector<std::pair<std::pair<smw_params*, int>, std::vector<size_t>>> groups;      // insertion-ordered (params, arity) -> trees
        for (size_t ti = 0; ti < trees.size(); ti++) {
            const ForestTree& t = trees[ti];
            if (j + 1 >= t.len.size()) continue;
            auto key = std::make_pair(t.p, t.arity);
            auto it = std::find_if(groups.begin(), groups.end(), [&](const auto& g) { return g.first == key; });
            if (it == groups.end()) { groups.push_back({key, {}}); it = groups.end() - 1; }
            it->second.push_back(ti);
        }
        for (auto& g : groups) {
            Launch L{g.first.first, g.first.second, segs.size(), 0, 0};
            for (size_t ti : g.second) {
                const ForestTree& t = trees[ti];
                uint64_t first = t.first_leaf; for (size_t q = 0; q <= j; q++) first /= (uint64_t)t.arity;      // first parent of level j + 1
                segs.push_back(NodeSeg{t.lv[j], t.lv[j + 1], (uint64_t)t.len[j], (uint64_t)t.len[j + 1], t.label, first, t.first_ds_level + (uint32_t)j, (uint32_t)t.arity, L.total});
                L.total += t.len[j + 1]; L.n++;
            }
            launches.push_back(L);
        }
    }
    if (segs.empty()) return SMW_OK;
    if (trees.size() == 1) {            // one tree: plain per-level launches, no segment table
        for (const NodeSeg& sg : segs) {
            int rc = launch_merkle_level(c, trees[0].p, sg.children, sg.n_children, sg.parents, sg.n_parents, trees[0].arity, sg.level, sg.label, 0, sg.pos_offset, 1, 0, 0);
            if (rc) return rc;
        }
        return SMW_OK;
    }
    NodeSeg* d_segs = nullptr;
    SMW_CUDA(c, cudaMalloc(&d_segs, segs.size() * sizeof(NodeSeg)));
    defer.push_back(d_segs);
    SMW_CUDA(c, cudaMemcpyAsync(d_segs, segs.data(), segs.size() * sizeof(NodeSeg), cudaMemcpyHostToDevice, c->stream));   // pageable source: staged before the call returns
    for (const Launch& L : launches) {
        int rc = SMW_OK;
        bool all_full = true;
        for (int i = 0; i < L.n; i++) { const NodeSeg& sg = segs[L.seg0 + i]; all_full = all_full && sg.n_children == sg.n_parents * (uint64_t)L.arity; }
        if (has_multi(c, L.p)) rc = launch_merkle_level_multi(c, L.p, d_segs + L.seg0, L.n, L.total, L.arity, all_full);
        else for (int i = 0; i < L.n && !rc; i++) {
            const NodeSeg& sg = segs[L.seg0 + i];
            rc = launch_merkle_level(c, L.p, sg.children, sg.n_children, sg.parents, sg.n_parents, L.arity, sg.level, sg.label, 0, sg.pos_offset, 1, 0, 0);
        }
        if (rc) return rc;
    }
    return SMW_OK;
}
smw_tree* tree_handle(smw_ctx* c, const ForestTree& t, bool own_leaves) {
    auto h = new smw_tree();
    h->ctx = c; h->params = t.p; h->arity = t.arity; h->tree_label = t.label;
    h->len = t.len; h->lv = t.lv; h->owned.assign(t.lv.size(), false); h->owned[0] = own_leaves;
    return h;
}
// level-by-level tree over a device level 0 (merkle/src/lib.rs:164-181); the levels above live in one allocation owned by the tree
int build_tree(smw_ctx* c, smw_params* p, int arity, uint64_t label, uint32_t* d_level0, bool own0, size_t n, smw_tree** out, uint32_t first_ds_level = 0) {
    std::vector<ForestTree> f(1);
    f[0] = ForestTree{p, arity, label, 0, first_ds_level, false, d_level0, n, {}, {}};
    uint32_t* arena = nullptr; std::vector<void*> defer;
    int rc = build_forest(c, f, &arena, defer);
    if (!defer.empty()) { cudaStreamSynchronize(c->stream); for (void* q : defer) cudaFree(q); }
    if (rc) { cudaStreamSynchronize(c->stream); pool_free(c, arena); if (own0) cudaFree(d_level0); return rc; }
    smw_tree* t = tree_handle(c, f[0], own0); t->arena = arena;
    *out = t;
    return SMW_OK;
}

} // namespace

extern "C" {

// ---------------------------------------------------------------------------------------- context
int32_t smw_ctx_create_on_stream(int32_t device, void* cuda_stream, smw_ctx** out) {
    if (!out) return SMW_ERR_INVALID_ARG;
    *out = nullptr;
    int ndev = 0;
    cudaError_t e = cudaGetDeviceCount(&ndev);
    if (e != cudaSuccess || ndev == 0) return SMW_ERR_CUDA;      // no CPU fallback
    if (device < 0 || device >= ndev) return SMW_ERR_INVALID_ARG;
    if (cudaSetDevice(device) != cudaSuccess) return SMW_ERR_CUDA;
    smw_ctx* c = new smw_ctx();
    c->device = device;
    { const char* e = getenv("SMW_TC_MIN"); if (e) c->tc_min_nodes = atoll(e) > 0 ? (uint64_t)atoll(e) : ~0ull; }     // SMW_TC_MIN=0: the tensor-core node hash is never taken
    { const char* rep = getenv("SMW_REP"); c->use_f52 = !(rep && strcmp(rep, "29") == 0); }     // SMW_REP=29: integer-pipe multiplier everywhere
    if (cuda_stream) { c->stream = (cudaStream_t)cuda_stream; c->own_stream = false; }
    else { if (cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking) != cudaSuccess) { delete c; return SMW_ERR_CUDA; } c->own_stream = true; }
    *out = c;
    return SMW_OK;
}
int32_t smw_ctx_create(int32_t device, smw_ctx** out) { return smw_ctx_create_on_stream(device, nullptr, out); }
static void free_params_internal(smw_params* p) { if (!p) return; if (p->tc && p->tc_free) p->tc_free(p->tc); cudaFree(p->d_tables); cudaFree(p->d_tables52); cudaFree(p->d_ext); cudaFree(p->d_prog); delete p; }
void smw_ctx_destroy(smw_ctx* c) {
    if (!c) return;
    cudaSetDevice(c->device);
    cudaStreamSynchronize(c->stream);
    free_params_internal(c->commitment_default); free_params_internal(c->transcript_default);
    for (auto& kv : c->by_width) free_params_internal(kv.second);
    for (auto& r : c->prof) { cudaEventDestroy(r.a); cudaEventDestroy(r.b); }
    for (auto e : c->prof_pool) cudaEventDestroy(e);
    smw::pool_release(c); for (auto& kv : c->pool_owned) cudaFree(kv.first); c->pool_owned.clear();
    cudaFree(c->d_twiddles); for (int i = 0; i < smw_ctx::N_SLOTS; i++) cudaFree(c->slot_ptr[i]);
    if (c->h_pinned) cudaFreeHost(c->h_pinned);
    for (auto st : c->side) cudaStreamDestroy(st);
    for (auto e : c->side_ev) cudaEventDestroy(e);
    if (c->fan_ev) cudaEventDestroy(c->fan_ev);
    if (c->own_stream) cudaStreamDestroy(c->stream);
    delete c;
}
int32_t smw_ctx_sync(smw_ctx* c) { if (!c) return SMW_ERR_INVALID_ARG; SMW_CUDA(c, cudaStreamSynchronize(c->stream)); return SMW_OK; }
int32_t smw_ctx_release_cached(smw_ctx* c) { if (!c) return SMW_ERR_INVALID_ARG; cudaSetDevice(c->device); SMW_CUDA(c, cudaStreamSynchronize(c->stream)); smw::pool_release(c); return SMW_OK; }
const char* smw_last_error(const smw_ctx* c) { return c ? c->last_error.c_str() : "null context"; }
uint64_t smw_ctx_launch_count(const smw_ctx* c) { return c ? c->launches : 0; }
void* smw_ctx_stream(const smw_ctx* c) { return c ? (void*)c->stream : nullptr; }
int32_t smw_ctx_profile_enable(smw_ctx* c, int32_t on) {
    if (!c) return SMW_ERR_INVALID_ARG;
    cudaSetDevice(c->device);
    SMW_CUDA(c, cudaStreamSynchronize(c->stream));
    for (auto& r : c->prof) { c->prof_pool.push_back(r.a); c->prof_pool.push_back(r.b); }
    c->prof.clear();
    c->prof_on = on != 0;
    return SMW_OK;
}
int32_t smw_ctx_profile_read(smw_ctx* c, double* ms_out, uint64_t* launches_out, int32_t n_classes) {
    if (!c || n_classes < 0) return SMW_ERR_INVALID_ARG;
    cudaSetDevice(c->device);
    SMW_CUDA(c, cudaStreamSynchronize(c->stream));
    for (int i = 0; i < n_classes; i++) { if (ms_out) ms_out[i] = 0; if (launches_out) launches_out[i] = 0; }
    for (auto& r : c->prof) {
        float ms = 0; cudaEventElapsedTime(&ms, r.a, r.b);
        if (r.cls < n_classes) { if (ms_out) ms_out[r.cls] += ms; if (launches_out) launches_out[r.cls] += r.n; }
        c->prof_pool.push_back(r.a); c->prof_pool.push_back(r.b);
    }
    c->prof.clear();
    return SMW_OK;
}
int32_t smw_diag_mac_rate(smw_ctx* c, double* mac_per_s, double* warp_instr_per_clk_per_sm) {
    if (!c) return SMW_ERR_INVALID_ARG;
    cudaSetDevice(c->device);
    return launch_diag_mac_rate(c, mac_per_s, warp_instr_per_clk_per_sm);
}

int32_t smw_diag_fp64_rate(smw_ctx* c, double* dfma_per_s, double* warp_instr_per_clk_per_sm) {
    if (!c) return SMW_ERR_INVALID_ARG;
    cudaSetDevice(c->device);
    return launch_diag_rate(c, 1, dfma_per_s, warp_instr_per_clk_per_sm);
}
int32_t smw_diag_pipe_rate(smw_ctx* c, int32_t fp64, double* out5) {      // out: lanes/s, warp-instr/clk/SM, SM Hz during the measurement, SM count, nominal max SM Hz
    if (!c || !out5) return SMW_ERR_INVALID_ARG;
    cudaSetDevice(c->device);
    return launch_diag_rate_ex(c, fp64 ? 1 : 0, out5, out5 + 1, out5 + 2, out5 + 3, out5 + 4);
}
int32_t smw_diag_perm_fp64_ops(const smw_params* p, uint64_t* ops_out) {       // 0 when this parameter set runs on the integer multiplier
    if (!p || !ops_out) return SMW_ERR_INVALID_ARG;
    *ops_out = (p->d_tables52 && p->ctx->use_f52) ? count_fp64_ops(p->h_prog, p->perm_first, p->perm_count) : 0;
    return SMW_OK;
}
int32_t smw_diag_f52_dot(smw_ctx* c, const void* a, const void* b, size_t n_out, int32_t k, void* out, int32_t mem) {
    if (!c || k < 1 || k > 200 || (n_out && (!a || !b || !out))) return SMW_ERR_INVALID_ARG;
    cudaSetDevice(c->device);
    In x(c, a, n_out * k, mem), y(c, b, n_out * k, mem); if (x.rc) return x.rc; if (y.rc) return y.rc;
    Out o(c, out, n_out, mem); if (o.rc) return o.rc;
    int rc = launch_f52_dot(c, x.d, y.d, n_out, k, o.d); if (rc) return rc;
    return o.finish();
}
int32_t smw_diag_perm_macs(const smw_params* p, uint64_t* macs_out) {
    if (!p || !macs_out) return SMW_ERR_INVALID_ARG;
    *macs_out = count_macs(p->h_prog, p->perm_first, p->perm_count);
    return SMW_OK;
}

// ---------------------------------------------------------------------------------------- field
int32_t smw_fr_from_canonical(smw_ctx* c, const void* in, void* out, size_t n, int32_t mem) {
    if (!c) return SMW_ERR_INVALID_ARG;
    cudaSetDevice(c->device);
    In i(c, in, n, mem); if (i.rc) return i.rc; Out o(c, out, n, mem); if (o.rc) return o.rc;
    int rc = launch_to_mont(c, i.d, o.d, n, 0); if (rc) return rc;
    return o.finish();
}
int32_t smw_fr_to_canonical(smw_ctx* c, const void* in, void* out, size_t n, int32_t mem) {
    if (!c) return SMW_ERR_INVALID_ARG;
    cudaSetDevice(c->device);
    In i(c, in, n, mem); if (i.rc) return i.rc; Out o(c, out, n, mem); if (o.rc) return o.rc;
    int rc = launch_to_mont(c, i.d, o.d, n, 1); if (rc) return rc;
    return o.finish();
}
int32_t smw_root_of_unity(smw_ctx* c, int32_t log_n, void* out_host) {
    if (!c || !out_host) return SMW_ERR_INVALID_ARG;
    cudaSetDevice(c->device);
    Out o(c, out_host, 1, SMW_HOST); if (o.rc) return o.rc;
    int rc = launch_root_of_unity(c, log_n, o.d); if (rc) return rc;
    return o.finish();
}

// ---------------------------------------------------------------------------------------- params
int32_t smw_params_from_seed(smw_ctx* c, const uint8_t* seed, size_t seed_len, int32_t t, int32_t rp, smw_params** out) {
    if (!c || !out || (!seed && seed_len)) return SMW_ERR_INVALID_ARG;
    cudaSetDevice(c->device);
    if (rp_for_width(t) < 0) return set_err(c, SMW_ERR_UNSUPPORTED_WIDTH, "unsupported Poseidon width");
    if (rp <= 0 || rp > 96 || (rp & 1)) return set_err(c, SMW_ERR_INVALID_ARG, "partial round count must be even and <= 96 (two-round blocks; lazy-reduction bound of the sparse form)");
    smw_params* p = new smw_params(); p->ctx = c; p->t = t; p->rp = rp;
    int rc = launch_derive_params(c, p, seed, (int)seed_len);
    if (rc) { free_params_internal(p); return rc; }
    *out = p; return SMW_OK;
}
int32_t smw_params_for_width(smw_ctx* c, int32_t t, smw_params** out) {
    if (!c || !out) return SMW_ERR_INVALID_ARG;
    int rp = rp_for_width(t);
    if (rp < 0) return set_err(c, SMW_ERR_UNSUPPORTED_WIDTH, "unsupported Poseidon width t; supported t in {9, 17, 33, 65, 129}");
    auto it = c->by_width.find(t);
    if (it != c->by_width.end()) { *out = it->second; return SMW_OK; }
    uint8_t seed[25]; memcpy(seed, "POSEIDON-PALLAS-T", 17);                 // poseidon/src/lib.rs:168-174 seed_for_t
    for (int i = 0; i < 8; i++) seed[17 + i] = (uint8_t)((uint64_t)t >> (8 * i));
    smw_params* p; int rc = smw_params_from_seed(c, seed, 25, t, rp, &p); if (rc) return rc;
    p->owned_by_ctx = true; c->by_width[t] = p; *out = p; return SMW_OK;
}
int32_t smw_params_for_arity(smw_ctx* c, int32_t arity, smw_params** out) {
    if (!c || !out) return SMW_ERR_INVALID_ARG;
    int t = width_for_arity(arity);
    if (t < 0) return set_err(c, SMW_ERR_UNSUPPORTED_ARITY, "unsupported Merkle arity; max supported = 128");
    return smw_params_for_width(c, t, out);
}
int32_t smw_params_commitment_default(smw_ctx* c, smw_params** out) {
    if (!c || !out) return SMW_ERR_INVALID_ARG;
    if (!c->commitment_default) {
        smw_params* p; int rc = smw_params_from_seed(c, (const uint8_t*)"POSEIDON-T17-X5-SEED", 20, 17, 64, &p); if (rc) return rc;
        p->owned_by_ctx = true; c->commitment_default = p;
    }
    *out = c->commitment_default; return SMW_OK;
}
int32_t smw_params_transcript_default(smw_ctx* c, smw_params** out) {
    if (!c || !out) return SMW_ERR_INVALID_ARG;
    if (!c->transcript_default) {
        smw_params* p; int rc = smw_params_from_seed(c, (const uint8_t*)"POSEIDON-T17-X5-TRANSCRIPT", 26, 17, 64, &p); if (rc) return rc;
        p->owned_by_ctx = true; c->transcript_default = p;
    }
    *out = c->transcript_default; return SMW_OK;
}
int32_t smw_params_from_tables(smw_ctx* c, int32_t t, int32_t rp, const void* mds, const void* rcf, const void* rcp, int32_t mem, smw_params** out) {
    if (!c || !out || !mds || !rcf || !rcp) return SMW_ERR_INVALID_ARG;
    cudaSetDevice(c->device);
    if (rp_for_width(t) < 0) return set_err(c, SMW_ERR_UNSUPPORTED_WIDTH, "unsupported Poseidon width");
    if (rp <= 0 || rp > 96 || (rp & 1)) return set_err(c, SMW_ERR_INVALID_ARG, "partial round count must be even and <= 96 (two-round blocks; lazy-reduction bound of the sparse form)");
    In a(c, mds, (size_t)t * t, mem), b(c, rcf, (size_t)8 * t, mem), d(c, rcp, (size_t)rp, mem);
    if (a.rc) return a.rc; if (b.rc) return b.rc; if (d.rc) return d.rc;
    smw_params* p = new smw_params(); p->ctx = c; p->t = t; p->rp = rp;
    int rc = launch_upload_params(c, p, a.d, b.d, d.d);
    if (rc) { free_params_internal(p); return rc; }
    *out = p; return SMW_OK;
}
int32_t smw_params_export(smw_params* p, void* mds, void* rcf, void* rcp) {
    if (!p) return SMW_ERR_INVALID_ARG;
    smw_ctx* c = p->ctx; cudaSetDevice(c->device);
    size_t n1 = (size_t)p->t * p->t, n2 = (size_t)8 * p->t, n3 = p->rp;
    if (mds) SMW_CUDA(c, cudaMemcpyAsync(mds, p->d_ext, n1 * FR, cudaMemcpyDeviceToHost, c->stream));
    if (rcf) SMW_CUDA(c, cudaMemcpyAsync(rcf, p->d_ext + n1 * 8, n2 * FR, cudaMemcpyDeviceToHost, c->stream));
    if (rcp) SMW_CUDA(c, cudaMemcpyAsync(rcp, p->d_ext + (n1 + n2) * 8, n3 * FR, cudaMemcpyDeviceToHost, c->stream));
    SMW_CUDA(c, cudaStreamSynchronize(c->stream));
    return SMW_OK;
}
int32_t smw_params_width(const smw_params* p) { return p ? p->t : -1; }
int32_t smw_params_partial_rounds(const smw_params* p) { return p ? p->rp : -1; }
void smw_params_free(smw_params* p) { if (!p || p->owned_by_ctx) return; cudaSetDevice(p->ctx->device); cudaStreamSynchronize(p->ctx->stream); free_params_internal(p); }

// ---------------------------------------------------------------------------------------- permutation / sponges
static int permute_impl(smw_ctx* c, const smw_params* p, void* states, size_t batch, int mem, bool sparse) {
    if (!c || !p || (!states && batch)) return SMW_ERR_INVALID_ARG;
    cudaSetDevice(c->device);
    size_t n = batch * p->t;
    if (mem == SMW_DEVICE) return launch_permute_batch(c, p, (uint32_t*)states, batch, sparse);
    In i(c, states, n, SMW_HOST); if (i.rc) return i.rc;
    int rc = launch_permute_batch(c, p, i.owned, batch, sparse); if (rc) return rc;
    return d2h(c, states, i.owned, n * FR);
}
int32_t smw_poseidon_permute(smw_ctx* c, const smw_params* p, void* states, size_t batch, int32_t mem) { return permute_impl(c, p, states, batch, mem, true); }
int32_t smw_poseidon_permute_dense(smw_ctx* c, const smw_params* p, void* states, size_t batch, int32_t mem) { return permute_impl(c, p, states, batch, mem, false); }
int32_t smw_hash_with_ds_dynamic(smw_ctx* c, const smw_params* p, const void* ds, size_t n_ds, const void* inputs, size_t n, void* out, int32_t mem) {
    if (!c || !p || !out) return SMW_ERR_INVALID_ARG;
    cudaSetDevice(c->device);
    In a(c, ds, n_ds, mem), b(c, inputs, n, mem); if (a.rc) return a.rc; if (b.rc) return b.rc;
    Out o(c, out, 1, mem); if (o.rc) return o.rc;
    int rc = launch_hash_ds_single(c, p, a.d, (int)n_ds, b.d, n, o.d); if (rc) return rc;
    return o.finish();
}
int32_t smw_hash_with_ds(smw_ctx* c, const smw_params* p, const void* inputs, size_t n, const void* ds_tag, void* out, int32_t mem) {
    if (!c || !p || !out || !ds_tag) return SMW_ERR_INVALID_ARG;
    cudaSetDevice(c->device);
    In a(c, inputs, n, mem), b(c, ds_tag, 1, mem); if (a.rc) return a.rc; if (b.rc) return b.rc;
    Out o(c, out, 1, mem); if (o.rc) return o.rc;
    int rc = launch_hash_legacy_single(c, p, a.d, n, b.d, o.d); if (rc) return rc;
    return o.finish();
}

// ---------------------------------------------------------------------------------------- Merkle
static int check_tree_args(smw_ctx* c, smw_params* p, int arity, size_t n) {
    if (!c || !p) return SMW_ERR_INVALID_ARG;
    if (n == 0) return set_err(c, SMW_ERR_EMPTY, "no leaves");
    if (arity < 1 || arity > 128) return set_err(c, SMW_ERR_UNSUPPORTED_ARITY, "unsupported Merkle arity; max supported = 128");
    if (width_for_arity(arity) != p->t) return set_err(c, SMW_ERR_ARITY_WIDTH, "arity incompatible with Poseidon width t");
    if (arity == 1 && n > 1) return set_err(c, SMW_ERR_UNSUPPORTED_ARITY, "arity 1 only admits a one-leaf tree (the reference does not terminate here)");
    return SMW_OK;
}
int32_t smw_merkle_new(smw_ctx* c, smw_params* p, int32_t arity, uint64_t label, const void* leaves, size_t n, int32_t mem, smw_tree** out) {
    if (!out) return SMW_ERR_INVALID_ARG;
    int rc = check_tree_args(c, p, arity, n); if (rc) return rc;
    if (!leaves) return SMW_ERR_INVALID_ARG;
    cudaSetDevice(c->device);
    // the tree owns a copy of the leaves, as MerkleTree::new takes the Vec by value (merkle/src/lib.rs:147)
    uint32_t* d0; SMW_CUDA(c, cudaMalloc(&d0, n * FR));
    { cudaError_t e = cudaMemcpyAsync(d0, leaves, n * FR, mem == SMW_DEVICE ? cudaMemcpyDeviceToDevice : cudaMemcpyHostToDevice, c->stream);
      if (e != cudaSuccess) { cudaFree(d0); return cuda_err(c, e, "copy leaves"); } }
    rc = build_tree(c, p, arity, label, d0, true, n, out);
    if (!rc && mem == SMW_HOST) SMW_CUDA(c, cudaStreamSynchronize(c->stream));      // the caller may reuse its host buffer as soon as the call returns
    return rc;
}
int32_t smw_merkle_new_pairs(smw_ctx* c, smw_params* p, int32_t arity, uint64_t label, const void* f, const void* cp, size_t n_f, size_t n_cp, int32_t mem, smw_tree** out) {
    if (!out || !c) return SMW_ERR_INVALID_ARG;
    if (n_f != n_cp) return set_err(c, SMW_ERR_LENGTH_MISMATCH, "f and cp length mismatch");
    int rc = check_tree_args(c, p, arity, n_f); if (rc) return rc;
    if (!f || !cp) return SMW_ERR_INVALID_ARG;
    cudaSetDevice(c->device);
    In a(c, f, n_f, mem), b(c, cp, n_f, mem); if (a.rc) return a.rc; if (b.rc) return b.rc;
    uint32_t* d0; SMW_CUDA(c, cudaMalloc(&d0, n_f * FR));
    rc = launch_pair_leaves(c, p, a.d, b.d, n_f, arity, label, 0, d0, 1);
    if (rc) { cudaFree(d0); return rc; }
    return build_tree(c, p, arity, label, d0, true, n_f, out);
}
int32_t smw_tree_root(smw_tree* t, void* out_host) { if (!t || !out_host) return SMW_ERR_INVALID_ARG; cudaSetDevice(t->ctx->device); return d2h(t->ctx, out_host, t->lv.back(), FR); }
int32_t smw_tree_num_levels(const smw_tree* t) { return t ? (int32_t)t->lv.size() : -1; }
size_t smw_tree_level_len(const smw_tree* t, int32_t level) { return (t && level >= 0 && level < (int)t->len.size()) ? t->len[level] : 0; }
int32_t smw_tree_level(smw_tree* t, int32_t level, void* out, int32_t mem) {
    if (!t || !out || level < 0 || level >= (int)t->lv.size()) return SMW_ERR_INVALID_ARG;
    smw_ctx* c = t->ctx; cudaSetDevice(c->device);
    SMW_CUDA(c, cudaMemcpyAsync(out, t->lv[level], t->len[level] * FR, mem == SMW_DEVICE ? cudaMemcpyDeviceToDevice : cudaMemcpyDeviceToHost, c->stream));
    SMW_CUDA(c, cudaStreamSynchronize(c->stream));
    return SMW_OK;
}
int32_t smw_tree_arity(const smw_tree* t) { return t ? t->arity : -1; }
void smw_tree_free(smw_tree* t) {
    if (!t) return;
    cudaSetDevice(t->ctx->device); cudaStreamSynchronize(t->ctx->stream);
    for (size_t i = 0; i < t->lv.size(); i++) if (t->owned[i]) cudaFree(t->lv[i]);
    smw::pool_free(t->ctx, t->arena);
    delete t;
}
int32_t smw_merkle_commit_columns(smw_ctx* c, smw_params* p, int32_t arity, uint64_t label0, uint64_t label_stride,
                                  const void* cols, size_t n, int32_t n_cols, int32_t mem, void* roots_out_host) {
    int rc = check_tree_args(c, p, arity, n); if (rc) return rc;
    if (!cols || !roots_out_host || n_cols <= 0) return SMW_ERR_INVALID_ARG;
    cudaSetDevice(c->device);
    In in(c, cols, n * (size_t)n_cols, mem); if (in.rc) return in.rc;
    if (n == 1) return d2h(c, roots_out_host, in.d, (size_t)n_cols * FR);   // a 1-leaf tree's root is the leaf
    // two ping-pong buffers sized for level 1 (context scratch slots: no allocation in steady state)
    size_t n1 = (n + arity - 1) / arity, n2 = (n1 + arity - 1) / arity;
    void *pa, *pb = nullptr;
    rc = ctx_scratch(c, smw_ctx::SLOT_TREE_A, n1 * (size_t)n_cols * FR, &pa); if (rc) return rc;
    if (n1 > 1) { rc = ctx_scratch(c, smw_ctx::SLOT_TREE_B, n2 * (size_t)n_cols * FR, &pb); if (rc) return rc; }
    uint32_t *bufA = (uint32_t*)pa, *bufB = (uint32_t*)pb;
    const uint32_t* cur = in.d; size_t cn = n, cstride = n; uint32_t level = 0; uint32_t* dst = bufA;
    while (cn > 1) {
        size_t nn = (cn + arity - 1) / arity;
        rc = launch_merkle_level(c, p, cur, cn, dst, nn, arity, level, label0, label_stride, 0, n_cols, cstride, nn);
        if (rc) break;
        cur = dst; cn = nn; cstride = nn; level++;
        dst = (dst == bufA) ? bufB : bufA;
    }
    if (!rc) rc = d2h(c, roots_out_host, cur, (size_t)n_cols * FR);
    return rc;
}
int32_t smw_commitment_commit(smw_ctx* c, uint64_t ds_tag, smw_params* params, const void* leaves, size_t n, int32_t mem, smw_tree** out) {
    if (!c) return SMW_ERR_INVALID_ARG;
    if (!params) { int rc = smw_params_commitment_default(c, &params); if (rc) return rc; }
    return smw_merkle_new(c, params, 16, ds_tag, leaves, n, mem, out);     // commitment/src/lib.rs:69-77: arity 16, tree_label = ds_tag
}

// ---------------------------------------------------------------------------------------- NTT / LDE
int32_t smw_ntt(smw_ctx* c, void* buf, int32_t log_n, int32_t inverse, int32_t n_cols, int32_t mem) {
    if (!c || !buf || n_cols < 0 || log_n < 0 || log_n > 32) return SMW_ERR_INVALID_ARG;
    cudaSetDevice(c->device);
    size_t n = (size_t)1 << log_n;
    if (mem == SMW_DEVICE) return launch_ntt(c, (uint32_t*)buf, log_n, inverse, n_cols, n);
    In i(c, buf, n * n_cols, SMW_HOST); if (i.rc) return i.rc;
    int rc = launch_ntt(c, i.owned, log_n, inverse, n_cols, n); if (rc) return rc;
    return d2h(c, buf, i.owned, n * n_cols * FR);
}
// ---- building blocks of the cross-GPU four-step NTT; host orchestration in stark_mlwe_b200/shard.py
int32_t smw_scale_powers(smw_ctx* c, void* buf, size_t n, const void* g_host, int32_t mem) {
    if (!c || (n && !buf) || !g_host) return SMW_ERR_INVALID_ARG;
    cudaSetDevice(c->device);
    In g(c, g_host, 1, SMW_HOST); if (g.rc) return g.rc;
    if (mem == SMW_DEVICE) { int rc = launch_scale_powers(c, (uint32_t*)buf, n, g.d); if (rc) return rc; SMW_CUDA(c, cudaStreamSynchronize(c->stream)); return SMW_OK; }
    In i(c, buf, n, SMW_HOST); if (i.rc) return i.rc;
    int rc = launch_scale_powers(c, i.owned, n, g.d); if (rc) return rc;
    return d2h(c, buf, i.owned, n * FR);
}
int32_t smw_dft_strided(smw_ctx* c, void* buf, int32_t log_g, size_t stride, size_t count, int32_t mem) {
    if (!c || !buf || log_g < 0 || log_g > 4 || count > stride) return SMW_ERR_INVALID_ARG;
    cudaSetDevice(c->device);
    const size_t G = (size_t)1 << log_g, total = (G - 1) * stride + count;
    int rc = SMW_OK;
    DevFree fr{{}, c->stream};
    uint32_t* d_w = nullptr;
    rc = make_wpow(c, log_g, &d_w, fr); if (rc) return rc;
    if (mem == SMW_DEVICE) return launch_dft_strided(c, (uint32_t*)buf, log_g, stride, count, d_w);      // DevFree drains the stream before freeing
    In i(c, buf, total, SMW_HOST); if (i.rc) return i.rc;
    rc = launch_dft_strided(c, i.owned, log_g, stride, count, d_w); if (rc) return rc;
    return d2h(c, buf, i.owned, total * FR);
}
int32_t smw_lde(smw_ctx* c, const void* in, int32_t log_n, int32_t log_blowup, const void* shift_host, void* out, int32_t n_cols, int32_t mem) {
    if (!c || !in || !out || !shift_host || n_cols < 0 || log_n < 0 || log_blowup < 0 || log_n + log_blowup > 32) return SMW_ERR_INVALID_ARG;
    cudaSetDevice(c->device);
    size_t n = (size_t)1 << log_n, N = n << log_blowup;
    In i(c, in, n * n_cols, mem), g(c, shift_host, 1, SMW_HOST); if (i.rc) return i.rc; if (g.rc) return g.rc;
    Out o(c, out, N * n_cols, mem); if (o.rc) return o.rc;
    int rc = launch_lde(c, i.d, log_n, log_blowup, g.d, o.d, n_cols); if (rc) return rc;
    return o.finish();
}
// Streams the columns through the device in chunks whose extended form fits `SMW_LDE_CHUNK_BYTES` (default 16 GiB; every chunk pays ~20 ms of latency-bound top-of-tree levels, so chunks are kept large):
// cfg4-sized inputs (2^24 rows x 32 columns per GPU, blow-up 8 = 137 GB extended) never exist in full, only the roots do.
// With host input the H2D copy of chunk i+1 runs on a second stream while chunk i is extended and committed.
int32_t smw_lde_commit(smw_ctx* c, const void* cols, int32_t log_n, int32_t log_blowup, const void* shift_host, int32_t n_cols,
                       uint64_t label0, int32_t mem, void* roots_out_host) {
    if (!c || !cols || !roots_out_host || !shift_host || n_cols <= 0 || log_n < 0 || log_blowup < 0 || log_n + log_blowup > 32) return SMW_ERR_INVALID_ARG;
    cudaSetDevice(c->device);
    Range rg("lde_commit");
    const size_t n = (size_t)1 << log_n, N = n << log_blowup;
    smw_params* p; int rc = smw_params_commitment_default(c, &p); if (rc) return rc;
    In g(c, shift_host, 1, SMW_HOST); if (g.rc) return g.rc;
    size_t budget = (size_t)16 << 30;
    if (const char* e = getenv("SMW_LDE_CHUNK_BYTES")) { size_t v = strtoull(e, nullptr, 10); if (v) budget = v; }
    const int chunk = (int)std::max<size_t>(1, std::min<size_t>((size_t)n_cols, budget / (N * FR)));
    const bool host = (mem == SMW_HOST);
    // column ranges processed one after the other.  Host input: the copy of range i + 1 runs under the compute of range i, so
    // a short first range (1/16 of the columns) lets the kernels start after a small copy instead of after the whole input.
    std::vector<std::pair<int, int>> ranges;      // (first column, count)
    for (int c0 = 0; c0 < n_cols; c0 += chunk) ranges.push_back({c0, std::min(chunk, n_cols - c0)});
    // The split is geometric (1/16, 3/16, the rest): the copy of a range must fit under the compute of the one before it, and
    // with a single 15/16 remainder it did not when the host link delivers less than ~40 GB/s (e2e 50 ms behind `value` on cfg2).
    if (host && n_cols >= 16 && ranges[0].second > n_cols / 16) {
        const int first = ranges[0].second, head = n_cols / 16, mid = std::min(first - head, 3 * head), rest = first - head - mid;
        std::vector<std::pair<int, int>> split = {{0, head}, {head, mid}};
        if (rest > 0) split.push_back({head + mid, rest});
        ranges.erase(ranges.begin()); ranges.insert(ranges.begin(), split.begin(), split.end());
    }
    const int n_chunks = (int)ranges.size();
    uint32_t* d_ext = nullptr; uint32_t* d_in[2] = {nullptr, nullptr};
    cudaStream_t copy = nullptr; cudaEvent_t copied[2] = {nullptr, nullptr}, consumed[2] = {nullptr, nullptr};
    auto cleanup = [&]() {
        cudaStreamSynchronize(c->stream);
        if (copy) { cudaStreamSynchronize(copy); cudaStreamDestroy(copy); }
        for (int i = 0; i < 2; i++) { if (copied[i]) cudaEventDestroy(copied[i]); if (consumed[i]) cudaEventDestroy(consumed[i]); }
    };
    { void* q; rc = ctx_scratch(c, smw_ctx::SLOT_EXT, N * (size_t)chunk * FR, &q); if (rc) return rc; d_ext = (uint32_t*)q; }
    cudaError_t e = cudaSuccess;
    if (host) {
        e = cudaStreamCreateWithFlags(&copy, cudaStreamNonBlocking);
        for (int i = 0; i < 2 && e == cudaSuccess; i++) {
            int widest = 0; for (int ci = i; ci < n_chunks; ci += 2) widest = std::max(widest, ranges[ci].second);      // ranges alternate between the two buffers
            if (widest) { void* q; if (ctx_scratch(c, smw_ctx::SLOT_IN0 + i, n * (size_t)widest * FR, &q)) e = cudaErrorMemoryAllocation; d_in[i] = (uint32_t*)q; }
            if (e == cudaSuccess) e = cudaEventCreateWithFlags(&copied[i], cudaEventDisableTiming);
            if (e == cudaSuccess) e = cudaEventCreateWithFlags(&consumed[i], cudaEventDisableTiming);
        }
        if (e != cudaSuccess) { cleanup(); return cuda_err(c, e, "smw_lde_commit staging"); }
    }
    auto enqueue_copy = [&](int ci) -> cudaError_t {      // host columns of range ci -> d_in[ci & 1] on the copy stream
        const int c0 = ranges[ci].first, nc = ranges[ci].second;
        cudaError_t er = cudaSuccess;
        if (ci >= 2) er = cudaStreamWaitEvent(copy, consumed[ci & 1], 0);
        if (er == cudaSuccess) er = cudaMemcpyAsync(d_in[ci & 1], (const uint8_t*)cols + (size_t)c0 * n * FR, (size_t)nc * n * FR, cudaMemcpyHostToDevice, copy);
        if (er == cudaSuccess) er = cudaEventRecord(copied[ci & 1], copy);
        return er;
    };
    if (host) e = enqueue_copy(0);
    for (int ci = 0; ci < n_chunks && !rc && e == cudaSuccess; ci++) {
        const int c0 = ranges[ci].first, nc = ranges[ci].second;
        const uint32_t* src;
        if (host) {
            if (ci + 1 < n_chunks) e = enqueue_copy(ci + 1);
            if (e == cudaSuccess) e = cudaStreamWaitEvent(c->stream, copied[ci & 1], 0);
            if (e != cudaSuccess) break;
            src = d_in[ci & 1];
        } else src = (const uint32_t*)cols + (size_t)c0 * n * 8;
        rc = launch_lde(c, src, log_n, log_blowup, g.d, d_ext, nc);
        if (host && !rc) e = cudaEventRecord(consumed[ci & 1], c->stream);
        if (!rc && e == cudaSuccess) rc = smw_merkle_commit_columns(c, p, 16, label0 + (uint64_t)c0, 1, d_ext, N, nc, SMW_DEVICE, (uint8_t*)roots_out_host + (size_t)c0 * FR);
    }
    if (!rc && e != cudaSuccess) rc = cuda_err(c, e, "smw_lde_commit: staged host copy (cudaMemcpyAsync / cudaStreamWaitEvent / cudaEventRecord)");
    cleanup();
    return rc;
}

// ---------------------------------------------------------------------------------------- transcript
// immediates of one transcript call -> the transcript's own device scratch (64 words) or, for a very long label, a temporary
static int tr_words(smw_transcript* t, const std::vector<Word>& w, uint32_t** d, bool* temp) {
    smw_ctx* c = t->ctx; *temp = false; *d = t->d_imm;
    if (w.empty()) return SMW_OK;
    if (w.size() > 64) { *temp = true; return upload_words(c, w, d); }
    SMW_CUDA(c, cudaMemcpyAsync(t->d_imm, w.data(), w.size() * FR, cudaMemcpyHostToDevice, c->stream));      // pageable source: staged before the call returns
    return SMW_OK;
}
int32_t smw_transcript_new(smw_ctx* c, const uint8_t* label, size_t len, smw_params* params, smw_transcript** out) {
    if (!c || !out || (!label && len)) return SMW_ERR_INVALID_ARG;
    cudaSetDevice(c->device);
    smw_params* p; int rc = get_transcript_params(c, params, &p); if (rc) return rc;
    auto t = std::make_unique<smw_transcript>(); t->ctx = c; t->params = p;
    SMW_CUDA(c, cudaMalloc(&t->d_state, (17 + 64 + 1) * FR));
    t->d_imm = t->d_state + 17 * 8; t->d_out = t->d_imm + 64 * 8;
    std::vector<Word> A; A.push_back(tag_word(TAG_INIT)); push_absorb_bytes(A, label, len);
    uint32_t* dA; bool temp; rc = tr_words(t.get(), A, &dA, &temp); if (rc) { cudaFree(t->d_state); return rc; }
    rc = launch_transcript_run(c, p, t->d_state, 1, 0, dA, (int)A.size(), nullptr, 0, nullptr, 0, 0, nullptr);
    if (temp) { cudaStreamSynchronize(c->stream); cudaFree(dA); }
    if (rc) { cudaStreamSynchronize(c->stream); cudaFree(t->d_state); return rc; }
    t->pos = advance_pos(0, A.size() - 1);
    *out = t.release(); return SMW_OK;
}
int32_t smw_transcript_absorb_bytes(smw_transcript* t, const uint8_t* bytes, size_t len) {
    if (!t || (!bytes && len)) return SMW_ERR_INVALID_ARG;
    smw_ctx* c = t->ctx; cudaSetDevice(c->device);
    std::vector<Word> A; push_absorb_bytes(A, bytes, len);
    uint32_t* dA; bool temp; int rc = tr_words(t, A, &dA, &temp); if (rc) return rc;
    rc = launch_transcript_run(c, t->params, t->d_state, 0, t->pos, dA, (int)A.size(), nullptr, 0, nullptr, 0, 0, nullptr);
    if (temp) { cudaStreamSynchronize(c->stream); cudaFree(dA); }
    if (!rc) t->pos = advance_pos(t->pos, A.size());
    return rc;
}
int32_t smw_transcript_absorb_fields(smw_transcript* t, const void* frs, size_t n, int32_t mem) {
    if (!t || (!frs && n)) return SMW_ERR_INVALID_ARG;
    if (n == 0) return SMW_OK;
    smw_ctx* c = t->ctx; cudaSetDevice(c->device);
    In i(c, frs, n, mem); if (i.rc) return i.rc;
    int rc = n >= COOP_ABSORB_MIN ? launch_transcript_absorb_coop(c, t->params, t->d_state, t->pos, i.d, n)
                                  : launch_transcript_run(c, t->params, t->d_state, 0, t->pos, nullptr, 0, i.d, n, nullptr, 0, 0, nullptr);
    if (!rc) t->pos = advance_pos(t->pos, n);
    return rc;
}
int32_t smw_transcript_absorb_field(smw_transcript* t, const void* fr_host) { return smw_transcript_absorb_fields(t, fr_host, 1, SMW_HOST); }
int32_t smw_transcript_challenge(smw_transcript* t, const uint8_t* label, size_t len, void* out_host) {
    if (!t || !out_host || (!label && len)) return SMW_ERR_INVALID_ARG;
    smw_ctx* c = t->ctx; cudaSetDevice(c->device);
    std::vector<Word> A; push_challenge(A, label, len);
    uint32_t* dA; bool temp; int rc = tr_words(t, A, &dA, &temp); if (rc) return rc;
    rc = launch_transcript_run(c, t->params, t->d_state, 0, t->pos, dA, (int)A.size(), nullptr, 0, nullptr, 0, 1, t->d_out);
    if (!rc) rc = d2h(c, out_host, t->d_out, FR);          // the one synchronisation: the challenge goes to the host
    if (temp) { cudaStreamSynchronize(c->stream); cudaFree(dA); }
    if (!rc) t->pos = 0;
    return rc;
}
int32_t smw_transcript_state(smw_transcript* t, void* out_17, int32_t* pos_out) {
    if (!t || !out_17) return SMW_ERR_INVALID_ARG;
    cudaSetDevice(t->ctx->device);
    if (pos_out) *pos_out = t->pos;
    return d2h(t->ctx, out_17, t->d_state, 17 * FR);
}
void smw_transcript_free(smw_transcript* t) { if (!t) return; cudaSetDevice(t->ctx->device); cudaStreamSynchronize(t->ctx->stream); cudaFree(t->d_state); delete t; }
int32_t smw_tr_hash_fields_tagged(smw_ctx* c, const uint8_t* tag, size_t tag_len, const void* xs, size_t n, int32_t mem, void* out_host) {
    if (!c || !out_host || (!tag && tag_len) || (!xs && n)) return SMW_ERR_INVALID_ARG;
    cudaSetDevice(c->device);
    smw_params* p; int rc = smw_params_transcript_default(c, &p); if (rc) return rc;
    In i(c, xs, n, mem); if (i.rc) return i.rc;
    Out o(c, out_host, 1, SMW_HOST); if (o.rc) return o.rc;
    rc = oneshot_hash(c, p, "FRI/FS", tag, tag_len, {}, i.d, n, "out", o.d); if (rc) return rc;
    return o.finish();
}
// `count` one-shot hashes with the same tag and k field elements each (k small: the whole pre-image fits one rate block), in ONE
// launch — e.g. the r * L index seeds of fri_prove_queries (deep_ali/src/fri.rs:184-191, 388-400): xs is count x k, out count Fr.
int32_t smw_tr_hash_fields_tagged_batch(smw_ctx* c, const uint8_t* tag, size_t tag_len, const void* xs, int32_t k, size_t count, int32_t mem, void* out) {
    if (!c || (!tag && tag_len) || k < 0 || (count && ((k && !xs) || !out))) return SMW_ERR_INVALID_ARG;
    if (count == 0) return SMW_OK;
    cudaSetDevice(c->device);
    smw_params* p; int rc = smw_params_transcript_default(c, &p); if (rc) return rc;
    std::vector<Word> seq;                                  // transcript/src/lib.rs:55-101 as a lane template
    push_absorb_bytes(seq, (const uint8_t*)"FRI/FS", 6);
    push_absorb_bytes(seq, tag, tag_len);
    const int lane0 = (int)seq.size();
    for (int i = 0; i < k; i++) seq.push_back(Word{});
    push_challenge(seq, (const uint8_t*)"out", 3);
    if (seq.size() > 16) return set_err(c, SMW_ERR_INVALID_ARG, "batched one-shot hash: the pre-image must fit one rate block (use smw_tr_hash_fields_tagged)");
    std::vector<Word> lanes(17, Word{});
    for (size_t i = 0; i < seq.size(); i++) lanes[i] = seq[i];
    lanes[16] = tag_word(TAG_INIT);
    DevFree fr{{}, c->stream};
    uint32_t* d_t; rc = upload_words(c, lanes, &d_t); fr.v.push_back(d_t); if (rc) return rc;
    In i(c, xs, count * (size_t)k, mem); if (i.rc) return i.rc;
    Out o(c, out, count, mem); if (o.rc) return o.rc;
    rc = launch_tr_hash_batch(c, p, d_t, lane0, k, i.d, count, o.d); if (rc) return rc;
    return o.finish();
}
// `count` independent one-shot hashes (e.g. the four ALI sponges of deep_ali_build_f0, deep_ali/src/fri.rs:541-548): each chain of
// launches runs on its own side stream, so the strictly serial sponges overlap each other.  xs[i]: n[i] elements; out: count Fr (host).
int32_t smw_tr_hash_fields_tagged_many(smw_ctx* c, int32_t count, const uint8_t* const* tags, const size_t* tag_lens, const void* const* xs, const size_t* n,
                                       int32_t mem, void* out_host) {
    if (!c || count < 0 || (count && (!tags || !tag_lens || !xs || !n || !out_host))) return SMW_ERR_INVALID_ARG;
    if (count == 0) return SMW_OK;
    cudaSetDevice(c->device);
    smw_params* p; int rc = smw_params_transcript_default(c, &p); if (rc) return rc;
    std::vector<std::unique_ptr<In>> ins;
    for (int i = 0; i < count; i++) { ins.emplace_back(new In(c, xs[i], n[i], mem)); if (ins.back()->rc) return ins.back()->rc; }
    uint32_t* d_out = nullptr; SMW_CUDA(c, cudaMalloc(&d_out, (size_t)count * FR));
    std::vector<void*> defer;
    {
        StreamFan fan(c, FAN_STREAMS);
        for (int i = 0; i < count && !rc; i++) {
            fan.enter(i);
            rc = oneshot_hash(c, p, "FRI/FS", tags[i], tag_lens[i], {}, ins[i]->d, n[i], "out", d_out + (size_t)i * 8, &defer);
            fan.leave();
        }
    }
    if (!rc) rc = d2h(c, out_host, d_out, (size_t)count * FR); else cudaStreamSynchronize(c->stream);
    for (void* q : defer) cudaFree(q);
    cudaFree(d_out);
    return rc;
}
// the 17-lane pre-image template of hash_leaf_pair (deep_ali/src/fri.rs:38-44); returns the lanes of f and s
static void leaf_template(std::vector<Word>& lanes, int* lane_f, int* lane_s) {
    std::vector<Word> seq;
    push_absorb_bytes(seq, (const uint8_t*)"FRI/leaf/poseidon", 17);
    push_absorb_bytes(seq, (const uint8_t*)"FRI/leaf", 8);
    *lane_f = (int)seq.size(); seq.push_back(Word{}); *lane_s = (int)seq.size(); seq.push_back(Word{});
    push_challenge(seq, (const uint8_t*)"leaf", 4);
    lanes.assign(17, Word{});
    for (size_t i = 0; i < seq.size(); i++) lanes[i] = seq[i];      // 9 lanes < rate 16: no intermediate permutation
    lanes[16] = tag_word(TAG_INIT);
}
// d_tmpl: the 17-lane pre-image template already on the device (upload_leaf_template), or nullptr to upload it for this call
static int upload_leaf_template(smw_ctx* c, uint32_t** d_t, int* lf, int* ls) {
    std::vector<Word> lanes; leaf_template(lanes, lf, ls);
    return upload_words(c, lanes, d_t);
}
static int hash_leaves_device(smw_ctx* c, const uint32_t* d_f, const uint32_t* d_s, size_t s_div, size_t n, uint32_t* d_out, const uint32_t* d_tmpl = nullptr) {
    smw_params* p; int rc = smw_params_transcript_default(c, &p); if (rc) return rc;
    int lf, ls;
    if (d_tmpl) { std::vector<Word> lanes; leaf_template(lanes, &lf, &ls); return launch_fri_leaf_hashed(c, p, d_tmpl, lf, ls, d_f, d_s, s_div, n, d_out); }
    uint32_t* d_t; rc = upload_leaf_template(c, &d_t, &lf, &ls); if (rc) return rc;
    rc = launch_fri_leaf_hashed(c, p, d_t, lf, ls, d_f, d_s, s_div, n, d_out);
    cudaStreamSynchronize(c->stream); cudaFree(d_t);
    return rc;
}
int32_t smw_hash_leaf_pair(smw_ctx* c, const void* f, const void* s, size_t n, void* out, int32_t mem) {
    if (!c || (n && (!f || !s || !out))) return SMW_ERR_INVALID_ARG;
    if (n == 0) return SMW_OK;
    cudaSetDevice(c->device);
    In a(c, f, n, mem), b(c, s, n, mem); if (a.rc) return a.rc; if (b.rc) return b.rc;
    Out o(c, out, n, mem); if (o.rc) return o.rc;
    int rc = hash_leaves_device(c, a.d, b.d, 1, n, o.d); if (rc) return rc;
    return o.finish();
}

// ---------------------------------------------------------------------------------------- FRI
int32_t smw_pick_arity_for_layer(size_t n, size_t m) {     // deep_ali/src/fri.rs:220-229
    if (m >= 128 && n % 128 == 0) return 128; if (m >= 64 && n % 64 == 0) return 64; if (m >= 32 && n % 32 == 0) return 32;
    if (m >= 16 && n % 16 == 0) return 16; if (m >= 8 && n % 8 == 0) return 8; if (m >= 4 && n % 4 == 0) return 4;
    if (n % 2 == 0) return 2; return 1;
}
static int sample_z_device(smw_ctx* c, uint64_t seed_z, const std::vector<uint64_t>& levels, const std::vector<uint64_t>& sizes, uint32_t* d_z) {
    smw_params* p; int rc = smw_params_transcript_default(c, &p); if (rc) return rc;
    int L = (int)sizes.size(); if (L == 0) return SMW_OK;
    std::vector<Word> all; all.reserve((size_t)L * 17);
    for (int l = 0; l < L; l++) {
        std::vector<Word> seq;
        push_absorb_bytes(seq, (const uint8_t*)"FRI/FS", 6);
        push_absorb_bytes(seq, (const uint8_t*)"FRI/z/l", 7);
        seq.push_back(word_from_u64(seed_z)); seq.push_back(word_from_u64(levels[l])); seq.push_back(word_from_u64(sizes[l]));
        push_challenge(seq, (const uint8_t*)"out", 3);
        std::vector<Word> lanes(17, Word{});
        for (size_t i = 0; i < seq.size(); i++) lanes[i] = seq[i];
        lanes[16] = tag_word(TAG_INIT);
        all.insert(all.end(), lanes.begin(), lanes.end());
    }
    uint32_t* d_t; uint64_t* d_sizes;
    rc = upload_words(c, all, &d_t); if (rc) return rc;
    std::vector<uint64_t> ls((size_t)2 * L);
    for (int l = 0; l < L; l++) { ls[2 * l] = levels[l]; ls[2 * l + 1] = sizes[l]; }
    SMW_CUDA(c, cudaMalloc(&d_sizes, (size_t)L * 16));
    SMW_CUDA(c, cudaMemcpyAsync(d_sizes, ls.data(), (size_t)L * 16, cudaMemcpyHostToDevice, c->stream));
    rc = launch_fri_sample_z(c, p, d_t, d_sizes, L, seed_z, d_z);
    cudaStreamSynchronize(c->stream); cudaFree(d_t); cudaFree(d_sizes);
    return rc;
}
int32_t smw_fri_sample_z_ell(smw_ctx* c, uint64_t seed_z, uint64_t level, uint64_t domain_size, void* out_host) {
    if (!c || !out_host) return SMW_ERR_INVALID_ARG;
    cudaSetDevice(c->device);
    Out o(c, out_host, 1, SMW_HOST); if (o.rc) return o.rc;
    int rc = sample_z_device(c, seed_z, {level}, {domain_size}, o.d); if (rc) return rc;
    return o.finish();
}
int32_t smw_fri_fold_layer(smw_ctx* c, const void* f, size_t n, const void* z_host, size_t m, void* out, int32_t mem) {
    if (!c || !z_host || (n && (!f || !out))) return SMW_ERR_INVALID_ARG;
    if (m < 2 || m > 128) return set_err(c, SMW_ERR_INVALID_ARG, "fold factor m must be in 2..128");
    if (n % m) return set_err(c, SMW_ERR_NOT_DIVISIBLE, "layer size must be divisible by m");
    cudaSetDevice(c->device);
    In a(c, f, n, mem), z(c, z_host, 1, SMW_HOST); if (a.rc) return a.rc; if (z.rc) return z.rc;
    Out o(c, out, n / m, mem); if (o.rc) return o.rc;
    int rc = launch_fri_fold(c, a.d, n, z.d, m, o.d); if (rc) return rc;
    return o.finish();
}
int32_t smw_compute_s_layer(smw_ctx* c, const void* f, size_t n, const void* z_host, size_t m, void* out, int32_t mem) {
    if (!c || !z_host || (n && (!f || !out))) return SMW_ERR_INVALID_ARG;
    if (m < 1 || m > 128) return set_err(c, SMW_ERR_INVALID_ARG, "fold factor m must be in 1..128");
    if (n % m) return set_err(c, SMW_ERR_NOT_DIVISIBLE, "layer size must be divisible by m");
    cudaSetDevice(c->device);
    In a(c, f, n, mem), z(c, z_host, 1, SMW_HOST); if (a.rc) return a.rc; if (z.rc) return z.rc;
    Out o(c, out, n, mem); if (o.rc) return o.rc;
    uint32_t* d_fold; SMW_CUDA(c, cudaMalloc(&d_fold, (n / m + 1) * FR));
    int rc = launch_fri_fold(c, a.d, n, z.d, m, d_fold);
    if (!rc) rc = launch_broadcast_div(c, d_fold, n, m, o.d);
    if (!rc) rc = o.finish();
    cudaStreamSynchronize(c->stream); cudaFree(d_fold);
    return rc;
}
void smw_fri_free(smw_fri* s) {
    if (!s) return;
    Range rg("fri_free");
    cudaSetDevice(s->ctx->device); cudaStreamSynchronize(s->ctx->stream);
    for (auto* t : s->tree) smw_tree_free(t);
    smw::pool_free(s->ctx, s->arena_f); smw::pool_free(s->ctx, s->arena_leaves); smw::pool_free(s->ctx, s->arena_levels);
    cudaFree(s->d_z);
    delete s;
}
} // extern "C"
namespace {
// The per-layer commitments of a FRI commit phase (deep_ali/src/fri.rs:267-301) for layers [0, K): leaf digests of every layer in
// one launch per leaf kind, then the layer trees as a forest (one launch per level index).  The layers are independent once
// every f_l exists (z_l is public, fri.rs:250).  n_l: elements of layer l held here (the whole layer, or this rank's slice
// starting at global index first_l); s_l[i] = f_{l+1}[i / m_l] is read in place of a materialised s layer (fri.rs:259-265).
struct FriLayers {
    std::vector<ForestTree> trees; uint32_t* arena_leaves = nullptr; uint32_t* arena_levels = nullptr;
};
// one (layer, slice) to commit: n_loc elements of layer `layer` starting at global index `first`; f_next = the matching slice of
// the next layer (nullptr for the last layer: s = 0)
struct LayerEntry { int layer; size_t n_glob, n_loc, m; uint64_t first; const uint32_t* f; const uint32_t* f_next; };
int fri_commit_entries(smw_ctx* c, const std::vector<LayerEntry>& E, uint64_t level0, bool whole_only, FriLayers& out, std::vector<void*>& defer) {
    if (E.empty()) return SMW_OK;
    Range rg("fri.leaves+forest");
    smw_params* trp; int rc = smw_params_transcript_default(c, &trp); if (rc) return rc;
    size_t total = 0; for (auto& e : E) total += e.n_loc;
    SMW_CUDA(c, pool_malloc(c, (void**)&out.arena_leaves, total * FR));
    uint32_t* d_tmpl = nullptr; int lf = 0, ls = 0;
    std::vector<LeafSeg> hashed_segs, pair_segs; uint64_t hashed_total = 0, pair_total = 0; smw_params* pair_p = nullptr; bool pair_uniform = true;
    size_t off = 0;
    for (auto& e : E) {
        const int arity = smw_pick_arity_for_layer(e.n_glob, e.m);                 // chosen from the GLOBAL layer size (fri.rs:272)
        const bool hashed = arity >= 8;                                           // fri.rs:275
        smw_params* p; rc = smw_params_for_arity(c, arity, &p); if (rc) return rc;
        uint32_t* leaves = out.arena_leaves + off * 8; off += e.n_loc;
        const uint64_t label = level0 + (uint64_t)e.layer;
        LeafSeg sg{e.f, e.f_next, leaves, (uint64_t)e.n_loc, label, e.first, (uint64_t)e.m, (uint32_t)arity, 0, 0};
        if (hashed) { sg.first = hashed_total; hashed_total += e.n_loc; hashed_segs.push_back(sg); }
        else { sg.first = pair_total; pair_total += e.n_loc; pair_segs.push_back(sg); if (pair_p && pair_p != p) pair_uniform = false; pair_p = p; }
        out.trees.push_back(ForestTree{p, arity, label, e.first, 0, whole_only, leaves, e.n_loc, {}, {}});
    }
    if (!hashed_segs.empty()) {
        rc = upload_leaf_template(c, &d_tmpl, &lf, &ls); if (rc) return rc;
        defer.push_back(d_tmpl);
        if (has_multi(c, trp)) {
            LeafSeg* d; SMW_CUDA(c, cudaMalloc(&d, hashed_segs.size() * sizeof(LeafSeg))); defer.push_back(d);
            SMW_CUDA(c, cudaMemcpyAsync(d, hashed_segs.data(), hashed_segs.size() * sizeof(LeafSeg), cudaMemcpyHostToDevice, c->stream));
            rc = launch_fri_leaf_hashed_multi(c, trp, d_tmpl, lf, ls, d, (int)hashed_segs.size(), hashed_total); if (rc) return rc;
        } else for (auto& sg : hashed_segs) { rc = launch_fri_leaf_hashed(c, trp, d_tmpl, lf, ls, sg.f, sg.cp, sg.cp_div, sg.n, sg.out); if (rc) return rc; }
    }
    if (!pair_segs.empty()) {
        if (pair_uniform && has_multi(c, pair_p)) {
            LeafSeg* d; SMW_CUDA(c, cudaMalloc(&d, pair_segs.size() * sizeof(LeafSeg))); defer.push_back(d);
            SMW_CUDA(c, cudaMemcpyAsync(d, pair_segs.data(), pair_segs.size() * sizeof(LeafSeg), cudaMemcpyHostToDevice, c->stream));
            rc = launch_pair_leaves_multi(c, pair_p, d, (int)pair_segs.size(), pair_total); if (rc) return rc;
        } else for (auto& sg : pair_segs) {
            smw_params* p; rc = smw_params_for_arity(c, (int)sg.arity, &p); if (rc) return rc;
            rc = launch_pair_leaves(c, p, sg.f, sg.cp, sg.n, (int)sg.arity, sg.label, sg.pos_offset, sg.out, sg.cp_div); if (rc) return rc;
        }
    }
    return build_forest(c, out.trees, &out.arena_levels, defer);
}
int fri_commit_layers(smw_ctx* c, int K, int L, const std::vector<size_t>& n_glob, const std::vector<size_t>& n_loc, const std::vector<size_t>& m,
                      const std::vector<uint64_t>& first, const std::vector<uint32_t*>& f, uint64_t level0, bool whole_only, FriLayers& out, std::vector<void*>& defer) {
    std::vector<LayerEntry> E;
    for (int l = 0; l < K; l++) E.push_back(LayerEntry{l, n_glob[l], n_loc[l], m[l], first[l], f[l], l < L ? f[l + 1] : nullptr});
    return fri_commit_entries(c, E, level0, whole_only, out, defer);
}
int fri_layer_sizes(smw_ctx* c, size_t n0, const uint64_t* schedule, int L, std::vector<size_t>& n, std::vector<size_t>& m) {
    n.assign(L + 1, 0); m.assign(L + 1, 1); n[0] = n0;          // fri.rs:145-155 — the reference asserts divisibility level by level
    for (int l = 0; l < L; l++) {
        m[l] = (size_t)schedule[l];
        if (m[l] < 2 || m[l] > 128) return set_err(c, SMW_ERR_INVALID_ARG, "fold factor m must be in 2..128");
        if (n[l] % m[l]) return set_err(c, SMW_ERR_NOT_DIVISIBLE, "layer size must be divisible by m");
        n[l + 1] = n[l] / m[l];
    }
    return SMW_OK;
}
} // namespace
extern "C" {
int32_t smw_fri_build_transcript(smw_ctx* c, const void* f0, size_t n0, const uint64_t* schedule, int32_t L, uint64_t seed_z, int32_t mem, smw_fri** out) {
    return smw_fri_build_transcript_from(c, f0, n0, schedule, L, seed_z, 0, mem, out);
}
int32_t smw_fri_build_transcript_from(smw_ctx* c, const void* f0, size_t n0, const uint64_t* schedule, int32_t L, uint64_t seed_z, uint64_t level0, int32_t mem, smw_fri** out) {
    if (!c || !out || !f0 || L < 0 || (L && !schedule) || n0 == 0) return SMW_ERR_INVALID_ARG;
    cudaSetDevice(c->device);
    Range rg("fri_build_transcript");
    std::vector<size_t> n, m; int rc = fri_layer_sizes(c, n0, schedule, L, n, m); if (rc) return rc;
    auto s = std::make_unique<smw_fri>(); s->ctx = c; s->L = L; s->n = n; s->m = m;
    s->arity.resize(L + 1); s->hashed.resize(L + 1); s->f.assign(L + 1, nullptr); s->tree.assign(L + 1, nullptr);
    std::vector<void*> defer;
    auto fail = [&](int code) { cudaStreamSynchronize(c->stream); for (void* q : defer) cudaFree(q); smw_fri_free(s.release()); return code; };
    { size_t total = 0; for (int l = 0; l <= L; l++) total += n[l];
      cudaError_t e = pool_malloc(c, (void**)&s->arena_f, total * FR); if (e != cudaSuccess) return fail(cuda_err(c, e, "cudaMalloc(fri layers)"));
      size_t off = 0; for (int l = 0; l <= L; l++) { s->f[l] = s->arena_f + off * 8; off += n[l]; } }
    { cudaError_t e = cudaMalloc(&s->d_z, (size_t)(L + 1) * FR); if (e != cudaSuccess) return fail(cuda_err(c, e, "cudaMalloc")); }
    { cudaError_t e = cudaMemcpyAsync(s->f[0], f0, n0 * FR, mem == SMW_DEVICE ? cudaMemcpyDeviceToDevice : cudaMemcpyHostToDevice, c->stream); if (e != cudaSuccess) return fail(cuda_err(c, e, "copy f0")); }
    // z_l depends only on public values (fri.rs:250): all levels in one launch, then the folds (fri.rs:254)
    std::vector<uint64_t> levels(L), sizes(L);
    for (int l = 0; l < L; l++) { levels[l] = level0 + (uint64_t)l; sizes[l] = (uint64_t)n[l]; }
    { Range r2("fri.sample_z"); rc = sample_z_device(c, seed_z, levels, sizes, s->d_z); if (rc) return fail(rc); }
    { Range r2("fri.folds"); for (int l = 0; l < L; l++) { rc = launch_fri_fold(c, s->f[l], n[l], s->d_z + (size_t)l * 8, m[l], s->f[l + 1]); if (rc) return fail(rc); } }
    FriLayers fl; std::vector<uint64_t> first(L + 1, 0);
    { Range r2("fri.commit_layers"); rc = fri_commit_layers(c, L + 1, L, n, n, m, first, s->f, level0, false, fl, defer); }
    s->arena_leaves = fl.arena_leaves; s->arena_levels = fl.arena_levels;
    if (rc) return fail(rc);
    for (int l = 0; l <= L; l++) { s->arity[l] = fl.trees[l].arity; s->hashed[l] = fl.trees[l].arity >= 8; s->tree[l] = tree_handle(c, fl.trees[l], false); }
    Range r3("fri.drain");
    cudaStreamSynchronize(c->stream);
    for (void* q : defer) cudaFree(q);
    *out = s.release();
    return SMW_OK;
}
int32_t smw_fri_num_layers(const smw_fri* s) { return s ? s->L + 1 : -1; }
int32_t smw_fri_layer_info(const smw_fri* s, int32_t l, size_t* n, size_t* m, int32_t* arity, int32_t* hashed) {
    if (!s || l < 0 || l > s->L) return SMW_ERR_INVALID_ARG;
    if (n) *n = s->n[l]; if (m) *m = s->m[l]; if (arity) *arity = s->arity[l]; if (hashed) *hashed = s->hashed[l];
    return SMW_OK;
}
int32_t smw_fri_layer_f(smw_fri* s, int32_t l, void* out, int32_t mem) {
    if (!s || !out || l < 0 || l > s->L) return SMW_ERR_INVALID_ARG;
    smw_ctx* c = s->ctx; cudaSetDevice(c->device);
    SMW_CUDA(c, cudaMemcpyAsync(out, s->f[l], s->n[l] * FR, mem == SMW_DEVICE ? cudaMemcpyDeviceToDevice : cudaMemcpyDeviceToHost, c->stream));
    SMW_CUDA(c, cudaStreamSynchronize(c->stream)); return SMW_OK;
}
int32_t smw_fri_layer_s(smw_fri* s, int32_t l, void* out, int32_t mem) {
    if (!s || !out || l < 0 || l > s->L) return SMW_ERR_INVALID_ARG;
    smw_ctx* c = s->ctx; cudaSetDevice(c->device);
    Out o(c, out, s->n[l], mem); if (o.rc) return o.rc;
    if (l == s->L) { SMW_CUDA(c, cudaMemsetAsync(o.d, 0, s->n[l] * FR, c->stream)); }
    else { int rc = launch_broadcast_div(c, s->f[l + 1], s->n[l], s->m[l], o.d); if (rc) return rc; }
    return o.finish();
}
int32_t smw_fri_roots(smw_fri* s, void* out_host) {
    if (!s || !out_host) return SMW_ERR_INVALID_ARG;
    for (int l = 0; l <= s->L; l++) { int rc = smw_tree_root(s->tree[l], (uint8_t*)out_host + (size_t)l * FR); if (rc) return rc; }
    return SMW_OK;
}
int32_t smw_fri_z(smw_fri* s, void* out_host) { if (!s || !out_host) return SMW_ERR_INVALID_ARG; if (s->L == 0) return SMW_OK; cudaSetDevice(s->ctx->device); return d2h(s->ctx, out_host, s->d_z, (size_t)s->L * FR); }
smw_tree* smw_fri_layer_tree(smw_fri* s, int32_t l) { return (s && l >= 0 && l <= s->L) ? s->tree[l] : nullptr; }

// ---------------------------------------------------------------------------------------- query phase helpers ("next" tier)
int32_t smw_index_from_seed(smw_ctx* c, const void* seeds, size_t n, uint64_t n_pow2, uint64_t* out_idx_host, int32_t mem) {
    if (!c || (n && (!seeds || !out_idx_host)) || n_pow2 == 0 || (n_pow2 & (n_pow2 - 1))) return SMW_ERR_INVALID_ARG;
    if (n == 0) return SMW_OK;
    cudaSetDevice(c->device);
    In a(c, seeds, n, mem); if (a.rc) return a.rc;
    uint64_t* d_o; SMW_CUDA(c, cudaMalloc(&d_o, n * 8));
    int rc = launch_index_from_seed(c, a.d, n, n_pow2, d_o);
    if (!rc) rc = d2h(c, out_idx_host, d_o, n * 8);
    cudaStreamSynchronize(c->stream); cudaFree(d_o);
    return rc;
}
int32_t smw_ali_sample_z_beta(smw_ctx* c, const void* fused_host, const void* roots_seed_host, uint64_t n0, void* z_out_host, void* beta_out_host) {
    if (!c || !fused_host || !roots_seed_host || !z_out_host || !beta_out_host) return SMW_ERR_INVALID_ARG;
    cudaSetDevice(c->device);
    In a(c, fused_host, 1, SMW_HOST), b(c, roots_seed_host, 1, SMW_HOST); if (a.rc) return a.rc; if (b.rc) return b.rc;
    Out z(c, z_out_host, 1, SMW_HOST), be(c, beta_out_host, 1, SMW_HOST); if (z.rc) return z.rc; if (be.rc) return be.rc;
    int rc = launch_ali_sample(c, a.d, b.d, n0, z.d, be.d); if (rc) return rc;
    rc = z.finish(); if (rc) return rc;
    return be.finish();
}

// ---------------------------------------------------------------------------------------- range-sharded FRI commit phase
// the fold is over contiguous blocks (fri.rs:95-100) and z_l is public (fri.rs:250), so the rank that holds
// f_0[first, first + n_local) folds its own slice through every layer with no exchange, hashes its own leaves and builds its
// own subtree with DS positions offset by the slice start (merkle/src/lib.rs:168-173) up to the last level whose nodes
// are complete inside the slice (the "cap").  The host all-gathers the caps (<= a few KB per layer) and every rank
// finishes the tree tops with smw_merkle_root_from_level; layers too small to split are rebuilt by every rank from the
// gathered slice of the first such layer with smw_fri_build_transcript_from.
struct smw_fri_shard {
    smw_ctx* ctx = nullptr;
    int K = 0;                                   // layers [0, K) are committed from local slices
    std::vector<uint32_t*> cap; std::vector<size_t> cap_n; std::vector<int> cap_level, arity;   // cap[l] points into arena_levels / arena_leaves
    uint32_t* tail = nullptr; size_t tail_n = 0;  // local slice of f_K (empty when every layer was local); points into arena_f
    uint32_t *arena_f = nullptr, *arena_leaves = nullptr, *arena_levels = nullptr;
};
void smw_fri_shard_free(smw_fri_shard* s) {
    if (!s) return;
    cudaSetDevice(s->ctx->device); cudaStreamSynchronize(s->ctx->stream);
    smw::pool_free(s->ctx, s->arena_f); smw::pool_free(s->ctx, s->arena_leaves); smw::pool_free(s->ctx, s->arena_levels);
    delete s;
}
int32_t smw_fri_shard_build(smw_ctx* c, const void* f0_slice, size_t n_local, size_t n0, size_t first_index, const uint64_t* schedule, int32_t L,
                            uint64_t seed_z, int32_t mem, smw_fri_shard** out) {
    if (!c || !out || !f0_slice || L < 0 || (L && !schedule) || n0 == 0 || n_local == 0 || n0 % n_local || first_index % n_local || first_index >= n0)
        return SMW_ERR_INVALID_ARG;
    cudaSetDevice(c->device);
    const size_t G = n0 / n_local;               // number of equal slices
    std::vector<size_t> n, m; int rc = fri_layer_sizes(c, n0, schedule, L, n, m); if (rc) return rc;
    // layer l is local while the slice of f_l folds into a whole slice of f_{l+1} (the last layer needs no fold)
    int K = 0;
    while (K <= L && n[K] % G == 0 && (K == L || (n[K] / G) % m[K] == 0)) K++;
    auto s = std::make_unique<smw_fri_shard>(); s->ctx = c; s->K = K;
    std::vector<void*> defer;
    auto fail = [&](int code) { cudaStreamSynchronize(c->stream); for (void* q : defer) cudaFree(q); smw_fri_shard_free(s.release()); return code; };
    const int n_f = (K <= L) ? K + 1 : K;         // local slices of f_0 .. f_K (f_K only when K <= L)
    std::vector<uint32_t*> f(L + 2, nullptr); std::vector<size_t> n_loc(L + 1, 0); std::vector<uint64_t> first(L + 1, 0);
    { size_t total = 0; for (int l = 0; l < n_f; l++) total += n[l] / G;
      cudaError_t e = pool_malloc(c, (void**)&s->arena_f, total * FR); if (e != cudaSuccess) return fail(cuda_err(c, e, "cudaMalloc(fri slices)"));
      size_t off = 0; for (int l = 0; l < n_f; l++) { f[l] = s->arena_f + off * 8; off += n[l] / G; } }
    for (int l = 0; l <= L; l++) { n_loc[l] = n[l] / G; first[l] = (uint64_t)(first_index / (n0 / n[l])); }     // slice of layer l: [first_l, first_l + n_l / G)
    { cudaError_t e = cudaMemcpyAsync(f[0], f0_slice, n_local * FR, mem == SMW_DEVICE ? cudaMemcpyDeviceToDevice : cudaMemcpyHostToDevice, c->stream);
      if (e != cudaSuccess) return fail(cuda_err(c, e, "copy f0 slice")); }
    uint32_t* d_z = nullptr; { cudaError_t e = cudaMalloc(&d_z, (size_t)(L + 1) * FR); if (e != cudaSuccess) return fail(cuda_err(c, e, "cudaMalloc")); defer.push_back(d_z); }
    std::vector<uint64_t> levels(L), sizes(L);
    for (int l = 0; l < L; l++) { levels[l] = (uint64_t)l; sizes[l] = (uint64_t)n[l]; }     // GLOBAL sizes: z_l is the same on every rank
    rc = sample_z_device(c, seed_z, levels, sizes, d_z); if (rc) return fail(rc);
    for (int l = 0; l + 1 < n_f; l++) { rc = launch_fri_fold(c, f[l], n_loc[l], d_z + (size_t)l * 8, m[l], f[l + 1]); if (rc) return fail(rc); }
    // whole sibling groups only: the ragged top of every layer tree belongs to the finisher (smw_merkle_root_from_level)
    FriLayers fl;
    rc = fri_commit_layers(c, K, L, n, n_loc, m, first, f, 0, true, fl, defer);
    s->arena_leaves = fl.arena_leaves; s->arena_levels = fl.arena_levels;
    if (rc) return fail(rc);
    for (int l = 0; l < K; l++) {
        const ForestTree& t = fl.trees[l];
        s->cap.push_back(t.lv.back()); s->cap_n.push_back(t.len.back()); s->cap_level.push_back((int)t.len.size() - 1); s->arity.push_back(t.arity);
    }
    if (K <= L) { s->tail = f[K]; s->tail_n = n_loc[K]; }
    cudaStreamSynchronize(c->stream);
    for (void* q : defer) cudaFree(q);
    *out = s.release();
    return SMW_OK;
}
int32_t smw_fri_shard_num_local_layers(const smw_fri_shard* s) { return s ? s->K : -1; }
int32_t smw_fri_shard_cap_info(const smw_fri_shard* s, int32_t l, size_t* n_nodes, int32_t* cap_level, int32_t* arity) {
    if (!s || l < 0 || l >= s->K) return SMW_ERR_INVALID_ARG;
    if (n_nodes) *n_nodes = s->cap_n[l]; if (cap_level) *cap_level = s->cap_level[l]; if (arity) *arity = s->arity[l];
    return SMW_OK;
}
int32_t smw_fri_shard_cap(smw_fri_shard* s, int32_t l, void* out, int32_t mem) {
    if (!s || !out || l < 0 || l >= s->K) return SMW_ERR_INVALID_ARG;
    smw_ctx* c = s->ctx; cudaSetDevice(c->device);
    SMW_CUDA(c, cudaMemcpyAsync(out, s->cap[l], s->cap_n[l] * FR, mem == SMW_DEVICE ? cudaMemcpyDeviceToDevice : cudaMemcpyDeviceToHost, c->stream));
    SMW_CUDA(c, cudaStreamSynchronize(c->stream)); return SMW_OK;
}
size_t smw_fri_shard_tail_len(const smw_fri_shard* s) { return s ? s->tail_n : 0; }
int32_t smw_fri_shard_tail(smw_fri_shard* s, void* out, int32_t mem) {
    if (!s || (s->tail_n && !out)) return SMW_ERR_INVALID_ARG;
    if (s->tail_n == 0) return SMW_OK;
    smw_ctx* c = s->ctx; cudaSetDevice(c->device);
    SMW_CUDA(c, cudaMemcpyAsync(out, s->tail, s->tail_n * FR, mem == SMW_DEVICE ? cudaMemcpyDeviceToDevice : cudaMemcpyDeviceToHost, c->stream));
    SMW_CUDA(c, cudaStreamSynchronize(c->stream)); return SMW_OK;
}
// ---- the range-sharded commit phase with the exchange inside the library: ONE all-gather per proof -------------------------
int32_t smw_fri_roots_sharded(smw_ctx* c, smw_comm* m, const void* chunks, size_t n_chunk, int32_t C, const uint64_t* first_indices,
                              size_t n0, const uint64_t* schedule, int32_t L, uint64_t seed_z, int32_t mem, void* roots_out_host) {
    if (!c || !chunks || !first_indices || !roots_out_host || C < 1 || L < 0 || (L && !schedule) || n0 == 0 || n_chunk == 0) return SMW_ERR_INVALID_ARG;
    const int world = m ? m->world : 1;
    const size_t V = (size_t)world * (size_t)C;                                   // slices in all
    if (n_chunk * V != n0) return set_err(c, SMW_ERR_INVALID_ARG, "world * n_chunks equal slices must cover n0");
    for (int q = 0; q < C; q++) if (first_indices[q] % n_chunk || first_indices[q] >= n0) return set_err(c, SMW_ERR_INVALID_ARG, "slice start must be a multiple of the slice length");
    Range rg("fri_roots_sharded");
    cudaSetDevice(c->device);
    std::vector<size_t> n, mm; int rc = fri_layer_sizes(c, n0, schedule, L, n, mm); if (rc) return rc;
    int K = 0;                                                                    // layers [0, K) are committed from slices (as smw_fri_shard_build)
    while (K <= L && n[K] % V == 0 && (K == L || (n[K] / V) % mm[K] == 0)) K++;
    const int n_f = (K <= L) ? K + 1 : K;
    std::vector<size_t> n_loc(L + 1, 0); for (int l = 0; l <= L; l++) n_loc[l] = n[l] / V;
    DevFree fr{{}, c->stream, c};
    std::vector<void*> defer; struct DeferFree { std::vector<void*>& v; cudaStream_t st; ~DeferFree() { cudaStreamSynchronize(st); for (void* q : v) cudaFree(q); } } dfree{defer, c->stream};
    // local slices: layer l holds the C slices back to back, so ONE fold launch per layer covers them all (a slice of a local
    // layer is a whole number of fold blocks)
    std::vector<uint32_t*> f(L + 2, nullptr);
    { size_t total = 0; for (int l = 0; l < n_f; l++) total += (size_t)C * n_loc[l];
      uint32_t* arena; SMW_CUDA(c, pool_malloc(c, (void**)&arena, total * FR)); fr.v.push_back(arena);
      size_t off = 0; for (int l = 0; l < n_f; l++) { f[l] = arena + off * 8; off += (size_t)C * n_loc[l]; } }
    SMW_CUDA(c, cudaMemcpyAsync(f[0], chunks, (size_t)C * n_chunk * FR, mem == SMW_DEVICE ? cudaMemcpyDeviceToDevice : cudaMemcpyHostToDevice, c->stream));
    uint32_t* d_z = nullptr; SMW_CUDA(c, cudaMalloc(&d_z, (size_t)(L + 1) * FR)); fr.v.push_back(d_z);
    std::vector<uint64_t> levels(L), sizes(L);
    for (int l = 0; l < L; l++) { levels[l] = (uint64_t)l; sizes[l] = (uint64_t)n[l]; }     // GLOBAL sizes: z_l is the same on every rank
    rc = sample_z_device(c, seed_z, levels, sizes, d_z); if (rc) return rc;
    for (int l = 0; l + 1 < n_f; l++) { rc = launch_fri_fold(c, f[l], (size_t)C * n_loc[l], d_z + (size_t)l * 8, mm[l], f[l + 1]); if (rc) return rc; }
    std::vector<LayerEntry> E;
    for (int l = 0; l < K; l++) for (int q = 0; q < C; q++)
        E.push_back(LayerEntry{l, n[l], n_loc[l], mm[l], first_indices[q] / (uint64_t)(n0 / n[l]), f[l] + (size_t)q * n_loc[l] * 8,
                               l < L ? f[l + 1] + (size_t)q * n_loc[l + 1] * 8 : nullptr});
    FriLayers fl;
    rc = fri_commit_entries(c, E, 0, true, fl, defer);
    fr.v.push_back(fl.arena_leaves); fr.v.push_back(fl.arena_levels);
    if (rc) return rc;
    // ---- one exchange: [slice starts | caps of every local layer | slices of f_K]
    std::vector<size_t> cap_n(K), off_l(K); std::vector<int> cap_level(K), arity(K);
    size_t send_n = (size_t)C;
    for (int l = 0; l < K; l++) { const ForestTree& t = fl.trees[(size_t)l * C]; cap_n[l] = t.len.back(); cap_level[l] = (int)t.len.size() - 1; arity[l] = t.arity; off_l[l] = send_n; send_n += (size_t)C * cap_n[l]; }
    const size_t off_tail = send_n, tail_n = (K <= L) ? n_loc[K] : 0;
    send_n += (size_t)C * tail_n;
    uint32_t *d_send, *d_recv;
    SMW_CUDA(c, pool_malloc(c, (void**)&d_send, send_n * FR)); fr.v.push_back(d_send);
    SMW_CUDA(c, pool_malloc(c, (void**)&d_recv, send_n * (size_t)world * FR)); fr.v.push_back(d_recv);
    { std::vector<uint64_t> hdr((size_t)C * 4, 0); for (int q = 0; q < C; q++) hdr[(size_t)q * 4] = first_indices[q];
      SMW_CUDA(c, cudaMemcpyAsync(d_send, hdr.data(), (size_t)C * FR, cudaMemcpyHostToDevice, c->stream)); }
    for (int l = 0; l < K; l++) for (int q = 0; q < C; q++)
        SMW_CUDA(c, cudaMemcpyAsync(d_send + (off_l[l] + (size_t)q * cap_n[l]) * 8, fl.trees[(size_t)l * C + q].lv.back(), cap_n[l] * FR, cudaMemcpyDeviceToDevice, c->stream));
    if (tail_n) SMW_CUDA(c, cudaMemcpyAsync(d_send + off_tail * 8, f[K], (size_t)C * tail_n * FR, cudaMemcpyDeviceToDevice, c->stream));
    rc = comm_all_gather_bytes(c, m, d_send, d_recv, send_n * FR); if (rc) return rc;
    // ---- global slice order from the gathered starts
    std::vector<uint64_t> hdr((size_t)world * C * 4);
    SMW_CUDA(c, cudaMemcpy2DAsync(hdr.data(), (size_t)C * FR, d_recv, send_n * FR, (size_t)C * FR, (size_t)world, cudaMemcpyDeviceToHost, c->stream));
    SMW_CUDA(c, cudaStreamSynchronize(c->stream));
    std::vector<std::pair<uint64_t, size_t>> order(V);                            // (start, rank * C + chunk)
    for (size_t v = 0; v < V; v++) order[v] = {hdr[v * 4], v};
    std::sort(order.begin(), order.end());
    for (size_t v = 0; v < V; v++) if (order[v].first != (uint64_t)v * n_chunk) return set_err(c, SMW_ERR_INVALID_ARG, "the slices of all ranks do not tile [0, n0)");
    size_t tops_n = 0; std::vector<size_t> top_off(K);
    for (int l = 0; l < K; l++) { top_off[l] = tops_n; tops_n += V * cap_n[l]; }
    const size_t tail_off = tops_n; tops_n += V * tail_n;
    std::vector<uint64_t> idx(tops_n);
    for (int l = 0; l < K; l++) for (size_t v = 0; v < V; v++) {
        const size_t r = order[v].second / C, q = order[v].second % C;
        for (size_t e = 0; e < cap_n[l]; e++) idx[top_off[l] + v * cap_n[l] + e] = r * send_n + off_l[l] + q * cap_n[l] + e;
    }
    for (size_t v = 0; v < V; v++) { const size_t r = order[v].second / C, q = order[v].second % C; for (size_t e = 0; e < tail_n; e++) idx[tail_off + v * tail_n + e] = r * send_n + off_tail + q * tail_n + e; }
    uint64_t* d_idx = nullptr; uint32_t* d_tops = nullptr;
    if (tops_n) {
        SMW_CUDA(c, cudaMalloc(&d_idx, tops_n * 8)); fr.v.push_back(d_idx);
        SMW_CUDA(c, pool_malloc(c, (void**)&d_tops, tops_n * FR)); fr.v.push_back(d_tops);
        SMW_CUDA(c, cudaMemcpyAsync(d_idx, idx.data(), tops_n * 8, cudaMemcpyHostToDevice, c->stream));
        rc = launch_gather(c, d_recv, d_idx, tops_n, d_tops); if (rc) return rc;
    }
    // ---- every rank finishes the tops of the local layers (a small forest) ...
    std::vector<ForestTree> tops;
    for (int l = 0; l < K; l++) {
        smw_params* p; rc = smw_params_for_arity(c, arity[l], &p); if (rc) return rc;
        tops.push_back(ForestTree{p, arity[l], (uint64_t)l, 0, (uint32_t)cap_level[l], false, d_tops + top_off[l] * 8, V * cap_n[l], {}, {}});
    }
    uint32_t* top_arena = nullptr;
    rc = build_forest(c, tops, &top_arena, defer); fr.v.push_back(top_arena); if (rc) return rc;
    uint32_t* d_roots; SMW_CUDA(c, cudaMalloc(&d_roots, (size_t)(L + 1) * FR)); fr.v.push_back(d_roots);
    for (int l = 0; l < K; l++) SMW_CUDA(c, cudaMemcpyAsync(d_roots + (size_t)l * 8, tops[l].lv.back(), FR, cudaMemcpyDeviceToDevice, c->stream));
    if (K) { rc = d2h(c, roots_out_host, d_roots, (size_t)K * FR); if (rc) return rc; }
    // ---- ... and rebuilds the layers too small to split from the gathered f_K
    if (K <= L) {
        smw_fri* sub = nullptr;
        rc = smw_fri_build_transcript_from(c, d_tops + tail_off * 8, n[K], schedule + K, L - K, seed_z, (uint64_t)K, SMW_DEVICE, &sub); if (rc) return rc;
        rc = smw_fri_roots(sub, (uint8_t*)roots_out_host + (size_t)K * FR);
        smw_fri_free(sub);
    }
    return rc;
}
int32_t smw_ntt_cyclic_to_chunks(smw_ctx* c, smw_comm* m, void* in_dev, void* out_dev, int32_t log_n) {
    if (!c || !in_dev || !out_dev || log_n < 0 || log_n > 32) return SMW_ERR_INVALID_ARG;
    const int G = m ? m->world : 1, rank = m ? m->rank : 0;
    int log_g = 0; while ((1 << log_g) < G) log_g++;
    if ((1 << log_g) != G || log_g > 4 || log_n < 2 * log_g) return set_err(c, SMW_ERR_INVALID_ARG, "ranks must be a power of two <= 16 and divide the local length");
    Range rg("ntt_cyclic_to_chunks");
    cudaSetDevice(c->device);
    const size_t M = ((size_t)1 << log_n) >> log_g;
    int rc = launch_ntt(c, (uint32_t*)in_dev, log_n - log_g, 0, 1, M); if (rc) return rc;       // NTT_M of the cyclic slice
    if (G == 1) { SMW_CUDA(c, cudaMemcpyAsync(out_dev, in_dev, M * FR, cudaMemcpyDeviceToDevice, c->stream)); return SMW_OK; }
    DevFree fr{{}, c->stream};
    uint32_t *d_w, *d_wr; SMW_CUDA(c, cudaMalloc(&d_w, FR)); fr.v.push_back(d_w); SMW_CUDA(c, cudaMalloc(&d_wr, FR)); fr.v.push_back(d_wr);
    rc = launch_root_of_unity(c, log_n, d_w); if (rc) return rc;
    rc = launch_fr_pow(c, d_w, (uint64_t)rank, d_wr); if (rc) return rc;                       // w_N^rank
    rc = launch_scale_powers(c, (uint32_t*)in_dev, M, d_wr); if (rc) return rc;                // twiddle: y[k2] *= w_N^(rank k2)
    rc = comm_all_to_all_bytes(c, m, in_dev, out_dev, (M / G) * FR); if (rc) return rc;        // rank r receives the k2-range [r M / G, (r + 1) M / G) of every rank
    uint32_t* d_wpow = nullptr; rc = make_wpow(c, log_g, &d_wpow, fr); if (rc) return rc;
    return launch_dft_strided(c, (uint32_t*)out_dev, log_g, M / G, M / G, d_wpow);             // length-G DFT across the received pieces
}
uint64_t smw_ctx_collective_count(const smw_ctx* c) { return c ? c->collectives : 0; }
int32_t smw_diag_tc_node_counts(const smw_ctx* c, const smw_params* p, uint64_t* out6) {
    if (!c || !p || !out6) return SMW_ERR_INVALID_ARG;
    for (int i = 0; i < 6; i++) out6[i] = 0;
    out6[5] = c->tc_min_nodes;
    if (c->use_f52 && smw::tc::usable(p)) smw::tc::node_counts(p->t, p->rp, out6);
    return SMW_OK;
}
int32_t smw_ctx_set_tc_min_nodes(smw_ctx* c, uint64_t n) { if (!c) return SMW_ERR_INVALID_ARG; c->tc_min_nodes = n ? n : ~0ull; return SMW_OK; }

// root of the tree whose DS-level-`level` children are `nodes` (level 0: the leaves / leaf digests themselves)
int32_t smw_merkle_root_from_level(smw_ctx* c, smw_params* p, int32_t arity, uint64_t label, int32_t level, const void* nodes, size_t n, int32_t mem, void* root_out_host) {
    if (!root_out_host || level < 0) return SMW_ERR_INVALID_ARG;
    int rc = check_tree_args(c, p, arity, n); if (rc) return rc;
    if (!nodes) return SMW_ERR_INVALID_ARG;
    cudaSetDevice(c->device);
    uint32_t* d0; SMW_CUDA(c, cudaMalloc(&d0, n * FR));
    { cudaError_t e = cudaMemcpyAsync(d0, nodes, n * FR, mem == SMW_DEVICE ? cudaMemcpyDeviceToDevice : cudaMemcpyHostToDevice, c->stream);
      if (e != cudaSuccess) { cudaFree(d0); return cuda_err(c, e, "copy nodes"); } }
    smw_tree* t = nullptr;
    rc = build_tree(c, p, arity, label, d0, true, n, &t, (uint32_t)level); if (rc) return rc;
    rc = smw_tree_root(t, root_out_host);
    smw_tree_free(t);
    return rc;
}

// ---------------------------------------------------------------------------------------- MF sum-check commit loop
int32_t smw_sumcheck_round_coeffs(smw_ctx* c, const void* layer, size_t n, void* c0_host, void* c1_host, int32_t mem) {
    if (!c || !c0_host || !c1_host || (n && !layer) || (n & 1)) return SMW_ERR_INVALID_ARG;
    cudaSetDevice(c->device);
    In a(c, layer, n, mem); if (a.rc) return a.rc;
    uint32_t* d_c; SMW_CUDA(c, cudaMalloc(&d_c, 2 * FR));
    int rc = launch_mf_coeffs(c, a.d, n, d_c);
    uint8_t tmp[64];
    if (!rc) rc = d2h(c, tmp, d_c, 64);
    if (!rc) { memcpy(c0_host, tmp, 32); memcpy(c1_host, tmp + 32, 32); }
    cudaFree(d_c);
    return rc;
}
// channel/src/lib.rs:592-598: new("SUMCHECK-MF/ROUND-CHAL"); absorb_bytes("SUMCHECK/MF/R"); absorb_bytes(u64le(i)); absorb_field(root); challenge("r_i")
static void mf_challenge_words(uint64_t round_idx, std::vector<Word>& A, std::vector<Word>& C) {
    A.clear(); C.clear();
    A.push_back(tag_word(TAG_INIT));
    push_absorb_bytes(A, (const uint8_t*)"SUMCHECK-MF/ROUND-CHAL", 22);
    push_absorb_bytes(A, (const uint8_t*)"SUMCHECK/MF/R", 13);
    uint8_t le[8]; for (int i = 0; i < 8; i++) le[i] = (uint8_t)(round_idx >> (8 * i));
    push_absorb_bytes(A, le, 8);
    push_challenge(C, (const uint8_t*)"r_i", 3);
}
static int mf_challenge_device(smw_ctx* c, smw_params* trp, uint64_t round_idx, const uint32_t* d_root, uint32_t* d_out) {
    std::vector<Word> A, C; mf_challenge_words(round_idx, A, C);
    uint32_t *dA, *dC; int rc = upload_words(c, A, &dA); if (rc) return rc; rc = upload_words(c, C, &dC); if (rc) { cudaFree(dA); return rc; }
    rc = launch_transcript_run(c, trp, nullptr, 1, 0, dA, (int)A.size(), d_root, 1, dC, (int)C.size(), 1, d_out);
    cudaStreamSynchronize(c->stream); cudaFree(dA); cudaFree(dC);
    return rc;
}
int32_t smw_mf_round_challenge_from_root(smw_ctx* c, uint64_t round_idx, const void* prev_root_host, smw_params* trp, void* out_host) {
    if (!c || !prev_root_host || !out_host) return SMW_ERR_INVALID_ARG;
    cudaSetDevice(c->device);
    smw_params* p; int rc = get_transcript_params(c, trp, &p); if (rc) return rc;
    In r(c, prev_root_host, 1, SMW_HOST); if (r.rc) return r.rc;
    Out o(c, out_host, 1, SMW_HOST); if (o.rc) return o.rc;
    rc = mf_challenge_device(c, p, round_idx, r.d, o.d); if (rc) return rc;
    return o.finish();
}
int32_t smw_mf_fold(smw_ctx* c, const void* cur, size_t n, const void* r_host, void* next, int32_t mem) {
    if (!c || !r_host || (n & 1) || (n && (!cur || !next))) return SMW_ERR_INVALID_ARG;
    cudaSetDevice(c->device);
    In a(c, cur, n, mem), r(c, r_host, 1, SMW_HOST); if (a.rc) return a.rc; if (r.rc) return r.rc;
    Out o(c, next, n / 2, mem); if (o.rc) return o.rc;
    int rc = launch_mf_fold(c, a.d, n, r.d, o.d); if (rc) return rc;
    return o.finish();
}
} // extern "C"
namespace {
// Device-resident pieces of the Merkle-folded sum-check commit loop (channel/src/lib.rs:600-657): roots, challenges, ping-pong
// layer and level buffers, the challenge pre-images of every round.  Nothing is allocated inside the loop.
struct MfBufs {
    uint32_t *d_roots = nullptr, *d_r = nullptr, *bufA = nullptr, *bufB = nullptr, *lvA = nullptr, *lvB = nullptr, *dA = nullptr, *dC = nullptr;
    int nA = 0, nC = 0;
};
int mf_alloc(smw_ctx* c, int k, size_t n_layer0, MfBufs& B, DevFree& fr) {
    SMW_CUDA(c, cudaMalloc(&B.d_roots, (size_t)(k + 1) * FR)); fr.v.push_back(B.d_roots);
    SMW_CUDA(c, cudaMalloc(&B.d_r, (size_t)(k + 1) * FR)); fr.v.push_back(B.d_r);
    SMW_CUDA(c, cudaMalloc(&B.bufA, (n_layer0 / 2 + 1) * FR)); fr.v.push_back(B.bufA);
    SMW_CUDA(c, cudaMalloc(&B.bufB, (n_layer0 / 4 + 1) * FR)); fr.v.push_back(B.bufB);
    SMW_CUDA(c, cudaMalloc(&B.lvA, ((n_layer0 + 15) / 16 + 1) * FR)); fr.v.push_back(B.lvA);
    SMW_CUDA(c, cudaMalloc(&B.lvB, ((n_layer0 + 255) / 256 + 1) * FR)); fr.v.push_back(B.lvB);
    std::vector<Word> allA, A, C;
    for (int i = 0; i < k; i++) { mf_challenge_words((uint64_t)i, A, C); B.nA = (int)A.size(); B.nC = (int)C.size(); allA.insert(allA.end(), A.begin(), A.end()); }
    if (k > 0) { int rc = upload_words(c, allA, &B.dA); fr.v.push_back(B.dA); if (rc) return rc; rc = upload_words(c, C, &B.dC); fr.v.push_back(B.dC); if (rc) return rc; }
    return SMW_OK;
}
// root of the arity-16 tree (tree_label = ds_tag) whose DS-level-`level` children are `nodes` -> d_root; uses the two level buffers
int mf_tree_top(smw_ctx* c, smw_params* cp, uint64_t ds_tag, const uint32_t* nodes, size_t cn, uint32_t level, uint64_t first, uint32_t* lvA, uint32_t* lvB, uint32_t* d_root) {
    if (cn == 1) { SMW_CUDA(c, cudaMemcpyAsync(d_root, nodes, FR, cudaMemcpyDeviceToDevice, c->stream)); return SMW_OK; }      // a one-leaf tree's root is the leaf
    const uint32_t* lv = nodes; uint32_t* dst = (nodes == lvA) ? lvB : lvA;
    while (cn > 1) {
        const size_t nn = (cn + 15) / 16; first /= 16;
        uint32_t* o = (nn == 1) ? d_root : dst;
        int rc = launch_merkle_level(c, cp, lv, cn, o, nn, 16, level, ds_tag, 0, first, 1, 0, 0); if (rc) return rc;
        lv = o; cn = nn; level++; dst = (dst == lvA) ? lvB : lvA;
    }
    return SMW_OK;
}
// rounds i0 .. k on one rank holding the whole layer `cur` (len elements); returns the final layer pointer (one element)
int mf_rounds(smw_ctx* c, smw_params* cp, smw_params* trp, uint64_t ds_tag, const uint32_t* cur, size_t len, int i0, int k, MfBufs& B, const uint32_t** final_out) {
    int rc = SMW_OK;
    for (int i = i0; i <= k && !rc; i++) {
        uint32_t* d_root = B.d_roots + (size_t)i * 8;
        rc = mf_tree_top(c, cp, ds_tag, cur, len, 0, 0, B.lvA, B.lvB, d_root);
        if (rc || i == k) break;
        // r_i from (i, root_i) stays on the device and feeds the fold directly: no host round trip per round
        rc = launch_transcript_run(c, trp, nullptr, 1, 0, B.dA + (size_t)i * B.nA * 8, B.nA, d_root, 1, B.dC, B.nC, 1, B.d_r + (size_t)i * 8); if (rc) break;
        uint32_t* nxt = (cur == B.bufA) ? B.bufB : B.bufA;
        rc = launch_mf_fold(c, cur, len, B.d_r + (size_t)i * 8, nxt); if (rc) break;
        cur = nxt; len /= 2;
    }
    *final_out = cur;
    return rc;
}
int mf_outputs(smw_ctx* c, int k, const MfBufs& B, const uint32_t* fin, void* roots_out_host, void* r_out_host, void* final_out_host) {
    int rc = d2h(c, roots_out_host, B.d_roots, (size_t)(k + 1) * FR);
    if (!rc && r_out_host && k > 0) rc = d2h(c, r_out_host, B.d_r, (size_t)k * FR);
    if (!rc && final_out_host) rc = d2h(c, final_out_host, fin, FR);
    return rc;
}
} // namespace
extern "C" {
// The whole commit loop of SumCheckMFProver (channel/src/lib.rs:600-657) as ONE chain of launches with no host round trip and
// no allocation inside: commit layer i (arity 16, tree_label = ds_tag; only the root is kept), r_i from (i, root_i) on the
// device, fold, repeat.  The rounds are strictly sequential (r_i needs root_i).
int32_t smw_mf_commit_phase(smw_ctx* c, uint64_t ds_tag, const void* table, size_t n, int32_t mem, void* roots_out_host, void* r_out_host, void* final_out_host) {
    if (!c || !table || !roots_out_host || n == 0 || (n & (n - 1))) return SMW_ERR_INVALID_ARG;
    Range rg("mf_commit_phase");
    cudaSetDevice(c->device);
    int k = 0; while (((size_t)1 << k) < n) k++;
    smw_params *cp, *trp; int rc = smw_params_commitment_default(c, &cp); if (rc) return rc; rc = smw_params_transcript_default(c, &trp); if (rc) return rc;
    In t0(c, table, n, mem); if (t0.rc) return t0.rc;
    DevFree fr{{}, c->stream}; MfBufs B;
    rc = mf_alloc(c, k, n, B, fr); if (rc) return rc;
    const uint32_t* fin = nullptr;
    rc = mf_rounds(c, cp, trp, ds_tag, t0.d, n, 0, k, B, &fin); if (rc) return rc;
    return mf_outputs(c, k, B, fin, roots_out_host, r_out_host, final_out_host);
}
// The same loop for a table split into `world` contiguous slices: while a slice holds at least one fold pair, the
// rank commits its slice up to the cap (whole sibling groups, DS positions offset by the slice start), ONE all-gather carries
// the caps, every rank finishes the top (so the root, and r_i, exist on every rank without a broadcast) and folds its own slice.
int32_t smw_mf_commit_phase_sharded(smw_ctx* c, smw_comm* m, uint64_t ds_tag, const void* table_slice, size_t n_local, int32_t mem,
                                    void* roots_out_host, void* r_out_host, void* final_out_host) {
    if (!c || !table_slice || !roots_out_host || n_local == 0 || (n_local & (n_local - 1))) return SMW_ERR_INVALID_ARG;
    const int world = m ? m->world : 1, rank = m ? m->rank : 0;
    if (world & (world - 1)) return set_err(c, SMW_ERR_INVALID_ARG, "the number of ranks must be a power of two");
    Range rg("mf_commit_phase_sharded");
    cudaSetDevice(c->device);
    const size_t n = n_local * (size_t)world;
    int k = 0; while (((size_t)1 << k) < n) k++;
    smw_params *cp, *trp; int rc = smw_params_commitment_default(c, &cp); if (rc) return rc; rc = smw_params_transcript_default(c, &trp); if (rc) return rc;
    In t0(c, table_slice, n_local, mem); if (t0.rc) return t0.rc;
    DevFree fr{{}, c->stream}; MfBufs B;
    rc = mf_alloc(c, k, std::max<size_t>(n_local, (size_t)world * 16), B, fr); if (rc) return rc;
    uint32_t *d_send = nullptr, *d_recv = nullptr;
    SMW_CUDA(c, cudaMalloc(&d_send, 16 * FR)); fr.v.push_back(d_send);
    SMW_CUDA(c, cudaMalloc(&d_recv, (size_t)world * 16 * FR)); fr.v.push_back(d_recv);
    const uint32_t* cur = t0.d; size_t len_loc = n_local; int i = 0;
    while (world > 1 && len_loc >= 2) {
        // local subtree: whole sibling groups only
        const uint32_t* lv = cur; size_t cn = len_loc; uint32_t level = 0; uint64_t first = (uint64_t)rank * len_loc; uint32_t* dst = B.lvA;
        while (cn >= 16 && cn % 16 == 0) {
            const size_t nn = cn / 16; first /= 16;
            rc = launch_merkle_level(c, cp, lv, cn, dst, nn, 16, level, ds_tag, 0, first, 1, 0, 0); if (rc) return rc;
            lv = dst; cn = nn; level++; dst = (dst == B.lvA) ? B.lvB : B.lvA;
        }
        // cap: cn <= 15 nodes (or the < 16 leaves themselves) per rank, in rank order = global order
        rc = comm_all_gather_bytes(c, m, lv, d_recv, cn * FR); if (rc) return rc;
        uint32_t* d_root = B.d_roots + (size_t)i * 8;
        rc = mf_tree_top(c, cp, ds_tag, d_recv, cn * (size_t)world, level, 0, B.lvA, B.lvB, d_root); if (rc) return rc;
        rc = launch_transcript_run(c, trp, nullptr, 1, 0, B.dA + (size_t)i * B.nA * 8, B.nA, d_root, 1, B.dC, B.nC, 1, B.d_r + (size_t)i * 8); if (rc) return rc;
        uint32_t* nxt = (cur == B.bufA) ? B.bufB : B.bufA;
        rc = launch_mf_fold(c, cur, len_loc, B.d_r + (size_t)i * 8, nxt); if (rc) return rc;
        cur = nxt; len_loc /= 2; i++;
    }
    size_t len = len_loc;
    if (world > 1) {        // one element per rank left: gather the layer, the remaining rounds run replicated
        SMW_CUDA(c, cudaMemcpyAsync(d_send, cur, FR, cudaMemcpyDeviceToDevice, c->stream));
        uint32_t* full = (cur == B.bufA) ? B.bufB : B.bufA;
        rc = comm_all_gather_bytes(c, m, d_send, full, FR); if (rc) return rc;
        cur = full; len = (size_t)world;
    }
    const uint32_t* fin = nullptr;
    rc = mf_rounds(c, cp, trp, ds_tag, cur, len, i, k, B, &fin); if (rc) return rc;
    return mf_outputs(c, k, B, fin, roots_out_host, r_out_host, final_out_host);
}

// ---------------------------------------------------------------------------------------- openings (merkle/src/lib.rs:246-315)
struct smw_proof {
    std::vector<uint64_t> indices;                       // unique, ascending
    std::vector<std::vector<uint8_t>> siblings;          // per level, 32 bytes per sibling
    std::vector<std::vector<uint8_t>> group_sizes;       // per level, child count of every touched parent
    int arity = 0;
};
int32_t smw_tree_open_many(smw_tree* t, const uint64_t* indices, size_t n, smw_proof** out) {
    if (!t || !out) return SMW_ERR_INVALID_ARG;
    smw_ctx* c = t->ctx; cudaSetDevice(c->device);
    if (n == 0 || !indices) return set_err(c, SMW_ERR_EMPTY_INDICES, "open_many: empty indices");
    const size_t arity = (size_t)t->arity; const int height = (int)t->lv.size() - 1;
    std::vector<uint64_t> cur(indices, indices + n);
    std::sort(cur.begin(), cur.end()); cur.erase(std::unique(cur.begin(), cur.end()), cur.end());
    if (cur.back() >= t->len[0]) return set_err(c, SMW_ERR_INVALID_ARG, "open_many: index out of range");
    auto pr = std::make_unique<smw_proof>(); pr->arity = t->arity; pr->indices = cur;
    pr->siblings.resize(height); pr->group_sizes.resize(height);
    for (int level = 0; level < height; level++) {
        const size_t level_len = t->len[level];
        std::vector<uint64_t> need;                      // node indices at this level to fetch, in proof order
        size_t i = 0;
        while (i < cur.size()) {                         // cur is sorted: equal-parent runs are the BTreeMap groups
            size_t parent = cur[i] / arity, j = i;
            while (j < cur.size() && cur[j] / arity == parent) j++;
            size_t base = parent * arity, end = std::min(base + arity, level_len), cc = end - base;
            pr->group_sizes[level].push_back((uint8_t)cc);
            size_t k = i;
            for (size_t pos = 0; pos < cc; pos++) { if (k < j && cur[k] % arity == pos) k++; else need.push_back(base + pos); }
            i = j;
        }
        pr->siblings[level].resize(need.size() * FR);
        if (!need.empty()) {
            uint64_t* d_idx; uint32_t* d_out;
            SMW_CUDA(c, cudaMalloc(&d_idx, need.size() * 8)); SMW_CUDA(c, cudaMalloc(&d_out, need.size() * FR));
            SMW_CUDA(c, cudaMemcpyAsync(d_idx, need.data(), need.size() * 8, cudaMemcpyHostToDevice, c->stream));
            int rc = launch_gather(c, t->lv[level], d_idx, need.size(), d_out);
            if (!rc) rc = d2h(c, pr->siblings[level].data(), d_out, need.size() * FR);
            cudaStreamSynchronize(c->stream); cudaFree(d_idx); cudaFree(d_out);
            if (rc) return rc;
        }
        for (auto& v : cur) v /= arity;
        cur.erase(std::unique(cur.begin(), cur.end()), cur.end());
    }
    *out = pr.release();
    return SMW_OK;
}
int32_t smw_proof_num_levels(const smw_proof* p) { return p ? (int32_t)p->siblings.size() : -1; }
size_t smw_proof_num_indices(const smw_proof* p) { return p ? p->indices.size() : 0; }
int32_t smw_proof_indices(const smw_proof* p, uint64_t* out) { if (!p || !out) return SMW_ERR_INVALID_ARG; memcpy(out, p->indices.data(), p->indices.size() * 8); return SMW_OK; }
size_t smw_proof_level_siblings(const smw_proof* p, int32_t l) { return (p && l >= 0 && l < (int)p->siblings.size()) ? p->siblings[l].size() / FR : 0; }
size_t smw_proof_level_groups(const smw_proof* p, int32_t l) { return (p && l >= 0 && l < (int)p->group_sizes.size()) ? p->group_sizes[l].size() : 0; }
int32_t smw_proof_level(const smw_proof* p, int32_t l, void* sib_out, uint8_t* grp_out) {
    if (!p || l < 0 || l >= (int)p->siblings.size()) return SMW_ERR_INVALID_ARG;
    if (sib_out) memcpy(sib_out, p->siblings[l].data(), p->siblings[l].size());
    if (grp_out) memcpy(grp_out, p->group_sizes[l].data(), p->group_sizes[l].size());
    return SMW_OK;
}
void smw_proof_free(smw_proof* p) { delete p; }

// ---------------------------------------------------------------------------------------- verification (merkle/src/lib.rs:587-701, 725-773)
// The index work of a multiproof check does not depend on the digests, so the host lays out EVERY level first — which pool
// entry each child of each recomputed parent comes from (opened value, sibling, parent of the level below) — uploads that
// once, and the device hashes level after level out of one pool: [pairs | leaves | siblings | parents level 0 | level 1 | ...].
} // extern "C"
namespace {
struct VerifyPlan {
    std::vector<uint64_t> child_idx, pos; std::vector<uint8_t> cnt;     // all levels concatenated; child_idx is [group][arity]
    std::vector<size_t> first_group, n_groups;                          // per level
    size_t pool_entries = 0, final_entry = 0;
};
// false = the reference returns false on the structure alone
bool plan_verify(const std::vector<uint64_t>& req, size_t leaf_base, size_t sib_base, int arity, int n_levels, const size_t* n_sib, const size_t* n_grp,
                 const uint8_t* group_sizes, VerifyPlan& pl) {
    std::vector<uint64_t> cur_idx = req; std::vector<size_t> cur_entry(req.size());
    for (size_t i = 0; i < req.size(); i++) cur_entry[i] = leaf_base + i;
    size_t sib_off = sib_base, total_sib = 0, goff = 0;
    for (int l = 0; l < n_levels; l++) total_sib += n_sib[l];
    size_t next_entry = sib_base + total_sib;
    for (int l = 0; l < n_levels; l++) {
        size_t groups = 0;
        for (size_t a = 0; a < cur_idx.size();) { const uint64_t par = cur_idx[a] / arity; while (a < cur_idx.size() && cur_idx[a] / arity == par) a++; groups++; }
        if (groups != n_grp[l]) return false;
        pl.first_group.push_back(pl.cnt.size()); pl.n_groups.push_back(groups);
        std::vector<uint64_t> nidx; std::vector<size_t> nent; nidx.reserve(groups); nent.reserve(groups);
        size_t off = 0, g = 0;
        for (size_t i = 0; i < cur_idx.size(); g++) {
            const uint64_t parent = cur_idx[i] / arity; size_t j = i; while (j < cur_idx.size() && cur_idx[j] / arity == parent) j++;
            const size_t cc = group_sizes[goff + g];
            if (cc == 0 || cc > (size_t)arity) return false;
            size_t k = i; const size_t row = pl.child_idx.size(); pl.child_idx.resize(row + arity, 0);
            for (size_t c = 0; c < cc; c++) {
                if (k < j && cur_idx[k] % arity == c) pl.child_idx[row + c] = cur_entry[k++];
                else { if (off >= n_sib[l]) return false; pl.child_idx[row + c] = sib_off + off++; }
            }
            pl.cnt.push_back((uint8_t)cc); pl.pos.push_back(parent);
            nidx.push_back(parent); nent.push_back(next_entry++);
            i = j;
        }
        if (off != n_sib[l]) return false;
        sib_off += n_sib[l]; goff += n_grp[l];
        cur_idx.swap(nidx); cur_entry.swap(nent);
    }
    if (cur_idx.size() != 1) return false;
    pl.pool_entries = next_entry; pl.final_entry = cur_entry[0];
    return true;
}
template <class V> int upload_vec(smw_ctx* c, const std::vector<V>& h, V** d) {
    *d = nullptr; if (h.empty()) return SMW_OK;
    SMW_CUDA(c, cudaMalloc(d, h.size() * sizeof(V)));
    SMW_CUDA(c, cudaMemcpyAsync(*d, h.data(), h.size() * sizeof(V), cudaMemcpyHostToDevice, c->stream));
    return SMW_OK;
}

// pairs != nullptr: 2 * nreq elements (f, cp interleaved, aligned with req) whose leaf digests are computed first; else leaves (nreq)
// legacy_tag_host != nullptr: the legacy sponge (hash_with_ds(children, ds_tag), merkle/src/lib.rs:487-584, 704-721) instead of the DS one
int verify_core(smw_ctx* c, smw_params* p, int arity, uint64_t label, const void* root_host, const std::vector<uint64_t>& req,
                const std::vector<uint8_t>& leaves, const std::vector<uint8_t>& pairs, int n_levels, const size_t* n_sib, const void* siblings_host,
                const size_t* n_grp, const uint8_t* group_sizes, int32_t* ok_out, const void* legacy_tag_host = nullptr) {
    const size_t nreq = req.size(), n_pairs = pairs.size() / FR, leaf_base = n_pairs, sib_base = leaf_base + nreq;
    VerifyPlan pl;
    if (!plan_verify(req, leaf_base, sib_base, arity, n_levels, n_sib, n_grp, group_sizes, pl)) { *ok_out = 0; return SMW_OK; }
    size_t total_sib = 0; for (int l = 0; l < n_levels; l++) total_sib += n_sib[l];
    DevFree fr{{}, c->stream};
    uint32_t* d_tag = nullptr;
    if (legacy_tag_host) { SMW_CUDA(c, cudaMalloc(&d_tag, FR)); fr.v.push_back(d_tag); SMW_CUDA(c, cudaMemcpyAsync(d_tag, legacy_tag_host, FR, cudaMemcpyHostToDevice, c->stream)); }
    uint32_t* pool = nullptr; SMW_CUDA(c, cudaMalloc(&pool, pl.pool_entries * FR)); fr.v.push_back(pool);
    if (n_pairs) SMW_CUDA(c, cudaMemcpyAsync(pool, pairs.data(), pairs.size(), cudaMemcpyHostToDevice, c->stream));
    else SMW_CUDA(c, cudaMemcpyAsync(pool + leaf_base * 8, leaves.data(), leaves.size(), cudaMemcpyHostToDevice, c->stream));
    if (total_sib) SMW_CUDA(c, cudaMemcpyAsync(pool + sib_base * 8, siblings_host, total_sib * FR, cudaMemcpyHostToDevice, c->stream));
    int rc;
    if (n_pairs) {        // encode_leaf_digest_ds, merkle/src/lib.rs:382-390 / 752-763: H(DS(arity, u32::MAX, idx, label), [f, cp])
        std::vector<uint64_t> ci(nreq * 2); std::vector<uint8_t> two(nreq, 2);
        for (size_t i = 0; i < 2 * nreq; i++) ci[i] = i;
        uint64_t *d_ci, *d_pos; uint8_t* d_two;
        rc = upload_vec(c, ci, &d_ci); fr.v.push_back(d_ci); if (rc) return rc;
        rc = upload_vec(c, req, &d_pos); fr.v.push_back(d_pos); if (rc) return rc;
        rc = upload_vec(c, two, &d_two); fr.v.push_back(d_two); if (rc) return rc;
        if (d_tag) rc = launch_legacy_nodes(c, p, 2, pool, nullptr, 2, d_ci, d_two, nreq, d_tag, pool + leaf_base * 8);       // encode_leaf_digest_legacy, :373-375
        else rc = launch_hash_groups(c, p, pool, d_ci, d_two, d_pos, nreq, arity, 2, 0xffffffffu, label, pool + leaf_base * 8);
        if (rc) return rc;
    }
    uint64_t *d_child = nullptr, *d_pos = nullptr; uint8_t* d_cnt = nullptr;
    rc = upload_vec(c, pl.child_idx, &d_child); fr.v.push_back(d_child); if (rc) return rc;
    rc = upload_vec(c, pl.pos, &d_pos); fr.v.push_back(d_pos); if (rc) return rc;
    rc = upload_vec(c, pl.cnt, &d_cnt); fr.v.push_back(d_cnt); if (rc) return rc;
    size_t out_entry = sib_base + total_sib;
    for (int l = 0; l < n_levels; l++) {
        const size_t g0 = pl.first_group[l], ng = pl.n_groups[l];
        if (d_tag) rc = launch_legacy_nodes(c, p, 2, pool, nullptr, (size_t)arity, d_child + g0 * (size_t)arity, d_cnt + g0, ng, d_tag, pool + out_entry * 8);
        else rc = launch_hash_groups(c, p, pool, d_child + g0 * (size_t)arity, d_cnt + g0, d_pos + g0, ng, arity, arity, (uint32_t)l, label, pool + out_entry * 8);
        if (rc) return rc;
        out_entry += ng;
    }
    uint8_t got[FR]; rc = d2h(c, got, pool + pl.final_entry * 8, FR); if (rc) return rc;
    *ok_out = memcmp(got, root_host, FR) == 0 ? 1 : 0;
    return SMW_OK;
}
// sort + dedup the requested indices and align the per-index payload (`per` elements each) with them: BTreeMap insert, the last duplicate wins
bool align_request(const uint64_t* indices, const void* payload, size_t n, size_t per, const uint64_t* proof_indices, size_t n_pidx,
                   std::vector<uint64_t>& req, std::vector<uint8_t>& aligned) {
    std::map<uint64_t, size_t> last; for (size_t i = 0; i < n; i++) last[indices[i]] = i;
    req.clear(); for (auto& kv : last) req.push_back(kv.first);
    aligned.resize(req.size() * per * FR);
    size_t k = 0; for (auto& kv : last) { memcpy(aligned.data() + k * per * FR, (const uint8_t*)payload + kv.second * per * FR, per * FR); k++; }
    if (n_pidx != req.size() || !proof_indices) return false;                    // proof.indices != req
    for (size_t i = 0; i < n_pidx; i++) if (proof_indices[i] != req[i]) return false;
    return true;
}
} // namespace
extern "C" {

int32_t smw_verify_many_ds(smw_ctx* c, smw_params* p, int32_t arity, uint64_t tree_label, const void* root_fr_host, const uint64_t* indices, const void* values_host, size_t n,
                           const uint64_t* proof_indices, size_t n_pidx, int32_t n_levels, const size_t* n_sib, const void* siblings_host,
                           const size_t* n_grp, const uint8_t* group_sizes, int32_t* ok_out) {
    if (!c || !p || !root_fr_host || !ok_out || n_levels < 0 || (n_levels && (!n_sib || !n_grp))) return SMW_ERR_INVALID_ARG;
    cudaSetDevice(c->device);
    *ok_out = 0;
    if (n == 0 || !indices || !values_host) return SMW_OK;                       // :595
    if (arity < 1 || arity > 128 || width_for_arity(arity) != p->t) return SMW_OK;   // :609-617 (arity 1 is the one-leaf final FRI layer; the reference divides by the arity)
    std::vector<uint64_t> req; std::vector<uint8_t> leaves;
    if (!align_request(indices, values_host, n, 1, proof_indices, n_pidx, req, leaves)) return SMW_OK;   // :601
    return verify_core(c, p, arity, tree_label, root_fr_host, req, leaves, {}, n_levels, n_sib, siblings_host, n_grp, group_sizes, ok_out);
}
int32_t smw_verify_pairs_ds(smw_ctx* c, smw_params* p, int32_t arity, uint64_t tree_label, const void* root_fr_host, const uint64_t* indices, const void* pairs_host, size_t n,
                            const uint64_t* proof_indices, size_t n_pidx, int32_t n_levels, const size_t* n_sib, const void* siblings_host,
                            const size_t* n_grp, const uint8_t* group_sizes, int32_t* ok_out) {
    if (!c || !p || !root_fr_host || !ok_out || n_levels < 0 || (n_levels && (!n_sib || !n_grp))) return SMW_ERR_INVALID_ARG;
    cudaSetDevice(c->device);
    *ok_out = 0;
    if (n == 0 || !indices || !pairs_host) return SMW_OK;                        // :733
    if (arity < 1 || arity > 128 || width_for_arity(arity) != p->t) return SMW_OK;   // :739-747
    std::vector<uint64_t> req; std::vector<uint8_t> pairs;
    if (!align_request(indices, pairs_host, n, 2, proof_indices, n_pidx, req, pairs)) return SMW_OK;
    return verify_core(c, p, arity, tree_label, root_fr_host, req, {}, pairs, n_levels, n_sib, siblings_host, n_grp, group_sizes, ok_out);
}

// ---------------------------------------------------------------------------------------- legacy trees
// merkle/src/lib.rs:196-223 new_legacy, :448-470 new_pairs_legacy: fixed t = 17, chunks of poseidon::RATE = 16, every node
// hash_with_ds(chunk, ds_tag) — one permutation, no DS label, no padding.  The tree handle is an ordinary smw_tree (arity 16),
// so smw_tree_open_many serves it (open_union_of_paths is shared, :246-325).
static int legacy_tree(smw_ctx* c, smw_params* p, const void* ds_tag_host, uint32_t* d_level0, size_t n, smw_tree** out) {
    DevFree fr{{}, c->stream};
    uint32_t* d_tag; SMW_CUDA(c, cudaMalloc(&d_tag, FR)); fr.v.push_back(d_tag);
    SMW_CUDA(c, cudaMemcpyAsync(d_tag, ds_tag_host, FR, cudaMemcpyHostToDevice, c->stream));
    auto t = std::make_unique<smw_tree>(); t->ctx = c; t->params = p; t->arity = 16; t->tree_label = 0;
    t->len = level_sizes(n, 16, false);
    size_t total = 0; for (size_t j = 1; j < t->len.size(); j++) total += t->len[j];
    if (total) { cudaError_t e = cudaMalloc(&t->arena, total * FR); if (e != cudaSuccess) { cudaFree(d_level0); return cuda_err(c, e, "cudaMalloc(tree levels)"); } }
    t->lv.push_back(d_level0); t->owned.push_back(true);
    size_t off = 0; for (size_t j = 1; j < t->len.size(); j++) { t->lv.push_back(t->arena + off * 8); t->owned.push_back(false); off += t->len[j]; }
    for (size_t j = 0; j + 1 < t->len.size(); j++) {
        int rc = launch_legacy_nodes(c, p, 0, t->lv[j], nullptr, t->len[j], nullptr, nullptr, t->len[j + 1], d_tag, t->lv[j + 1]);
        if (rc) { smw_tree_free(t.release()); return rc; }
    }
    *out = t.release();
    return SMW_OK;
}
static int legacy_params(smw_ctx* c, smw_params** p) {
    if (!*p) { int rc = smw_params_commitment_default(c, p); if (rc) return rc; }          // merkle/src/lib.rs:364-370 default_params: "POSEIDON-T17-X5-SEED"
    if ((*p)->t != 17) return set_err(c, SMW_ERR_UNSUPPORTED_WIDTH, "legacy trees are fixed at t = 17");
    return SMW_OK;
}
int32_t smw_merkle_new_legacy(smw_ctx* c, smw_params* p, const void* ds_tag_host, const void* leaves, size_t n, int32_t mem, smw_tree** out) {
    if (!c || !out || !ds_tag_host) return SMW_ERR_INVALID_ARG;
    if (n == 0) return set_err(c, SMW_ERR_EMPTY, "no leaves");
    if (!leaves) return SMW_ERR_INVALID_ARG;
    cudaSetDevice(c->device);
    int rc = legacy_params(c, &p); if (rc) return rc;
    uint32_t* d0; SMW_CUDA(c, cudaMalloc(&d0, n * FR));
    { cudaError_t e = cudaMemcpyAsync(d0, leaves, n * FR, mem == SMW_DEVICE ? cudaMemcpyDeviceToDevice : cudaMemcpyHostToDevice, c->stream);
      if (e != cudaSuccess) { cudaFree(d0); return cuda_err(c, e, "copy leaves"); } }
    return legacy_tree(c, p, ds_tag_host, d0, n, out);
}
int32_t smw_merkle_new_pairs_legacy(smw_ctx* c, smw_params* p, const void* ds_tag_host, const void* f, const void* cp, size_t n_f, size_t n_cp, int32_t mem, smw_tree** out) {
    if (!c || !out || !ds_tag_host) return SMW_ERR_INVALID_ARG;
    if (n_f != n_cp) return set_err(c, SMW_ERR_LENGTH_MISMATCH, "f and cp length mismatch");
    if (n_f == 0) return set_err(c, SMW_ERR_EMPTY, "no leaves");
    if (!f || !cp) return SMW_ERR_INVALID_ARG;
    cudaSetDevice(c->device);
    int rc = legacy_params(c, &p); if (rc) return rc;
    In a(c, f, n_f, mem), b(c, cp, n_f, mem), tg(c, ds_tag_host, 1, SMW_HOST); if (a.rc) return a.rc; if (b.rc) return b.rc; if (tg.rc) return tg.rc;
    uint32_t* d0; SMW_CUDA(c, cudaMalloc(&d0, n_f * FR));
    rc = launch_legacy_nodes(c, p, 1, a.d, b.d, n_f, nullptr, nullptr, n_f, tg.d, d0);            // encode_leaf_digest_legacy (:373-375)
    if (rc) { cudaFree(d0); return rc; }
    return legacy_tree(c, p, ds_tag_host, d0, n_f, out);
}
int32_t smw_verify_many_legacy(smw_ctx* c, smw_params* p, int32_t arity, const void* ds_tag_host, const void* root_fr_host, const uint64_t* indices, const void* values_host, size_t n,
                               const uint64_t* proof_indices, size_t n_pidx, int32_t n_levels, const size_t* n_sib, const void* siblings_host,
                               const size_t* n_grp, const uint8_t* group_sizes, int32_t* ok_out) {
    if (!c || !ds_tag_host || !root_fr_host || !ok_out || n_levels < 0 || (n_levels && (!n_sib || !n_grp))) return SMW_ERR_INVALID_ARG;
    cudaSetDevice(c->device);
    *ok_out = 0;
    int rc = legacy_params(c, &p); if (rc) return rc;
    if (n == 0 || !indices || !values_host) return SMW_OK;                       // :495
    if (arity < 1 || arity > 16) return SMW_OK;                                  // more children than the rate cannot be one hash_with_ds block of a tree node
    std::vector<uint64_t> req; std::vector<uint8_t> leaves;
    if (!align_request(indices, values_host, n, 1, proof_indices, n_pidx, req, leaves)) return SMW_OK;   // :502
    return verify_core(c, p, arity, 0, root_fr_host, req, leaves, {}, n_levels, n_sib, siblings_host, n_grp, group_sizes, ok_out, ds_tag_host);
}
int32_t smw_verify_pairs_legacy(smw_ctx* c, smw_params* p, int32_t arity, const void* ds_tag_host, const void* root_fr_host, const uint64_t* indices, const void* pairs_host, size_t n,
                                const uint64_t* proof_indices, size_t n_pidx, int32_t n_levels, const size_t* n_sib, const void* siblings_host,
                                const size_t* n_grp, const uint8_t* group_sizes, int32_t* ok_out) {
    if (!c || !ds_tag_host || !root_fr_host || !ok_out || n_levels < 0 || (n_levels && (!n_sib || !n_grp))) return SMW_ERR_INVALID_ARG;
    cudaSetDevice(c->device);
    *ok_out = 0;
    int rc = legacy_params(c, &p); if (rc) return rc;
    if (n == 0 || !indices || !pairs_host) return SMW_OK;                        // :712
    if (arity < 1 || arity > 16) return SMW_OK;
    std::vector<uint64_t> req; std::vector<uint8_t> pairs;
    if (!align_request(indices, pairs_host, n, 2, proof_indices, n_pidx, req, pairs)) return SMW_OK;
    return verify_core(c, p, arity, 0, root_fr_host, req, {}, pairs, n_levels, n_sib, siblings_host, n_grp, group_sizes, ok_out, ds_tag_host);
}

// ---------------------------------------------------------------------------------------- DEEP-ALI merge
int32_t smw_deep_ali_merge_evals(smw_ctx* c, const void* a, const void* s, const void* e, const void* t, size_t n, const void* omega_host, const void* z_host, void* f0_out, int32_t mem) {
    if (!c || !a || !s || !e || !t || !omega_host || !z_host || !f0_out || n < 2) return SMW_ERR_INVALID_ARG;
    cudaSetDevice(c->device);
    In ia(c, a, n, mem), is(c, s, n, mem), ie(c, e, n, mem), it(c, t, n, mem), iw(c, omega_host, 1, SMW_HOST), iz(c, z_host, 1, SMW_HOST);
    if (ia.rc) return ia.rc; if (is.rc) return is.rc; if (ie.rc) return ie.rc; if (it.rc) return it.rc; if (iw.rc) return iw.rc; if (iz.rc) return iz.rc;
    Out o(c, f0_out, n, mem); if (o.rc) return o.rc;
    int rc = launch_deep_ali_merge(c, ia.d, is.d, ie.d, it.d, n, iw.d, iz.d, o.d); if (rc) return rc;
    return o.finish();
}

} // extern "C"
