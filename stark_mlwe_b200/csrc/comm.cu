// comm.cu — the multi-GPU data plane of the library: a communicator handle over NCCL (NVLink 5 / NVSwitch on the box) that the
// sharded entry points of include/smw.h take, so that a host in any language shards the path without re-implementing the
// exchange (round 1 did it in Python over torch.distributed, through host memory).
//
// NCCL is bound at run time (dlopen of libnccl.so.2): the library loads on a machine without NCCL, a process that already
// holds an NCCL (e.g. the one PyTorch bundles) shares it instead of loading a second copy, and every entry point that needs a
// communicator fails with SMW_ERR_COMM when there is none.  Bootstrap is NCCL's own: rank 0 makes a 128-byte unique id
// (smw_comm_unique_id), the host carries it to the other ranks by whatever means it has, every rank calls smw_comm_init.
#include "smw_internal.h"
#include <dlfcn.h>
#include <cstring>

namespace smw {

static bool load_nccl(NcclApi& a, std::string& err) {
    const char* names[] = {"libnccl.so.2", "libnccl.so"};
    void* h = nullptr;
    for (const char* n : names) { h = dlopen(n, RTLD_NOW | RTLD_GLOBAL); if (h) break; }
    if (!h) { err = std::string("dlopen(libnccl.so.2): ") + (dlerror() ? dlerror() : "not found"); return false; }
#define SMW_NCCL_SYM(field, sym) a.field = reinterpret_cast<decltype(a.field)>(dlsym(h, sym)); if (!a.field) { err = std::string("dlsym(") + sym + ") failed"; return false; }
    SMW_NCCL_SYM(GetUniqueId, "ncclGetUniqueId")
    SMW_NCCL_SYM(CommInitRank, "ncclCommInitRank")
    SMW_NCCL_SYM(CommDestroy, "ncclCommDestroy")
    SMW_NCCL_SYM(AllGather, "ncclAllGather")
    SMW_NCCL_SYM(Send, "ncclSend")
    SMW_NCCL_SYM(Recv, "ncclRecv")
    SMW_NCCL_SYM(GroupStart, "ncclGroupStart")
    SMW_NCCL_SYM(GroupEnd, "ncclGroupEnd")
    SMW_NCCL_SYM(GetErrorString, "ncclGetErrorString")
#undef SMW_NCCL_SYM
    return true;
}
const NcclApi* nccl_api(std::string* err) {
    static NcclApi api; static int state = 0; static std::string why; static std::mutex mu;
    std::lock_guard<std::mutex> lk(mu);
    if (state == 0) state = load_nccl(api, why) ? 1 : -1;
    if (state < 0) { if (err) *err = why; return nullptr; }
    return &api;
}
int nccl_err(smw_ctx* c, const NcclApi* a, ncclResult_t r, const char* what) {
    if (c) c->last_error = std::string(what) + ": " + (a ? a->GetErrorString(r) : "NCCL unavailable");
    return SMW_ERR_COMM;
}

} // namespace smw

using namespace smw;

extern "C" {

int32_t smw_comm_unique_id(uint8_t* out128) {
    if (!out128) return SMW_ERR_INVALID_ARG;
    std::string err; const NcclApi* a = nccl_api(&err); if (!a) return SMW_ERR_COMM;
    ncclUniqueId id; static_assert(sizeof(ncclUniqueId) == 128, "ncclUniqueId is 128 bytes");
    if (a->GetUniqueId(&id) != ncclSuccess) return SMW_ERR_COMM;
    memcpy(out128, &id, 128);
    return SMW_OK;
}
int32_t smw_comm_init(smw_ctx* c, const uint8_t* id128, int32_t world, int32_t rank, smw_comm** out) {
    if (!c || !id128 || !out || world < 1 || rank < 0 || rank >= world) return SMW_ERR_INVALID_ARG;
    *out = nullptr;
    cudaSetDevice(c->device);
    std::string err; const NcclApi* a = nccl_api(&err); if (!a) return set_err(c, SMW_ERR_COMM, err.c_str());
    ncclUniqueId id; memcpy(&id, id128, 128);
    ncclComm_t comm;
    ncclResult_t r = a->CommInitRank(&comm, world, id, rank);
    if (r != ncclSuccess) return nccl_err(c, a, r, "ncclCommInitRank");
    smw_comm* m = new smw_comm(); m->ctx = c; m->comm = comm; m->rank = rank; m->world = world;
    *out = m;
    return SMW_OK;
}
int32_t smw_comm_rank(const smw_comm* m) { return m ? m->rank : 0; }
int32_t smw_comm_world(const smw_comm* m) { return m ? m->world : 1; }
void smw_comm_free(smw_comm* m) {
    if (!m) return;
    cudaSetDevice(m->ctx->device); cudaStreamSynchronize(m->ctx->stream);
    if (const NcclApi* a = nccl_api(nullptr)) a->CommDestroy(m->comm);
    delete m;
}
// recv[r * n_fr .. (r + 1) * n_fr) = rank r's send[0 .. n_fr); device buffers, on the context's stream (no host synchronisation)
int32_t smw_comm_all_gather(smw_comm* m, const void* send_dev, void* recv_dev, size_t n_fr) {
    if (!m || (n_fr && (!send_dev || !recv_dev))) return SMW_ERR_INVALID_ARG;
    smw_ctx* c = m->ctx; cudaSetDevice(c->device);
    return comm_all_gather_bytes(c, m, send_dev, recv_dev, n_fr * 32);
}

} // extern "C"

namespace smw {
int comm_all_gather_bytes(smw_ctx* c, smw_comm* m, const void* send, void* recv, size_t bytes) {
    if (bytes == 0) return SMW_OK;
    if (!m || m->world == 1) { if (send != recv) SMW_CUDA(c, cudaMemcpyAsync(recv, send, bytes, cudaMemcpyDeviceToDevice, c->stream)); return SMW_OK; }
    const NcclApi* a = nccl_api(nullptr); if (!a) return set_err(c, SMW_ERR_COMM, "NCCL unavailable");
    ncclResult_t r = a->AllGather(send, recv, bytes, ncclUint8, m->comm, c->stream);
    if (r != ncclSuccess) return nccl_err(c, a, r, "ncclAllGather");
    c->collectives++;
    return SMW_OK;
}
// piece p of `send` (piece_bytes each) goes to rank p; piece p of `recv` comes from rank p
int comm_all_to_all_bytes(smw_ctx* c, smw_comm* m, const void* send, void* recv, size_t piece_bytes) {
    if (piece_bytes == 0) return SMW_OK;
    if (!m || m->world == 1) { SMW_CUDA(c, cudaMemcpyAsync(recv, send, piece_bytes, cudaMemcpyDeviceToDevice, c->stream)); return SMW_OK; }
    const NcclApi* a = nccl_api(nullptr); if (!a) return set_err(c, SMW_ERR_COMM, "NCCL unavailable");
    ncclResult_t r = a->GroupStart(); if (r != ncclSuccess) return nccl_err(c, a, r, "ncclGroupStart");
    for (int p = 0; p < m->world && r == ncclSuccess; p++) {
        r = a->Send((const uint8_t*)send + (size_t)p * piece_bytes, piece_bytes, ncclUint8, p, m->comm, c->stream);
        if (r == ncclSuccess) r = a->Recv((uint8_t*)recv + (size_t)p * piece_bytes, piece_bytes, ncclUint8, p, m->comm, c->stream);
    }
    ncclResult_t e = a->GroupEnd();
    if (r != ncclSuccess) return nccl_err(c, a, r, "ncclSend/ncclRecv");
    if (e != ncclSuccess) return nccl_err(c, a, e, "ncclGroupEnd");
    c->collectives++;
    return SMW_OK;
}
} // namespace smw
