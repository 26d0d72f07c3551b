// comm.cu — multi-GPU plumbing of libvxrt: one handle (= one process, one GPU) per rank, the world
// replicated, the image split into row bands (vxrt_create: row_begin/row_end).  The data path has two
// exchange steps and nothing else: the final RGBA8 bands are gathered over NVLink, and voxel edit
// lists are broadcast so that every replica rebuilds its own occupancy field.
// The reference has no multi-GPU path at all (single Vulkan queue, pipeline.rs:229-235).
//
// NCCL is bound at run time (dlopen of libnccl.so.2) so that libvxrt.so loads on a single-GPU host
// without NCCL; only the types come from <nccl.h>.
#include "vxrt_internal.h"
#include <dlfcn.h>
#include <nccl.h>
#include <string.h>

namespace {

struct NcclApi {
    void* lib = nullptr;
    ncclResult_t (*GetUniqueId)(ncclUniqueId*) = nullptr;
    ncclResult_t (*CommInitRank)(ncclComm_t*, int, ncclUniqueId, int) = nullptr;
    ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
    ncclResult_t (*Send)(const void*, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*Recv)(void*, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*Broadcast)(const void*, void*, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*AllGather)(const void*, void*, size_t, ncclDataType_t, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*GroupStart)() = nullptr;
    ncclResult_t (*GroupEnd)() = nullptr;
    const char* (*GetErrorString)(ncclResult_t) = nullptr;
    std::string err;
};
NcclApi g_nccl;

bool load_nccl() {
    if (g_nccl.lib) return true;
    // RTLD_NOLOAD first: reuse the copy a host process (e.g. torch) already mapped
    void* lib = dlopen("libnccl.so.2", RTLD_NOW | RTLD_NOLOAD);
    if (!lib) lib = dlopen("libnccl.so.2", RTLD_NOW | RTLD_GLOBAL);
    if (!lib) lib = dlopen("libnccl.so", RTLD_NOW | RTLD_GLOBAL);
    if (!lib) { g_nccl.err = std::string("dlopen(libnccl.so.2) failed: ") + dlerror(); return false; }
#define SYM(field, name)                                                                     \
    g_nccl.field = reinterpret_cast<decltype(g_nccl.field)>(dlsym(lib, name));               \
    if (!g_nccl.field) { g_nccl.err = std::string("libnccl lacks ") + name; return false; }
    SYM(GetUniqueId, "ncclGetUniqueId") SYM(CommInitRank, "ncclCommInitRank") SYM(CommDestroy, "ncclCommDestroy")
    SYM(Send, "ncclSend") SYM(Recv, "ncclRecv") SYM(Broadcast, "ncclBroadcast") SYM(AllGather, "ncclAllGather")
    SYM(GroupStart, "ncclGroupStart") SYM(GroupEnd, "ncclGroupEnd") SYM(GetErrorString, "ncclGetErrorString")
#undef SYM
    g_nccl.lib = lib;
    return true;
}

#define CU(call)                                                                              \
    do {                                                                                      \
        cudaError_t e_ = (call);                                                              \
        if (e_ != cudaSuccess) { h->err = std::string(#call) + ": " + cudaGetErrorString(e_); return VXRT_ERR_CUDA; } \
    } while (0)
#define NC(call)                                                                              \
    do {                                                                                      \
        ncclResult_t r_ = g_nccl.call;                                                        \
        if (r_ != ncclSuccess) { h->err = std::string("nccl" #call ": ") + g_nccl.GetErrorString(r_); return VXRT_ERR_COMM; } \
    } while (0)

}  // namespace

void vxrt_comm_release(VxrtHandle* h) {
    if (h->comm && g_nccl.lib) g_nccl.CommDestroy((ncclComm_t)h->comm);
    h->comm = nullptr;
}

extern "C" {

int vxrt_comm_unique_id(void* id_out, size_t nbytes) {
    if (!id_out || nbytes < sizeof(ncclUniqueId)) return VXRT_ERR_SIZE;
    if (!load_nccl()) return VXRT_ERR_COMM;
    ncclUniqueId id;
    if (g_nccl.GetUniqueId(&id) != ncclSuccess) return VXRT_ERR_COMM;
    memcpy(id_out, &id, sizeof id);
    return VXRT_OK;
}

int vxrt_comm_init(VxrtHandle* h, const void* id, size_t nbytes, int rank, int nranks) {
    if (!h) return VXRT_ERR_INVALID;
    if (!id || nbytes < sizeof(ncclUniqueId)) { h->err = "vxrt_comm_init: id must be VXRT_COMM_ID_BYTES bytes"; return VXRT_ERR_SIZE; }
    if (nranks < 1 || rank < 0 || rank >= nranks) { h->err = "vxrt_comm_init: bad rank / nranks"; return VXRT_ERR_INVALID; }
    if (h->comm) { h->err = "vxrt_comm_init: communicator already initialised"; return VXRT_ERR_STATE; }
    if (!load_nccl()) { h->err = g_nccl.err; return VXRT_ERR_COMM; }
    CU(cudaSetDevice(h->cfg.device));
    ncclUniqueId uid;
    memcpy(&uid, id, sizeof uid);
    ncclComm_t comm;
    NC(CommInitRank(&comm, nranks, uid, rank));
    h->comm = comm; h->rank = rank; h->nranks = nranks;
    // every rank learns every band: all-gather of (row0, row1)
    int32_t* d_bands = nullptr;
    CU(cudaMalloc((void**)&d_bands, (size_t)nranks * 2 * sizeof(int32_t)));
    const int32_t mine[2] = {h->row0, h->row1};
    CU(cudaMemcpyAsync(d_bands + 2 * rank, mine, sizeof mine, cudaMemcpyHostToDevice, h->stream));
    NC(AllGather(d_bands + 2 * rank, d_bands, 2, ncclInt32, comm, h->stream));
    h->bands.assign((size_t)nranks * 2, 0);
    CU(cudaMemcpyAsync(h->bands.data(), d_bands, (size_t)nranks * 2 * sizeof(int32_t), cudaMemcpyDeviceToHost, h->stream));
    CU(cudaStreamSynchronize(h->stream));
    CU(cudaFree(d_bands));
    return VXRT_OK;
}

// Output row of G-buffer row y is height-1-y (finalize.comp:60-63), so band [r0,r1) is the contiguous
// byte range [(H-r1)*W*4, (H-r0)*W*4) of the RGBA8 image.
int vxrt_comm_gather_output(VxrtHandle* h, int root) {
    if (!h) return VXRT_ERR_INVALID;
    if (!h->comm) { h->err = "vxrt_comm_gather_output: vxrt_comm_init not called"; return VXRT_ERR_STATE; }
    if (root >= h->nranks) { h->err = "vxrt_comm_gather_output: bad root"; return VXRT_ERR_INVALID; }
    if (h->nranks == 1) return VXRT_OK;
    CU(cudaSetDevice(h->cfg.device));
    ncclComm_t comm = (ncclComm_t)h->comm;
    const size_t row_bytes = (size_t)h->cfg.width * 4;
    uint8_t* img = reinterpret_cast<uint8_t*>(h->d_output);
    auto off = [&](int r) { return (size_t)(h->cfg.height - h->bands[2 * r + 1]) * row_bytes; };
    auto len = [&](int r) { return (size_t)(h->bands[2 * r + 1] - h->bands[2 * r]) * row_bytes; };
    NC(GroupStart());
    for (int r = 0; r < h->nranks; r++) {
        if (r == h->rank) continue;
        const bool i_collect = root < 0 || root == h->rank, r_collects = root < 0 || root == r;
        if (r_collects) NC(Send(img + off(h->rank), len(h->rank), ncclUint8, r, comm, h->stream));
        if (i_collect) NC(Recv(img + off(r), len(r), ncclUint8, r, comm, h->stream));
    }
    NC(GroupEnd());
    return VXRT_OK;
}

int vxrt_comm_broadcast_edits(VxrtHandle* h, const VxrtEdit* edits, size_t n, int root) {
    if (!h) return VXRT_ERR_INVALID;
    if (!h->comm) { h->err = "vxrt_comm_broadcast_edits: vxrt_comm_init not called"; return VXRT_ERR_STATE; }
    if (root < 0 || root >= h->nranks) { h->err = "vxrt_comm_broadcast_edits: bad root"; return VXRT_ERR_INVALID; }
    if (h->rank == root && n && !edits) { h->err = "vxrt_comm_broadcast_edits: edits is NULL on the root"; return VXRT_ERR_INVALID; }
    if (n == 0) return VXRT_OK;
    CU(cudaSetDevice(h->cfg.device));
    if (n > h->edits_cap) {
        if (h->d_edits) CU(cudaFree(h->d_edits));
        h->d_edits = nullptr; h->edits_cap = 0;
        size_t cap = 1024; while (cap < n) cap *= 2;
        CU(cudaMalloc(&h->d_edits, cap * sizeof(VxrtEdit)));
        h->edits_cap = cap;
    }
    if (h->rank == root) CU(cudaMemcpyAsync(h->d_edits, edits, n * sizeof(VxrtEdit), cudaMemcpyHostToDevice, h->stream));
    NC(Broadcast(h->d_edits, h->d_edits, n * sizeof(VxrtEdit), ncclUint8, root, (ncclComm_t)h->comm, h->stream));
    const int rc = vxrt_apply_edits_device(h, h->d_edits, n);
    if (rc) return rc;
    CU(cudaStreamSynchronize(h->stream));
    return VXRT_OK;
}

}  // extern "C"
