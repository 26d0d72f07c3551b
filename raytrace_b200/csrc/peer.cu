// peer.cu — the multi-GPU frame over NVLink peer memory (no NCCL on the per-frame path).
//
// One handle per GPU, world replicated, image split into row bands.  After vxrt_peer_attach a
// frame is:
//
//   trace OWN rows only  ->  k_halo_push: the rows of lighting/depth/normal a neighbour's denoise needs are
//   stored straight into that neighbour's images over NVLink, followed by a release-store of the frame number
//   into its flag word  ->  k_peer_sync waits for the neighbours' flags  ->  the six denoise passes on the
//   shrinking row ranges  ->  k_finalize writes the band straight into the ROOT's output image (peer stores:
//   compute and gather are one kernel)  ->  flag to the root.
//
// So a band never re-traces its 117-row denoise halo (the NCCL-only variant in comm.cu + api.cu does: at 8 x 136
// rows that is 2.7x the rays), and the only things on the wire are the halo rows (11 B/pixel) and the final RGBA8
// band.  Ordering is by monotone frame counters in device memory: data stores -> __threadfence_system ->
// st.release.sys of the counter; consumers spin with ld.acquire.sys (bounded: a timeout sets an error word
// instead of hanging the GPU).  Back-pressure flags (CONSUMED, RELEASED) keep a fast rank from overwriting rows
// or output buffers a slow rank is still reading.
//
// Peer memory is mapped with CUDA IPC when the ranks are processes (torchrun), or used directly when both
// handles live in one process (tests: two bands on one GPU exercise the same kernels and flags).
// The reference has no multi-GPU path (single Vulkan queue, pipeline.rs:229-235).
#include "vxrt_internal.h"
#include <string.h>
#include <unistd.h>
#include <algorithm>
#include <utility>
#include "peer_dev.cuh"

namespace {

#define CU(call)                                                                              \
    do {                                                                                      \
        cudaError_t e_ = (call);                                                              \
        if (e_ != cudaSuccess) { h->err = std::string(#call) + ": " + cudaGetErrorString(e_); return VXRT_ERR_CUDA; } \
    } while (0)

// flag words (uint32 frame counters) in every handle's d_peer_flags
enum { F_ARRIVED = 0, F_CONSUMED = VXRT_PEER_MAX_RANKS, F_BAND_DONE = 2 * VXRT_PEER_MAX_RANKS, F_RELEASED = 3 * VXRT_PEER_MAX_RANKS,
       F_ERROR = 3 * VXRT_PEER_MAX_RANKS + 1, F_FIN_TICKET = 3 * VXRT_PEER_MAX_RANKS + 2, F_BARRIER = 4 * VXRT_PEER_MAX_RANKS,
       F_PUSH_TICKET = 5 * VXRT_PEER_MAX_RANKS, F_COUNT = 5 * VXRT_PEER_MAX_RANKS + VXRT_PEER_MAX_LINKS };

struct PeerBlob {                        // what vxrt_peer_export writes; VXRT_PEER_BLOB_BYTES on the wire
    uint32_t magic, version;
    int32_t pid, device, width, height, row0, row1;
    uint32_t flags;
    uint32_t _pad;
    uint64_t raw[6];                     // same-process fast path: the device pointers themselves
    cudaIpcMemHandle_t ipc[6];           // lighting, depth, normal, out_buf[0], out_buf[1], flags
};
static_assert(sizeof(PeerBlob) <= VXRT_PEER_BLOB_BYTES, "PeerBlob must fit VXRT_PEER_BLOB_BYTES");
constexpr uint32_t kMagic = 0x56585250u;   // "VXRP"

struct SyncList { const uint32_t* wait[VXRT_PEER_MAX_LINKS]; uint32_t* post[VXRT_PEER_MAX_LINKS]; int n_wait, n_post; };

// one warp: lane i waits on wait[i]; then lane i posts to post[i]
__global__ void k_peer_sync(SyncList s, uint32_t wait_value, uint32_t post_value, uint32_t* err) {
    const int i = threadIdx.x;
    if (i < s.n_wait) vx_spin_until(s.wait[i], wait_value, err);
    __syncwarp();
    __threadfence_system();
    if (i < s.n_post) vx_st_release_sys(s.post[i], post_value);
}

// post first, then wait: every rank announces itself before looking at the others
__global__ void k_peer_barrier(SyncList s, uint32_t value, uint32_t* err) {
    const int i = threadIdx.x;
    __threadfence_system();
    if (i < s.n_post) vx_st_release_sys(s.post[i], value);
    if (i < s.n_wait) vx_spin_until(s.wait[i], value, err);
}

struct PushLink {                        // rows [r0,r1) of my band go to one neighbour
    ushort4* lighting; uint16_t* depth; uint8_t* normal;   // the neighbour's images (peer memory)
    uint32_t* arrived;                   // its F_ARRIVED + my rank
    const uint32_t* consumed;            // MY F_CONSUMED + its rank: it has finished the halo of the previous frame
    int r0, r1;
};
struct PushParams {
    PushLink link[VXRT_PEER_MAX_LINKS]; int n;
    const ushort4* lighting; const uint16_t* depth; const uint8_t* normal;
    int width; uint32_t frame; uint32_t* ticket; uint32_t* err;
};

// 16-byte vector copy of a byte range (src/dst share their offset inside 256-byte-aligned allocations)
__device__ void copy_range(uint8_t* dst, const uint8_t* src, size_t n, size_t tid, size_t nthreads) {
    const size_t head = min(n, (size_t)((16 - ((uintptr_t)src & 15)) & 15));
    for (size_t i = tid; i < head; i += nthreads) dst[i] = src[i];
    const size_t body = (n - head) / 16;
    const uint4* s4 = reinterpret_cast<const uint4*>(src + head);
    uint4* d4 = reinterpret_cast<uint4*>(dst + head);
    for (size_t i = tid; i < body; i += nthreads) d4[i] = s4[i];
    for (size_t i = head + body * 16 + tid; i < n; i += nthreads) dst[i] = src[i];
}

// blockIdx.y = link.  Wait until the neighbour has consumed the previous frame's halo, store this frame's rows
// into its images, fence, and let the last CTA of the link publish the frame number.
__global__ void __launch_bounds__(256) k_halo_push(PushParams p) {
    const PushLink& L = p.link[blockIdx.y];
    __shared__ bool last;
    if (threadIdx.x == 0) vx_spin_until(L.consumed, p.frame - 1, p.err);
    __syncthreads();
    const size_t tid = (size_t)blockIdx.x * blockDim.x + threadIdx.x, nt = (size_t)gridDim.x * blockDim.x;
    const size_t a = (size_t)L.r0 * p.width, n = (size_t)(L.r1 - L.r0) * p.width;
    copy_range(reinterpret_cast<uint8_t*>(L.lighting + a), reinterpret_cast<const uint8_t*>(p.lighting + a), n * 8, tid, nt);
    copy_range(reinterpret_cast<uint8_t*>(L.depth + a), reinterpret_cast<const uint8_t*>(p.depth + a), n * 2, tid, nt);
    copy_range(L.normal + a, p.normal + a, n, tid, nt);
    __syncthreads();                                       // the CTA's stores happen-before thread 0's fence
    if (threadIdx.x == 0) { __threadfence_system(); last = atomicAdd(p.ticket + blockIdx.y, 1u) == gridDim.x - 1; }
    __syncthreads();
    if (last && threadIdx.x == 0) {
        p.ticket[blockIdx.y] = 0;
        __threadfence_system();
        vx_st_release_sys(L.arrived, p.frame);
    }
}

int open_peer(VxrtHandle* h, const PeerBlob& b, int which, void** out) {
    if (b.pid == (int32_t)getpid()) {                      // same process: the pointer is valid as is
        *out = reinterpret_cast<void*>(b.raw[which]);
        if (b.device != h->cfg.device) {
            int can = 0;
            CU(cudaDeviceCanAccessPeer(&can, h->cfg.device, b.device));
            if (!can) { h->err = "vxrt_peer_attach: devices cannot access each other's memory"; return VXRT_ERR_COMM; }
            cudaError_t e = cudaDeviceEnablePeerAccess(b.device, 0);
            if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled) { h->err = std::string("cudaDeviceEnablePeerAccess: ") + cudaGetErrorString(e); return VXRT_ERR_CUDA; }
            cudaGetLastError();
        }
        return VXRT_OK;
    }
    cudaError_t e = cudaIpcOpenMemHandle(out, b.ipc[which], cudaIpcMemLazyEnablePeerAccess);
    if (e != cudaSuccess) { h->err = std::string("cudaIpcOpenMemHandle: ") + cudaGetErrorString(e); cudaGetLastError(); return VXRT_ERR_COMM; }
    h->peer_ipc_opened.push_back(*out);
    return VXRT_OK;
}

}  // namespace

// CUDA loads a kernel's code at its first launch, and that load cannot complete while another kernel is resident
// and spinning on a flag (peer.cu).  Every kernel of this file is therefore loaded when a handle is created.
#define VXRT_PRELOAD(k) do { cudaFuncAttributes a_; cudaError_t e_ = cudaFuncGetAttributes(&a_, k); if (e_ != cudaSuccess) return e_; } while (0)
cudaError_t vxrt_preload_peer() { VXRT_PRELOAD(k_peer_sync); VXRT_PRELOAD(k_peer_barrier); VXRT_PRELOAD(k_halo_push); return cudaSuccess; }

void vxrt_peer_release(VxrtHandle* h) {
    for (void* p : h->peer_ipc_opened) cudaIpcCloseMemHandle(p);
    h->peer_ipc_opened.clear();
    if (h->peer_stream) { cudaStreamDestroy(h->peer_stream); h->peer_stream = nullptr; }
    if (h->ev_peer) { cudaEventDestroy(h->ev_peer); h->ev_peer = nullptr; }
    if (h->d_peer_flags) { cudaFree(h->d_peer_flags); h->d_peer_flags = nullptr; }
    h->peer_on = false;
}

// ---- called by vxrt_render_frame (api.cu) when peer mode is on ---------------------------------------------

// start of a frame on the root: once the read-back that used this frame's output buffer has drained, tell every
// rank it may store frame `f` into it
int vxrt_peer_begin_frame(VxrtHandle* h, int out_idx) {
    h->peer_frame++;
    if (h->rank != h->peer_root || !(h->cfg.flags & VXRT_FLAG_FINALIZE)) return VXRT_OK;
    CU(cudaEventRecord(h->ev_peer, h->stream));                          // earlier work of this handle (sync read-backs)
    CU(cudaStreamWaitEvent(h->peer_stream, h->ev_peer, 0));
    if (h->copy_pending[out_idx]) CU(cudaStreamWaitEvent(h->peer_stream, h->ev_copy_done[out_idx], 0));
    SyncList s; memset(&s, 0, sizeof s);
    for (int r = 0; r < h->nranks; r++) if (r != h->rank) s.post[s.n_post++] = h->peer_flags[r] + F_RELEASED;
    k_peer_sync<<<1, 32, 0, h->peer_stream>>>(s, 0, h->peer_frame, h->d_peer_flags + F_ERROR);
    CU(cudaGetLastError());
    return VXRT_OK;
}

// Rows of this band some neighbour's denoise needs, merged into disjoint ascending intervals (a0,b0,a1,b1,...).
// vxrt_render_frame traces these first and pushes them, then traces the interior, so a neighbour's halo is normally
// on its way long before that neighbour finishes its own trace pass.
int vxrt_peer_boundary_rows(VxrtHandle* h, int32_t* iv /* 2 * VXRT_PEER_MAX_LINKS */) {
    if (!(h->cfg.flags & VXRT_FLAG_DENOISE)) return 0;
    std::vector<std::pair<int, int>> v;
    for (const VxrtPeerLink& d : h->peer_dst) v.push_back({d.r0, d.r1});
    std::sort(v.begin(), v.end());
    int n = 0;
    for (const auto& p : v) {
        if (n && p.first <= iv[2 * n - 1]) { if (p.second > iv[2 * n - 1]) iv[2 * n - 1] = p.second; }
        else { iv[2 * n] = p.first; iv[2 * n + 1] = p.second; n++; }
    }
    return n;
}

// after the boundary rows are traced: store them into the neighbours' images
int vxrt_peer_push_halo(VxrtHandle* h) {
    if (!(h->cfg.flags & VXRT_FLAG_DENOISE)) return VXRT_OK;
    if (!h->peer_dst.empty()) {
        PushParams p; memset(&p, 0, sizeof p);
        int rows = 0;
        for (size_t i = 0; i < h->peer_dst.size(); i++) {
            const VxrtPeerLink& d = h->peer_dst[i];
            PushLink& L = p.link[p.n++];
            L.lighting = (ushort4*)d.lighting; L.depth = (uint16_t*)d.depth; L.normal = (uint8_t*)d.normal;
            L.arrived = h->peer_flags[d.rank] + F_ARRIVED + h->rank;
            L.consumed = h->d_peer_flags + F_CONSUMED + d.rank;
            L.r0 = d.r0; L.r1 = d.r1;
            rows = rows > d.r1 - d.r0 ? rows : d.r1 - d.r0;
        }
        p.lighting = h->d_lighting; p.depth = h->d_depth; p.normal = h->d_normal;
        p.width = h->cfg.width; p.frame = h->peer_frame; p.ticket = h->d_peer_flags + F_PUSH_TICKET; p.err = h->d_peer_flags + F_ERROR;
        const size_t bytes = (size_t)rows * h->cfg.width * 8;
        int ctas = (int)((bytes / 16 + 256 * 4 - 1) / (256 * 4));         // ~4 vectors per thread
        ctas = ctas < 1 ? 1 : ctas > 148 ? 148 : ctas;
        k_halo_push<<<dim3(ctas, p.n), 256, 0, h->stream>>>(p);
        CU(cudaGetLastError());
    }
    return VXRT_OK;
}

// before the first denoise pass: the neighbours' rows must have landed
int vxrt_peer_wait_halo(VxrtHandle* h) {
    if (!(h->cfg.flags & VXRT_FLAG_DENOISE)) return VXRT_OK;
    if (!h->peer_src.empty()) {
        SyncList s; memset(&s, 0, sizeof s);
        for (const VxrtPeerLink& l : h->peer_src) s.wait[s.n_wait++] = h->d_peer_flags + F_ARRIVED + l.rank;
        k_peer_sync<<<1, 32, 0, h->stream>>>(s, h->peer_frame, 0, h->d_peer_flags + F_ERROR);
        CU(cudaGetLastError());
    }
    return VXRT_OK;
}

// The frame's last signals ride in k_finalize (post.cu): its prologue waits for the root's RELEASED word (non-root),
// its last CTA posts CONSUMED to the halo sources and BAND_DONE to the root.  Returns the tail to hand to the launch.
VxrtPeerTail vxrt_peer_finalize_tail(VxrtHandle* h) {
    VxrtPeerTail t; memset(&t, 0, sizeof t);
    t.value = h->peer_frame; t.err = h->d_peer_flags + F_ERROR; t.ticket = h->d_peer_flags + F_FIN_TICKET;
    if (h->cfg.flags & VXRT_FLAG_DENOISE)
        for (const VxrtPeerLink& l : h->peer_src) t.post[t.n_post++] = h->peer_flags[l.rank] + F_CONSUMED + h->rank;
    if (h->rank != h->peer_root) {
        t.wait = h->d_peer_flags + F_RELEASED;
        t.post[t.n_post++] = h->peer_flags[h->peer_root] + F_BAND_DONE + h->rank;
    }
    return t;
}

// without a finalize pass the CONSUMED words still have to be posted after the last denoise pass
int vxrt_peer_after_denoise(VxrtHandle* h) {
    if ((h->cfg.flags & VXRT_FLAG_FINALIZE) || !(h->cfg.flags & VXRT_FLAG_DENOISE) || h->peer_src.empty()) return VXRT_OK;
    SyncList s; memset(&s, 0, sizeof s);
    for (const VxrtPeerLink& l : h->peer_src) s.post[s.n_post++] = h->peer_flags[l.rank] + F_CONSUMED + h->rank;
    k_peer_sync<<<1, 32, 0, h->stream>>>(s, 0, h->peer_frame, h->d_peer_flags + F_ERROR);
    CU(cudaGetLastError());
    return VXRT_OK;
}

// where k_finalize stores this rank's band: the root's output image of the same parity
uchar4* vxrt_peer_output(VxrtHandle* h, int out_idx) {
    return h->rank == h->peer_root ? h->d_out_buf[out_idx] : (uchar4*)h->peer_root_out[out_idx];
}

// after k_finalize on the root: wait for every band
int vxrt_peer_end_frame(VxrtHandle* h) {
    if (!(h->cfg.flags & VXRT_FLAG_FINALIZE) || h->rank != h->peer_root || h->nranks == 1) return VXRT_OK;
    SyncList s; memset(&s, 0, sizeof s);
    for (int r = 0; r < h->nranks; r++) if (r != h->rank) s.wait[s.n_wait++] = h->d_peer_flags + F_BAND_DONE + r;
    k_peer_sync<<<1, 32, 0, h->stream>>>(s, h->peer_frame, 0, h->d_peer_flags + F_ERROR);
    CU(cudaGetLastError());
    return VXRT_OK;
}

int vxrt_peer_error_word(VxrtHandle* h, uint32_t* out) {
    *out = 0;
    if (!h->peer_on) return VXRT_OK;
    CU(cudaMemcpyAsync(out, h->d_peer_flags + F_ERROR, 4, cudaMemcpyDeviceToHost, h->stream));
    CU(cudaStreamSynchronize(h->stream));
    return VXRT_OK;
}

void vxrt_peer_clear_error(VxrtHandle* h) {
    if (h->peer_on) cudaMemsetAsync(h->d_peer_flags + F_ERROR, 0, 4, h->stream);
}

extern "C" {

int vxrt_peer_export(VxrtHandle* h, void* blob, size_t nbytes) {
    if (!h) return VXRT_ERR_INVALID;
    if (!blob || nbytes < VXRT_PEER_BLOB_BYTES) { h->err = "vxrt_peer_export: blob must be VXRT_PEER_BLOB_BYTES bytes"; return VXRT_ERR_SIZE; }
    if (h->peer_on) { h->err = "vxrt_peer_export: already attached"; return VXRT_ERR_STATE; }
    CU(cudaSetDevice(h->cfg.device));
    if (!h->d_peer_flags) {
        CU(cudaMalloc((void**)&h->d_peer_flags, F_COUNT * sizeof(uint32_t)));
        CU(cudaMemset(h->d_peer_flags, 0, F_COUNT * sizeof(uint32_t)));
    }
    PeerBlob b; memset(&b, 0, sizeof b);
    b.magic = kMagic; b.version = VXRT_ABI_VERSION; b.pid = (int32_t)getpid(); b.device = h->cfg.device;
    b.width = h->cfg.width; b.height = h->cfg.height; b.row0 = h->row0; b.row1 = h->row1; b.flags = h->cfg.flags;
    void* ptrs[6] = {h->d_lighting, h->d_depth, h->d_normal, h->d_out_buf[0], h->d_out_buf[1], h->d_peer_flags};
    for (int i = 0; i < 6; i++) {
        b.raw[i] = (uint64_t)(uintptr_t)ptrs[i];
        CU(cudaIpcGetMemHandle(&b.ipc[i], ptrs[i]));
    }
    memset(blob, 0, VXRT_PEER_BLOB_BYTES);
    memcpy(blob, &b, sizeof b);
    return VXRT_OK;
}

// device-side barrier over all ranks on the handle's stream (benchmarks: align the ranks before a timed frame)
int vxrt_peer_barrier(VxrtHandle* h) {
    if (!h) return VXRT_ERR_INVALID;
    if (!h->peer_on) { h->err = "vxrt_peer_barrier: vxrt_peer_attach not called"; return VXRT_ERR_STATE; }
    CU(cudaSetDevice(h->cfg.device));
    h->peer_barriers++;
    SyncList s; memset(&s, 0, sizeof s);
    for (int r = 0; r < h->nranks; r++) {
        if (r == h->rank) continue;
        s.post[s.n_post++] = h->peer_flags[r] + F_BARRIER + h->rank;
        s.wait[s.n_wait++] = h->d_peer_flags + F_BARRIER + r;
    }
    k_peer_barrier<<<1, 32, 0, h->stream>>>(s, h->peer_barriers, h->d_peer_flags + F_ERROR);
    CU(cudaGetLastError());
    return VXRT_OK;
}

int vxrt_peer_attach(VxrtHandle* h, const void* blobs, size_t nbytes, int rank, int nranks, int root) {
    if (!h) return VXRT_ERR_INVALID;
    if (nranks < 1 || nranks > VXRT_PEER_MAX_RANKS || rank < 0 || rank >= nranks || root < 0 || root >= nranks) {
        h->err = "vxrt_peer_attach: bad rank / nranks / root"; return VXRT_ERR_INVALID;
    }
    if (!blobs || nbytes < (size_t)nranks * VXRT_PEER_BLOB_BYTES) { h->err = "vxrt_peer_attach: need nranks blobs"; return VXRT_ERR_SIZE; }
    if (h->peer_on) { h->err = "vxrt_peer_attach: already attached"; return VXRT_ERR_STATE; }
    if (h->cfg.flags & VXRT_FLAG_REPROJECT) { h->err = "vxrt_peer_attach: VXRT_FLAG_REPROJECT needs the whole frame on one handle"; return VXRT_ERR_INVALID; }
    if (!h->d_peer_flags) { h->err = "vxrt_peer_attach: vxrt_peer_export not called on this handle"; return VXRT_ERR_STATE; }
    CU(cudaSetDevice(h->cfg.device));
    std::vector<PeerBlob> all(nranks);
    for (int r = 0; r < nranks; r++) {
        memcpy(&all[r], (const uint8_t*)blobs + (size_t)r * VXRT_PEER_BLOB_BYTES, sizeof(PeerBlob));
        const PeerBlob& b = all[r];
        if (b.magic != kMagic || b.version != VXRT_ABI_VERSION) { h->err = "vxrt_peer_attach: bad blob"; return VXRT_ERR_INVALID; }
        if (b.width != h->cfg.width || b.height != h->cfg.height || ((b.flags ^ h->cfg.flags) & (VXRT_FLAG_DENOISE | VXRT_FLAG_FINALIZE))) {
            h->err = "vxrt_peer_attach: ranks disagree on the frame size or passes"; return VXRT_ERR_INVALID;
        }
    }
    if (all[rank].row0 != h->row0 || all[rank].row1 != h->row1 || all[rank].raw[5] != (uint64_t)(uintptr_t)h->d_peer_flags) {
        h->err = "vxrt_peer_attach: blob[rank] is not this handle's export"; return VXRT_ERR_INVALID;
    }
    h->bands.assign((size_t)nranks * 2, 0);
    for (int r = 0; r < nranks; r++) { h->bands[2 * r] = all[r].row0; h->bands[2 * r + 1] = all[r].row1; }
    h->peer_flags.assign(nranks, nullptr);
    h->peer_flags[rank] = h->d_peer_flags;
    h->peer_dst.clear(); h->peer_src.clear();
    int rc;
    for (int r = 0; r < nranks; r++) {
        if (r == rank) continue;
        int32_t plan[14];
        vxrt_plan_passes(h->cfg.height, all[r].row0, all[r].row1, h->cfg.flags, plan);
        // rows of MY band inside r's trace range: r needs them; rows of r's band inside MY trace range: I need them
        const int d0 = h->row0 > plan[0] ? h->row0 : plan[0], d1 = h->row1 < plan[1] ? h->row1 : plan[1];
        const int s0 = all[r].row0 > h->plan[0] ? all[r].row0 : h->plan[0], s1 = all[r].row1 < h->plan[1] ? all[r].row1 : h->plan[1];
        const bool dst = (h->cfg.flags & VXRT_FLAG_DENOISE) && d0 < d1, src = (h->cfg.flags & VXRT_FLAG_DENOISE) && s0 < s1;
        const bool need_flags = true;                      // vxrt_peer_barrier posts to every rank
        if (need_flags) {
            void* f = nullptr;
            if ((rc = open_peer(h, all[r], 5, &f))) return rc;
            h->peer_flags[r] = (uint32_t*)f;
        }
        if (dst) {
            if (h->peer_dst.size() >= VXRT_PEER_MAX_LINKS) { h->err = "vxrt_peer_attach: bands thinner than the denoise halo need too many links"; return VXRT_ERR_INVALID; }
            VxrtPeerLink l; l.rank = r; l.r0 = d0; l.r1 = d1;
            if ((rc = open_peer(h, all[r], 0, &l.lighting)) || (rc = open_peer(h, all[r], 1, &l.depth)) || (rc = open_peer(h, all[r], 2, &l.normal))) return rc;
            h->peer_dst.push_back(l);
        }
        if (src) {
            if (h->peer_src.size() >= VXRT_PEER_MAX_LINKS) { h->err = "vxrt_peer_attach: too many halo sources"; return VXRT_ERR_INVALID; }
            VxrtPeerLink l; l.rank = r; l.r0 = s0; l.r1 = s1; l.lighting = l.depth = l.normal = nullptr;
            h->peer_src.push_back(l);
        }
        if (r == root && (h->cfg.flags & VXRT_FLAG_FINALIZE)) {
            if ((rc = open_peer(h, all[r], 3, &h->peer_root_out[0])) || (rc = open_peer(h, all[r], 4, &h->peer_root_out[1]))) return rc;
        }
    }
    if (nranks - 1 > VXRT_PEER_MAX_LINKS && rank == root) { h->err = "vxrt_peer_attach: too many ranks for one sync list"; return VXRT_ERR_INVALID; }
    CU(cudaStreamCreateWithFlags(&h->peer_stream, cudaStreamNonBlocking));
    CU(cudaEventCreateWithFlags(&h->ev_peer, cudaEventDisableTiming));
    h->rank = rank; h->nranks = nranks; h->peer_root = root; h->peer_frame = 0;
    h->peer_on = true;
    // the band no longer traces its own halo: the neighbours deliver it
    h->trace_row0 = h->row0; h->trace_row1 = h->row1;
    return VXRT_OK;
}

}  // extern "C"
