// vxrt_internal.h — handle layout and launch prototypes shared by the .cu files of libvxrt.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <string>
#include <vector>
#include "../../include/vxrt.h"

#define VXRT_NOISE_SIZE 512
#define VXRT_PEER_MAX_RANKS 16
#define VXRT_PEER_MAX_LINKS 16
#define VXRT_CHUNK 64
// value tables behind the four diffuse tables in TraceParams::lut (api.cu value_tables; trace_fast.cu (E12))
#define VXRT_LUT_JITTER 1024            // [256] n / 255 * 0.05   (raytrace.comp:186)
#define VXRT_LUT_ALBEDO 1280            // [128] n / 127          (:156-158)
#define VXRT_LUT_SCREEN 1408            // [width] px / width * 2 - 1, then [height] py / height * 2 - 1   (:296-297)

// trace_wave.cu: buffers of the FULL-lighting tree run as a wavefront (level 0 = the rays that leave the primary hit, level 1 =
// the rays that leave the first-bounce hit).  An ENTRY is a pixel that reached the level; its two rays (sun, bounce) are rays
// 2e and 2e + 1 of the level.
struct WaveBufs {
    uint32_t* counts;                    // [level] entries pushed, [2 + level] rays queued for the march kernel, [4 + level] next ray it hands out
    float4* ent_o[2];                    // xyz = offset hit position the two rays start from, w = px | py << 16
    uint32_t* ent_m[2];                  // face code of the hit | noise.x << 8 | noise.y << 16 | (level 1: sun1 saw the sky) << 24
    uint32_t* ent_w;                     // level 1: voxel word of the first-bounce hit
    // march queue of a level, one slot per ray that outlived its first steps (trace_wave_dev.cuh wave_start_ray):
    float4* rq_p[2];                     // position, w = ray number (2e: shadow ray of entry e, 2e + 1: its bounce ray)
    float4* rq_d[2];                     // direction as trace_ray marches it (normalised again, :83), w = occupancy texel at the position
    float4* rq_l[2];                     // 1 / |direction| per axis (:88), w = iterations of :113's limit to go
    float4* res_p[2];                    // per ENTRY, bounce ray that hit: offset stop position (:166-180), w = texel offset of the stop voxel
    uint32_t* res_f[2];                  // per RAY: ended in the sky (bit 0) | face code << 1 | (E4) start in a solid / NaN << 4
    uint32_t cap;                        // entries per level the buffers hold (pixels of the band)
};

// Per-frame constants handed to the kernels by value (fits the 4 KB parameter space).
struct TraceParams {
    // frame block (structs.rs:3-31), decoded
    float ox, oy, oz, fx, fy, fz, ux, uy, uz, rx, ry, rz;
    float lrx, lry, lrz;                 // window centre as float (raytrace.comp:104)
    float sun_x, sun_y, sun_z;           // raytrace.comp:317, evaluated once on the host
    float sl_r, sl_g, sl_b;              // sun_color(), raytrace.comp:318
    // sample_sky terms that depend only on the sun (:272-274, :279): mix(40,10,sl)/2, mix(5,1,sl), max(sl,0.1)
    float sky_horizon_exp_half, sky_halo_exp, sky_bright_scale;
    float noise_base_x, noise_base_y;    // raytrace.comp:302-303
    // geometry
    int width, height, row0, row1;       // rows [row0,row1) are traced
    int own_row0, own_row1;              // rows whose rays enter the frame counters (the band without its halo)
    int nx, ny, nz;                      // window dims (powers of two)
    int lx, ly, lz;                      // log2 of the dims
    float hx, hy, hz;                    // half dims (pos_offset, raytrace.comp:105)
    int max_steps, mode;
    int fast;                            // 1: k_trace_fast's preconditions hold for this frame (trace_fast.cu (E5))
    float sky_lo[3], sky_hi[3];          // (E8): |pos - lr| >= H  <=>  pos <= sky_lo || pos >= sky_hi, per axis
    uint32_t flags;
    float inv_frames;                    // accumulation: 1/(frames so far incl. this one) is NOT used; see k_trace
    // device pointers
    const uint8_t* minefield;            // linear [nz][ny][nx]
    const uint32_t* materials;           // linear [nz][ny][nx]
    const uint8_t* noise;                // 512*512*4
    const float* lut;                    // 4*256: sin1, cos1, sin2, cos2 (raytrace.comp:190-197), then the VXRT_LUT_* tables
    ushort4* lighting; uint16_t* depth; uint8_t* normal; uchar4* albedo; uchar4* emission; uchar4* fog;
    int32_t* hit_voxel; uint8_t* hit_face; uint32_t* hit_word; float* hit_t; uint32_t* steps; uint8_t* rays;
    float4* accum;
    float4* hit_pos;                     // VXRT_FLAG_REPROJECT: primary.position (offset, :166-180) + face bits; {0,0,0,16} on air
    unsigned long long* counters;        // [0]=rays [1]=steps [2]=hits [3]=noise fetches [4]=limit exhausted
    unsigned long long one2;             // {1.0f, 1.0f}: a multiplicand ptxas cannot see through (trace_fast.cu (E14))
    // occupancy field as a 3-D texture over a block-linear cudaArray (layout A/B): 0 = linear array + LDG,
    // 1 = integer texel coordinates, 2 = normalised coordinates with wrap addressing (the unit does the floor and the mod),
    // 3 = plain memory in 8x8x8 bricks (field_tex then holds the device address of the bricked copy), 4 = plain memory at 4 bits
    // per voxel, x-neighbours paired (field_tex = the address of that copy)
    int field; unsigned long long field_tex; float inx, iny, inz;
    void* pool_ctx;                      // 16 bytes per pixel for k_trace_pool (trace_pool.cu), or NULL
    void* pool_rings;                    // its per-warp rings (vxrt_trace_pool_ring_bytes)
    WaveBufs wave;                       // counts == NULL: no wavefront buffers on this handle
    int trace_form;                      // VXRT_TRACE_* (vxrt_set_trace_form)
};

// temporal.cu: old_origin / old_transform_c0..2 of the frame block (pipeline.rs:213-227) and the images of the pass
struct ReprojectParams {
    float oox, ooy, ooz, c0x, c0y, c0z, c1x, c1y, c1z, c2x, c2y, c2z;
    int width, height;
    ushort4* lighting; ushort4* completed;
    const float4* hit_pos_cur; const float4* hit_pos_prev; const float4* hist_prev; float4* hist_cur;
};

// one halo link of the peer-memory frame (peer.cu): rows [r0,r1) go to / come from `rank`
struct VxrtPeerLink { int rank, r0, r1; void *lighting, *depth, *normal; };

// cross-GPU signals carried by k_finalize in peer mode (peer.cu): wait word checked by every CTA before it stores
// into the root's image, flag words posted by the last CTA after its stores are fenced
struct VxrtPeerTail { const uint32_t* wait; uint32_t* post[VXRT_PEER_MAX_LINKS + 1]; int n_post; uint32_t value; uint32_t* ticket; uint32_t* err; };

struct VxrtHandle {
    VxrtConfig cfg;
    int row0, row1;                      // rendered G-buffer rows
    int trace_row0, trace_row1;          // rows traced (band + denoise halo)
    int32_t plan[14];                    // vxrt_plan_passes: traced rows, then the rows of each denoise pass
    cudaStream_t stream;
    cudaEvent_t ev[5];
    // pipelined read-back: the final image is double-buffered so frame k's D2H copy (copy_stream) overlaps frame k+1
    cudaStream_t copy_stream; cudaEvent_t ev_frame_done[2], ev_copy_done[2]; bool copy_pending[2];
    uchar4* d_out_buf[2]; int out_idx; unsigned long long tickets, ticket_of_buf[2];
    std::string err;
    // world.  A handle made by vxrt_create_shared borrows d_materials / d_minefield / d_noise from world_owner (same
    // device) and never writes them; the owner orders its world writes against the borrowers' frames (api.cu world_write_*)
    VxrtHandle* world_owner = nullptr; std::vector<VxrtHandle*> sharers;
    cudaEvent_t ev_world = nullptr; unsigned long long world_version = 0, world_seen = 0;
    // VXRT_FLAG_ACCUMULATE on a borrower: the running sum is the owner's (accum_borrowed), so frames dealt alternately
    // to owner and borrowers form ONE accumulation chain; a trace pass waits for the previous writer's (ev_accum, kept
    // on the owner) and the passes after it overlap the next frame's trace pass
    bool accum_borrowed = false; cudaEvent_t ev_accum = nullptr; VxrtHandle* accum_last = nullptr;
    uint32_t* d_materials; uint8_t* d_minefield; uint8_t* d_noise; float* d_lut;
    uint8_t* d_minefield_base = nullptr; // the allocation d_minefield lies in (api.cu: the field never crosses a 4 GiB line)
    // layout A/B (VXRT_FIELD=tex|texn at create): a copy of the occupancy field in a block-linear cudaArray, refreshed after every
    // world write, read through texture objects [0] = integer coordinates, [1] = normalised + wrap
    int field_mode = 0; cudaArray_t field_arr = nullptr; cudaTextureObject_t field_tex[2] = {0, 0};
    uint8_t* d_field_nib = nullptr;      // VXRT_FIELD=nib: the same at 4 bits per voxel (half the bytes)
    uint8_t* d_field_brick = nullptr;    // VXRT_FIELD=brick: the same copy in plain memory as 8x8x8-voxel bricks of 512 contiguous bytes
    std::vector<uint8_t> h_noise;        // host copy: the seed texel is read on the host (raytrace.comp:298-303)
    uint32_t* d_dirty;                   // one flag per 64^3 chunk
    void* d_edits; size_t edits_cap;
    void* d_edit_table = nullptr; size_t edit_table_slots = 0;     // last-wins resolution of duplicate voxels in an edit list (world.cu)
    // vxrt_apply_edits stages the caller's list through two page-locked buffers, so it returns without waiting for the
    // frames still queued on the stream (ev_edits[i]: the copy out of buffer i has been issued and completed)
    void* h_edits[2] = {nullptr, nullptr}; size_t h_edits_cap = 0; cudaEvent_t ev_edits[2] = {nullptr, nullptr}; int edit_slot = 0;
    // frame images
    ushort4 *d_lighting, *d_pong, *d_lighting_raw;
    uint16_t* d_depth; uint8_t* d_normal; uchar4 *d_albedo, *d_emission, *d_fog, *d_output;
    int32_t* d_hit_voxel; uint8_t* d_hit_face; uint32_t* d_hit_word; float* d_hit_t; uint32_t* d_steps; uint8_t* d_rays;
    float4* d_accum;
    void* d_pool_ctx = nullptr;          // k_trace_pool's per-pixel records (FULL lighting without hit buffers)
    int trace_form = 0;                  // VXRT_TRACE_* preference of this handle
    float steps_per_ray = 0.0f;          // of the last frame whose counters the caller read (vxrt_frame_stats); 0 = unknown
    bool wave_frame = false;             // the last frame's trace pass ran as a wavefront
    void* d_wave = nullptr; WaveBufs wave = {};   // trace_wave.cu: entry / ray / result arrays of the wavefront form (FULL lighting)
    void* d_pool_rings = nullptr;        // and its per-warp rings
    // temporal reprojection (temporal.cu): hit records and {mean light, count} of this frame and the one before
    float4* d_hit_pos[2] = {nullptr, nullptr}; float4* d_hist[2] = {nullptr, nullptr}; ushort4* d_completed = nullptr; int hist_idx = 0;
    unsigned long long* d_counters;
    VxrtFrameParams frame; bool have_frame, have_world, have_noise, rendered;
    uint32_t frames_accumulated;
    // multi-GPU (comm.cu)
    void* comm; int rank, nranks; std::vector<int32_t> bands;   // bands[2r], bands[2r+1] = rows of rank r
    // multi-GPU over peer memory (peer.cu)
    bool peer_on = false; int peer_root = 0; uint32_t peer_frame = 0, peer_barriers = 0;
    uint32_t* d_peer_flags = nullptr;                            // this handle's flag words (exported)
    std::vector<uint32_t*> peer_flags;                           // every rank's flag words as seen from this device
    std::vector<VxrtPeerLink> peer_dst, peer_src;
    void* peer_root_out[2] = {nullptr, nullptr};                 // the root's two output images
    std::vector<void*> peer_ipc_opened;
    cudaStream_t peer_stream = nullptr; cudaEvent_t ev_peer = nullptr;
    // the frame as a CUDA graph, one per output image (api.cu): everything after the host-side waits, replayed with the trace
    // node's parameters set per frame; re-captured when the trace kernel or its grid changes
    struct FrameGraph { cudaGraph_t graph = nullptr; cudaGraphExec_t exec = nullptr; cudaGraphNode_t trace_node = nullptr; const void* func = nullptr; dim3 grid, block; size_t smem = 0; };
    FrameGraph graphs[2]; bool graph_failed = false;
};
// error text of calls that have no handle (read back with vxrt_last_error(NULL))
void vxrt_set_global_error(const char* msg);
void vxrt_comm_release(VxrtHandle* h);
void vxrt_peer_release(VxrtHandle* h);
int vxrt_peer_begin_frame(VxrtHandle* h, int out_idx);
int vxrt_peer_boundary_rows(VxrtHandle* h, int32_t* iv);
int vxrt_peer_push_halo(VxrtHandle* h);
int vxrt_peer_wait_halo(VxrtHandle* h);
int vxrt_peer_after_denoise(VxrtHandle* h);
VxrtPeerTail vxrt_peer_finalize_tail(VxrtHandle* h);
uchar4* vxrt_peer_output(VxrtHandle* h, int out_idx);
int vxrt_peer_end_frame(VxrtHandle* h);
int vxrt_peer_error_word(VxrtHandle* h, uint32_t* out);
void vxrt_peer_clear_error(VxrtHandle* h);

// module preload (cudaFuncGetAttributes on every kernel of a file)
cudaError_t vxrt_preload_trace(); cudaError_t vxrt_preload_trace_fast(); cudaError_t vxrt_preload_post();
cudaError_t vxrt_preload_world(); cudaError_t vxrt_preload_peer(); cudaError_t vxrt_preload_temporal();

// launchers (each returns the cudaGetLastError() of its launch)
cudaError_t vxrt_launch_trace(const TraceParams& p, cudaStream_t s);        // dispatches on p.fast
cudaError_t vxrt_launch_trace_fast(const TraceParams& p, cudaStream_t s);
// the kernel (and launch shape) vxrt_launch_trace runs for these parameters: the kernel node of a captured frame (api.cu)
const void* vxrt_trace_kernel(const TraceParams& p, dim3* grid, dim3* block, size_t* smem);
const void* vxrt_trace_pool_kernel(const TraceParams& p, dim3* grid, dim3* block, size_t* smem);   // trace_pool.cu; NULL = not applicable
cudaError_t vxrt_preload_trace_pool();
size_t vxrt_trace_pool_ring_bytes(int sms);
const void* vxrt_trace_fast_kernel(const TraceParams& p, dim3* grid, dim3* block);
const void* vxrt_trace_fast_wave_kernel(const TraceParams& p, dim3* grid, dim3* block);   // the primary stage of the wavefront form, or NULL
// trace_wave.cu: the FULL-lighting trace pass as a wavefront (primary stage, march, resolve, march, resolve)
bool vxrt_trace_wave_applies(const TraceParams& p);
cudaError_t vxrt_launch_trace_wave(const TraceParams& p, cudaStream_t s);
cudaError_t vxrt_preload_trace_wave();
size_t vxrt_trace_wave_bytes(size_t entries);
void vxrt_trace_wave_carve(WaveBufs* w, void* base, size_t entries);
// finalize.comp folded into the last denoise pass (post.cu k_denoise<.., FIN>): the images finalize reads besides lighting/depth
struct VxrtFinalizeFold { const uchar4* albedo; const uchar4* emission; const uchar4* fog; const uint8_t* noise; uchar4* out; int row0, row1; };
cudaError_t vxrt_launch_denoise(const ushort4* src, ushort4* dst, const uint16_t* depth, const uint8_t* normal,
                                int width, int height, int size, int swap, int row0, int row1, cudaStream_t s, bool pdl,
                                const VxrtFinalizeFold* fold = nullptr);
cudaError_t vxrt_launch_finalize(const uchar4* albedo, const uchar4* emission, const uchar4* fog, const ushort4* lighting,
                                 const uint16_t* depth, const uint8_t* noise, uchar4* out, int width, int height,
                                 int row0, int row1, cudaStream_t s, const VxrtPeerTail* tail = nullptr);
cudaError_t vxrt_launch_reproject(const ReprojectParams& p, cudaStream_t s);
cudaError_t vxrt_launch_row_cost(const uint32_t* steps, const uint8_t* rays, const uint8_t* normal, int width, int height,
                                 float* out /* 3 per row: steps, rays, filtered pixels */, cudaStream_t s);
cudaError_t vxrt_launch_build_minefield(uint32_t* materials, uint8_t* minefield, int nx, int ny, int nz,
                                        const uint32_t* dirty /*NULL = all chunks*/, cudaStream_t s);
cudaError_t vxrt_launch_apply_edits(uint32_t* materials, uint32_t* dirty, const VxrtEdit* d_edits, size_t n,
                                    int nx, int ny, int nz, void* table /* 12 bytes per slot */, size_t table_slots /* pow2 >= 4n */, cudaStream_t s);
cudaError_t vxrt_launch_compare_chunk(const uint8_t* minefield, const uint8_t* d_stored, int nx, int ny, int cx, int cy, int cz,
                                      unsigned long long* d_out /* [0] mismatching bytes, [1] first offset */, cudaStream_t s);
cudaError_t vxrt_launch_field_to_nibbles(const uint8_t* minefield, uint8_t* nibbles, int nx, int ny, int nz, const int* box, cudaStream_t s);
cudaError_t vxrt_launch_field_to_bricks(const uint8_t* minefield, uint8_t* bricks, int nx, int ny, int nz, const int* box, cudaStream_t s);
cudaError_t vxrt_launch_generate(uint32_t* materials, int nx, int ny, int nz, int kind, uint32_t seed, cudaStream_t s);
