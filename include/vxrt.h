/*
 * vxrt.h — C ABI of the B200-native voxel ray-tracing hot path.
 *
 * This is the drop-in boundary for the reference's `src/render` module
 * (someguynamedjosh/raytrace).  Each entry point names the reference interface
 * it replaces (file:line under /root/reference).  Plain pointers and sizes
 * only; the caller owns every host pointer, the library owns all device memory.
 * Every call returns VXRT_OK (0) or a negative VxrtStatus; vxrt_last_error()
 * gives the text.  A handle is bound to one GPU and one CUDA stream and is not
 * thread-safe (the reference is single-threaded: Rc<Core>, render/mod.rs:1,39).
 *
 * There is NO CPU fallback: without a CUDA device vxrt_create fails with
 * VXRT_ERR_CUDA.
 */
#ifndef VXRT_H
#define VXRT_H
#include <stddef.h>
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

#define VXRT_ABI_VERSION 1

typedef enum VxrtStatus {
    VXRT_OK = 0,
    VXRT_ERR_INVALID = -1,   /* bad argument / configuration             */
    VXRT_ERR_CUDA = -2,      /* CUDA runtime error (text in last_error)   */
    VXRT_ERR_STATE = -3,     /* call made in the wrong order              */
    VXRT_ERR_SIZE = -4,      /* host buffer too small                     */
    VXRT_ERR_COMM = -5       /* NCCL unavailable or an NCCL call failed   */
} VxrtStatus;

/* Truncations of raytrace.comp:main (:320-350).  Only FULL is reference behaviour. */
typedef enum VxrtMode { VXRT_MODE_PRIMARY = 0, VXRT_MODE_SHADOW = 1, VXRT_MODE_FULL = 2 } VxrtMode;

/* vxrt_create flags */
#define VXRT_FLAG_HIT_BUFFERS   0x1u  /* keep per-pixel hit voxel/face/word/t/steps/rays (parity + stats) */
#define VXRT_FLAG_DENOISE       0x2u  /* run the 6 bilateral passes (pipeline.rs:98-115)                    */
#define VXRT_FLAG_FINALIZE      0x4u  /* run finalize.comp -> RGBA8 output (pipeline.rs:117-123)            */
#define VXRT_FLAG_ACCUMULATE    0x8u  /* temporal accumulation of `light` (extension)          */
#define VXRT_FLAG_PONG_SWAP     0x10u /* reproduce descriptor_sets.rs:37-39 (odd passes read swapped images)*/
#define VXRT_FLAG_GENERAL_MARCH 0x40u /* force the general march kernel (normally chosen per frame when a coordinate
                                         could reach 2^21 or the window has more than 2^32 voxels)               */
#define VXRT_FLAG_REPROJECT     0x80u /* temporal reprojection under camera motion (extension): reads the
                                         frame block's old_origin / old_transform_c0..2, which the reference fills every
                                         frame (pipeline.rs:213-227) and its shaders never read; see vxrt_reset_history */
/* 0x20 is unassigned (a reference-rows mode was declared here and never built: every pixel row is rendered, deviation D1) */
#define VXRT_FLAG_RAY_HANDOFF   0x100u /* FULL lighting: a pixel entering the second bounce level gives its diffuse ray to a
                                          finished lane of its warp, so both rays of the level march in one round
                                          (trace_fast.cu (E16)).  Same rays, same arithmetic, same images and frame
                                          counters; ignored with VXRT_FLAG_HIT_BUFFERS (per-pixel counters) and on row
                                          bands.  Off by default: measured 0.94x .. 1.06x depending on the scene */

/* Replaces the compile-time constants of src/render/constants.rs:9-10,22-31 and the
 * shader constants raytrace.comp:37-60, which become run-time parameters. */
typedef struct VxrtConfig {
    uint32_t abi_version;        /* VXRT_ABI_VERSION                                        */
    int32_t  width, height;      /* frame size (constants.rs:9-10: 1024x1024)               */
    int32_t  nx, ny, nz;         /* voxel window, powers of two >= 64 (ROOT_BLOCK_SIZE 256) */
    int32_t  mode;               /* VxrtMode                                                */
    uint32_t flags;              /* VXRT_FLAG_*                                             */
    int32_t  max_steps;          /* raytrace.comp:109, 0 -> 2048                            */
    int32_t  device;             /* CUDA device ordinal                                     */
    int32_t  row_begin, row_end; /* G-buffer rows this handle renders (image-space shard);
                                    0,0 -> whole frame.  Output rows are height-1-row.      */
    int32_t  reserved[4];
} VxrtConfig;

/* Byte-compatible with `RaytraceUniformData` (#[repr(C)], 192 B,
 * src/render/pipeline/structs.rs:3-31) so the Rust side can pass &uniform_data. */
typedef struct VxrtFrameParams {
    float    sun_angle;              /*   0 */
    uint32_t seed;                   /*   4 */
    uint64_t _padding0;              /*   8 */
    float    origin[3];  uint32_t _padding1;   /*  16 */
    float    forward[3]; uint32_t _padding2;   /*  32 */
    float    up[3];      uint32_t _padding3;   /*  48  (pre-scaled by 0.4, pipeline.rs:198) */
    float    right[3];   uint32_t _padding4;   /*  64  (pre-scaled by 0.4, pipeline.rs:199) */
    float    old_origin[3];       uint32_t _padding5;  /*  80  unread by the reference shaders; read by VXRT_FLAG_REPROJECT */
    float    old_transform_c0[3]; uint32_t _padding6;  /*  96  columns of inverse([right*0.4 | up*0.4 | forward]) of the */
    float    old_transform_c1[3]; uint32_t _padding7;  /* 112  previous frame (pipeline.rs:215-226); unread by the        */
    float    old_transform_c2[3]; uint32_t _padding8;  /* 128  reference shaders; read by VXRT_FLAG_REPROJECT            */
    int32_t  region_offset[3];    uint32_t _padding9;  /* 144  unread */
    int32_t  rotation[3];         uint32_t _padding10; /* 160  shader name: lr (window centre) */
    int32_t  space_offset[3];     uint32_t _padding11; /* 176  unread */
} VxrtFrameParams;

/* One voxel edit in WORLD coordinates (window texel = (world + N/2) mod N). */
typedef struct VxrtEdit { int32_t x, y, z; uint32_t word; } VxrtEdit;

typedef enum VxrtBufferId {
    VXRT_BUF_OUTPUT = 0,     /* RGBA8   w*h*4   final image, y flipped (finalize.comp:60-63)   */
    VXRT_BUF_LIGHTING = 1,   /* RGBA16  w*h*8   lighting_buffer after the last pass run        */
    VXRT_BUF_DEPTH = 2,      /* R16UI   w*h*2                                                  */
    VXRT_BUF_NORMAL = 3,     /* R8UI    w*h                                                    */
    VXRT_BUF_ALBEDO = 4,     /* RGBA8   w*h*4                                                  */
    VXRT_BUF_EMISSION = 5,   /* RGBA8   w*h*4                                                  */
    VXRT_BUF_FOG = 6,        /* RGBA8   w*h*4                                                  */
    VXRT_BUF_HIT_VOXEL = 7,  /* int32x3 w*h*12  floor(position) at loop exit (HIT_BUFFERS)     */
    VXRT_BUF_HIT_FACE = 8,   /* u8      w*h     0..5, 16 = air                                 */
    VXRT_BUF_HIT_WORD = 9,   /* u32     w*h*4   packed material word, 0 on air                 */
    VXRT_BUF_HIT_T = 10,     /* f32     w*h*4   primary hit distance                           */
    VXRT_BUF_STEPS = 11,     /* u32     w*h*4   march iterations over all rays of the pixel    */
    VXRT_BUF_RAYS = 12,      /* u8      w*h     trace_ray calls of the pixel                   */
    VXRT_BUF_ACCUM = 13,     /* f32x4   w*h*16  running sum of light + frame count             */
    VXRT_BUF_LIGHTING_RAW = 14, /* RGBA16 w*h*8 lighting as written by the trace pass          */
    VXRT_BUF_HIT_POS = 15,   /* f32x4   w*h*16  primary.position after the :166-180 offset, w = bits of the int face
                                                code; {0,0,0,16} on air (REPROJECT)                               */
    VXRT_BUF_HISTORY = 16,   /* f32x4   w*h*16  mean light over the reprojected samples + sample count (REPROJECT) */
    VXRT_BUF_COMPLETED = 17, /* RGBA16  w*h*8   completed_buffer (raytrace.comp:19): mean/16, a = count/MAX_SAMPLES */
    VXRT_BUF_COUNT_
} VxrtBufferId;

/* Counters of the last rendered frame over the handle's own rows [row_begin,row_end); rays traced for
 * the denoise halo of a band are not counted. */
typedef struct VxrtFrameStats {
    uint64_t rays, steps, hits, noise_fetches;
    uint64_t limit_exhausted;    /* rays that ran out of max_steps (reference UB) */
    float    trace_ms, denoise_ms, finalize_ms, total_ms;   /* CUDA-event times of the last frame */
} VxrtFrameStats;

typedef struct VxrtHandle VxrtHandle;

/* ---- construction / teardown -------------------------------------------------------- */
/* replaces render::create_instance (src/render/mod.rs:36-43) + Pipeline::new (pipeline.rs:36-76)
 * + RenderData::create (render_data.rs:165-201), minus window/swapchain. */
int  vxrt_create(const VxrtConfig* cfg, VxrtHandle** out);
/* A second handle on the owner's GPU that BORROWS the owner's world (voxel words, occupancy field, blue noise: no
 * second copy in HBM) and has its own stream and frame images.  No reference equivalent: the reference keeps one
 * frame in flight (pipeline.rs:162-172).  Frames that do not depend on each other (a camera path) can be submitted
 * alternately to the owner and a borrower, so that the tail of one frame's passes overlaps the next frame's trace
 * pass on the same GPU (measured: +23 % frames/s on BASELINE config 2 with two handles).  cfg->device and the window
 * dims must equal the owner's.  World writes (uploads, generate, edits, blue noise) go through the owner only — on a
 * borrower they fail with VXRT_ERR_STATE; the owner waits for the borrowers' frames in flight before it writes, and a
 * borrower's next frame waits for the write.  Destroy borrowers before their owner (a borrower that outlives it has
 * no world and fails with VXRT_ERR_STATE).
 * VXRT_FLAG_ACCUMULATE on a borrower (the owner must have it, same image and rows): the running sum is the owner's, so
 * frames submitted alternately to owner and borrowers form ONE accumulation chain, exactly the frames of a single
 * handle - each trace pass is ordered after the previous frame's, the denoise/finalize passes of frame k overlap the
 * trace pass of frame k+1; vxrt_reset_accumulation on any handle of the group resets that sum. */
int  vxrt_create_shared(const VxrtConfig* cfg, VxrtHandle* owner, VxrtHandle** out);
/* replaces Drop for Pipeline/Core (pipeline.rs:258-277, core.rs:88-112) */
void vxrt_destroy(VxrtHandle* h);
const char* vxrt_last_error(const VxrtHandle* h);   /* h may be NULL: error of the last failed create */
int  vxrt_abi_version(void);

/* ---- world upload ------------------------------------------------------------------- */
/* One 64^3 chunk at window chunk coords (cx,cy,cz) in [0, n/64): replaces the per-chunk
 * copy_materials/copy_minefield of make_world_upload_buffers (render_data.rs:221-241).
 * materials: 64^3 u32, x fastest (PackedChunkData, chunk.rs:52-56).
 * minefield: 64^3 u8 or NULL -> built on the GPU with the pack_into rule (chunk.rs:125-184). */
int  vxrt_upload_chunk(VxrtHandle* h, int cx, int cy, int cz, const uint32_t* materials, const uint8_t* minefield);
/* Whole window at once ([nz][ny][nx], x fastest); minefield NULL -> built on the GPU. */
int  vxrt_upload_world(VxrtHandle* h, const uint32_t* materials, const uint8_t* minefield);
/* A 16-thick slab along `axis` at window texel offset `offset` (multiple of 16): replaces
 * TerrainUploadManager::upload_slice's copy_buffer_to_image_offset (terrain_upload.rs:212-267).
 * Slab layout: x fastest, extents = window dims with `axis` replaced by 16. */
int  vxrt_upload_slice(VxrtHandle* h, int axis, int offset, const uint32_t* materials, const uint8_t* minefield);
/* Seeded synthetic worlds generated on the device (kind 0 = heightmap, 1 = biome). */
int  vxrt_generate_world(VxrtHandle* h, int kind, uint32_t seed);
/* Rebuild every chunk's minefield from the material words (pack_into, chunk.rs:125-184). */
int  vxrt_rebuild_minefield(VxrtHandle* h);
int  vxrt_read_world(VxrtHandle* h, uint32_t* materials, uint8_t* minefield);   /* either may be NULL */
/* Stored-chunk check (chunk_storage.rs:56-68: a chunk file carries materials || minefield): compares the 64^3 occupancy bytes
 * of a file (vxrt_chunk_decode) with the ones the device holds for window chunk (cx,cy,cz) - after vxrt_upload_chunk(materials,
 * NULL) those are the ones k_build_minefield derived from the file's words with the pack_into rule (chunk.rs:125-184).
 * *mismatches = differing bytes; *first_offset (may be NULL) = offset of the first one in the chunk, -1 if none. */
int  vxrt_validate_chunk(VxrtHandle* h, int cx, int cy, int cz, const uint8_t* stored_minefield, uint64_t* mismatches, int64_t* first_offset);
/* replaces create_blue_noise (render_data.rs:110-133): 512*512 RGBA8 */
int  vxrt_upload_blue_noise(VxrtHandle* h, const uint8_t* rgba, size_t nbytes);

/* ---- edits (no reference equivalent beyond set_block + pack_into, chunk.rs:115-184) -- */
/* Writes the words, then rebuilds the minefield of every touched 64^3 chunk.  A voxel named more than once takes the word of
 * its LAST entry in the list (set_block applied in list order, chunk.rs:115-117) on every replica.  Asynchronous on the
 * handle's stream; the list is copied (into page-locked staging) before the call returns, so the caller may reuse it at once. */
int  vxrt_apply_edits(VxrtHandle* h, const VxrtEdit* edits, size_t n);
/* Same with the edit list already on the device (e.g. after an NCCL broadcast). */
int  vxrt_apply_edits_device(VxrtHandle* h, const void* d_edits, size_t n);

/* No counterpart in the reference.  The edit path for callers that may not allocate
 * (code under CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc): scatters the n edits of a DEVICE list into
 * the caller's own window arrays (d_materials u32 / d_minefield u8, [nz][ny][nx], x fastest) with vxrt_apply_edits' last-entry-wins
 * rule and rebuilds the occupancy field of every touched 64^3 chunk (pack_into rule), asynchronously on `cuda_stream` (a
 * cudaStream_t; NULL = the default stream).  It calls no cudaMalloc / cudaFree: its scratch (last-wins table + per-chunk dirty
 * flags, 896 KiB) is one file-scope device array that every GPU that loads the library holds until exit.  Limits: n <=
 * VXRT_STATIC_MAX_EDITS, window chunks (nx/64 * ny/64 * nz/64) <= VXRT_STATIC_MAX_CHUNKS; beyond them VXRT_ERR_SIZE (text from
 * vxrt_last_error(NULL)).  Every call on a device shares the array, which is cleared at the start of each call: calls that overlap
 * on the GPU (different streams without ordering) race. */
#define VXRT_STATIC_MAX_EDITS  16384
#define VXRT_STATIC_MAX_CHUNKS 32768
int  vxrt_apply_edits_static(int device, uint32_t* d_materials, uint8_t* d_minefield, int nx, int ny, int nz,
                             const VxrtEdit* d_edits, size_t n, void* cuda_stream);

/* ---- per frame ---------------------------------------------------------------------- */
/* replaces the uniform fill + mapped-buffer memcpy of draw_frame (pipeline.rs:195-211) */
int  vxrt_set_frame(VxrtHandle* h, const VxrtFrameParams* p);
/* replaces the submit of the pre-recorded command buffer (pipeline.rs:78-132,229-235):
 * trace -> [6x denoise] -> [finalize], asynchronous on the handle's stream. */
int  vxrt_render_frame(VxrtHandle* h);
/* replaces wait_for_fences (pipeline.rs:162-172) */
int  vxrt_sync(VxrtHandle* h);
int  vxrt_reset_accumulation(VxrtHandle* h);
/* Temporal reprojection (VXRT_FLAG_REPROJECT; extension — the reference only prepares its inputs).  Between the trace
 * pass and the denoise passes, per pixel:
 *   L = lighting texel * 16;  n = 0, Hm = 0
 *   hit: r = primary.position - old_origin; v = c0*r.x + c1*r.y + c2*r.z; if v.z > 0: f = ((v.xy/v.z + 1) * 0.5) * (W,H),
 *        q = floor(f + 0.5); if q is inside the frame and the previous frame's hit at q has the same face code and the
 *        same floor(position): n = min(count_prev(q), MAX_SAMPLES - 1), Hm = mean_prev(q)        (MAX_SAMPLES = 8)
 *   mean = (Hm*n + L) / (n + 1);  history = {mean, n + 1};  lighting = mean / 16;  completed = {mean / 16, (n+1)/8}
 * so a pixel that sees the same voxel face as the previous frame keeps up to 8 frames of light, and a disoccluded
 * one starts again at 1.  The whole frame must live on one handle (no row bands, no peer mode) and the flag excludes
 * VXRT_FLAG_ACCUMULATE.  vxrt_reset_history forgets the previous frame (camera cut, world edit). */
int  vxrt_reset_history(VxrtHandle* h);

/* ---- output ------------------------------------------------------------------------- */
/* replaces queue_present (pipeline.rs:237-254): the frame is a headless device buffer. */
int  vxrt_read_output(VxrtHandle* h, uint8_t* rgba8, size_t nbytes);
/* Pipelined read-back (no reference equivalent: the reference presents from the GPU).  The final image is
 * double-buffered on the device: vxrt_read_output_async enqueues the D2H copy of the frame rendered last on a
 * second stream and returns a ticket at once, so the next vxrt_render_frame overlaps the copy;
 * vxrt_wait_output(ticket) blocks until that copy (and every earlier one) has landed in `rgba8`.
 * The copy only runs at DMA speed when `rgba8` is page-locked: register the caller's buffer once with
 * vxrt_host_register (cudaHostRegister underneath). */
int  vxrt_read_output_async(VxrtHandle* h, uint8_t* rgba8, size_t nbytes, uint64_t* ticket);
int  vxrt_wait_output(VxrtHandle* h, uint64_t ticket);
int  vxrt_host_register(VxrtHandle* h, void* ptr, size_t nbytes);
int  vxrt_host_unregister(VxrtHandle* h, void* ptr);
int  vxrt_read_buffer(VxrtHandle* h, int buffer_id, void* dst, size_t nbytes);
size_t vxrt_buffer_bytes(const VxrtHandle* h, int buffer_id);
void* vxrt_device_buffer(VxrtHandle* h, int buffer_id);     /* zero-copy consumers (NCCL gather) */
void* vxrt_cuda_stream(VxrtHandle* h);                      /* cudaStream_t of the handle         */
int  vxrt_frame_stats(VxrtHandle* h, VxrtFrameStats* out);  /* syncs the stream                   */
/* Kernels one vxrt_render_frame launches on this handle outside peer mode: the trace pass, the six denoise passes, k_finalize
 * (VXRT_FOLD=1 in the environment folds it into the last denoise launch), k_reproject with VXRT_FLAG_REPROJECT. */
int  vxrt_launches_per_frame(const VxrtHandle* h);
/* Form of the trace pass on this handle.  All forms march the same rays with the same fp32 operations (images, running sum and
 * frame counters are equal bit for bit); they differ in how rays are dealt to lanes.  The reference has one form, the compute
 * dispatch of raytrace.comp (pipeline.rs:88-96).  AUTO is the sequential kernel; under FULL lighting it hands the second bounce
 * ray of a pixel to a finished lane of its warp once the caller's vxrt_frame_stats calls have shown more than 28 march steps per
 * ray (VXRT_NO_AUTO_HANDOFF=1 turns that off; VXRT_WAVE=1 / VXRT_POOL=1 in the environment make AUTO pick those forms). */
#define VXRT_TRACE_AUTO       0
#define VXRT_TRACE_SEQUENTIAL 1   /* k_trace_fast / k_trace_general: a warp keeps its 32 pixels through their rays            */
#define VXRT_TRACE_POOL       2   /* k_trace_pool: per-warp ray rings inside one kernel (FULL lighting; measured slower)       */
#define VXRT_TRACE_WAVEFRONT  3   /* trace_wave.cu: primary stage, then per level a ray-queue march and a resolve pass (FULL) */
int  vxrt_set_trace_form(VxrtHandle* h, int form);

/* ---- image-space sharding plan: pure host arithmetic, usable without a device ------------------------ */
/* rows[2r], rows[2r+1] = band [begin,end) of rank r: contiguous, edges on multiples of 8 rows
 * (the 8x8 shader group, constants.rs:33). */
int  vxrt_plan_bands(int height, int nranks, int32_t* rows /* 2*nranks */);
/* Load-balanced variant: contiguous bands (edges on multiples of 8 rows) that equalise the modelled cost of a band,
 * its denoise halo included.  trace_cost[y] / denoise_cost[y] (per pass) are per-row costs in march-step units as
 * vxrt_row_costs returns them for a representative frame (sky rows are cheap, ground rows are not: equal-height
 * bands leave the rank that owns the ground with most of the work). */
int  vxrt_plan_bands_weighted(int height, int nranks, uint32_t flags, const float* trace_cost, const float* denoise_cost,
                              int32_t* rows /* 2*nranks */);
/* Per-row cost of the frame rendered last (needs VXRT_FLAG_HIT_BUFFERS): trace_cost[y] = steps + 7 rays + 16.7 W,
 * denoise_cost[y] = 15 filtered pixels + 0.45 W; rows outside the handle's band read 0. */
int  vxrt_row_costs(VxrtHandle* h, float* trace_cost, float* denoise_cost, size_t n);
/* For a handle rendering [row_begin,row_end): rows[0..1] = rows the trace pass must cover (band + 117-row
 * denoise halo, clipped), rows[2+2k..3+2k] = rows denoise pass k (sizes 1,2,4,8,8,16, pipeline.rs:103)
 * must cover so that the final band only ever read valid rows. */
int  vxrt_plan_passes(int height, int row_begin, int row_end, uint32_t flags, int32_t* rows /* 14 */);

/* ---- multi-GPU (no reference equivalent: the reference drives one Vulkan queue, pipeline.rs:229-235) --
 * One handle per GPU/process; the world is replicated and each handle renders the row band given
 * at vxrt_create.  NCCL (bound at run time) moves only the final image bands and edit lists. */
#define VXRT_COMM_ID_BYTES 128
/* On one rank: an NCCL unique id to hand to every rank by any host channel. */
int  vxrt_comm_unique_id(void* id_out, size_t nbytes);
int  vxrt_comm_init(VxrtHandle* h, const void* id, size_t nbytes, int rank, int nranks);
/* After vxrt_render_frame: every rank's RGBA8 band lands in root's VXRT_BUF_OUTPUT (root < 0: in
 * every rank's).  Asynchronous on the handle's stream. */
int  vxrt_comm_gather_output(VxrtHandle* h, int root);
/* Broadcast root's host edit list (n given on every rank) and apply it on every replica. */
int  vxrt_comm_broadcast_edits(VxrtHandle* h, const VxrtEdit* edits, size_t n, int root);


/* ---- chunk files (host only; no device needed) --------------------------------------------------------
 * The reference's on-disk chunk format (src/world/chunk_storage.rs:37-68): file name = three {:016X} fields of the
 * signed chunk coordinates; contents = LZ4 frame of materials (u32 LE x 64^3) followed by minefield (u8 x 64^3).
 * vxrt_chunk_decode output feeds vxrt_upload_chunk directly. */
int    vxrt_chunk_file_name(int64_t cx, int64_t cy, int64_t cz, char* out /* >= 49 bytes */, size_t nbytes);
size_t vxrt_chunk_encode_bound(void);
int    vxrt_chunk_encode(const uint32_t* materials, const uint8_t* minefield, void* out, size_t cap, size_t* nbytes);
int    vxrt_chunk_decode(const void* file, size_t nbytes, uint32_t* materials, uint8_t* minefield);

/* ---- multi-GPU frame over NVLink peer memory (no NCCL on the per-frame path; peer.cu) ---------------------
 * After vxrt_peer_attach a band traces only its own rows: the rows of lighting/depth/normal a neighbour's
 * denoise needs are stored straight into that neighbour's images, the final RGBA8 band is stored by the
 * finalize kernel straight into the root's output image, and ordering is by frame counters in device memory.
 * Every rank must call vxrt_render_frame the same number of times.  Ranks may be processes (CUDA IPC) or
 * handles of one process.
 *   1. every rank: vxrt_peer_export(h, blob)          (VXRT_PEER_BLOB_BYTES bytes)
 *   2. all-gather the blobs by any host channel (rank order)
 *   3. every rank: vxrt_peer_attach(h, blobs, nranks * VXRT_PEER_BLOB_BYTES, rank, nranks, root)
 * The root's vxrt_read_output / vxrt_read_output_async then return the whole frame. */
#define VXRT_PEER_BLOB_BYTES 512
int  vxrt_peer_export(VxrtHandle* h, void* blob, size_t nbytes);
int  vxrt_peer_attach(VxrtHandle* h, const void* blobs, size_t nbytes, int rank, int nranks, int root);
/* Device-side barrier over all ranks, enqueued on the handle's stream (no host round trip): benchmarks use it to
 * align the ranks before a timed frame. */
int  vxrt_peer_barrier(VxrtHandle* h);

/* ---- self test ---------------------------------------------------------------------------------------
 * The trace kernel writes IEEE divisions that share a divisor as one refined reciprocal plus two FMAs per quotient
 * (trace_fast.cu (E13)).  This runs that sequence and __fdiv_rn on n pseudo-random operand sets inside the guarded
 * range on `device` and returns the number of results that differ in any bit (must be 0).  Test hook: the
 * reference has no counterpart. */
int  vxrt_selftest_division(int device, uint64_t n, uint64_t seed, uint64_t* mismatches);

#ifdef __cplusplus
}
#endif
#endif
