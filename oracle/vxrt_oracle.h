/*
 * vxrt_oracle.h — C interface of the CPU ORACLE.
 *
 * TEST INFRASTRUCTURE ONLY.  This is a plain-CPU restatement of the reference's
 * GLSL compute passes and of `pack_into`, used as the parity checker by tests/,
 * by __graft_entry__.smoke() and by bench.py's cpu_baseline / --impl reference
 * legs.  Nothing under raytrace_b200/ may include, link or call it: the product
 * path is the CUDA library behind include/vxrt.h and it has no CPU fallback.
 *
 * PARITY STATUS: the reference has no test, fixture or golden vector on this
 * path and its GLSL cannot be executed here through any
 * graphics API ("unavailable on this box").  The oracle is pinned instead by
 *   (i)  known answers derived from reference code (tests/test_oracle_known.py),
 *   (ii) oracle/_ref: the reference's own shader text, mechanically wrapped and
 *        compiled as C++ (oracle/make_ref.py), compared output-for-output.
 *
 * All paths cited are relative to /root/reference.
 */
#ifndef VXRT_ORACLE_H
#define VXRT_ORACLE_H
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

enum { ORC_MODE_PRIMARY = 0, ORC_MODE_SHADOW = 1, ORC_MODE_FULL = 2 };

/* Read-only scene: the two 3-D textures + blue noise (render_data.rs:54-133). */
typedef struct OrcScene {
    int32_t nx, ny, nz;          /* window dims, powers of two (reference: 256^3) */
    const uint32_t* materials;   /* [nz][ny][nx] packed voxel words, x fastest  */
    const uint8_t*  minefield;   /* [nz][ny][nx] occupancy-LOD codes 0..6        */
    const uint8_t*  blue_noise;  /* 512*512*4 RGBA8                              */
    int32_t max_steps;           /* reference: 2048 (raytrace.comp:109)          */
} OrcScene;

/* Per-pixel outputs.  Any pointer may be NULL (not written). Images are
 * row-major [height][width]. */
typedef struct OrcBuffers {
    int32_t width, height;
    /* G-buffer, formats as render_data.rs:166-189 */
    uint16_t* lighting;   /* RGBA16_UNORM  w*h*4 */
    uint16_t* depth;      /* R16_UINT      w*h   */
    uint8_t*  normal;     /* R8_UINT       w*h   */
    uint8_t*  albedo;     /* RGBA8_UNORM   w*h*4 */
    uint8_t*  emission;   /* RGBA8_UNORM   w*h*4 */
    uint8_t*  fog;        /* RGBA8_UNORM   w*h*4 */
    /* primary-hit record (integer parity targets) */
    int32_t*  hit_voxel;  /* w*h*3, floor(position) at loop exit, pre-offset */
    uint8_t*  hit_face;   /* w*h, 0..5, 16 = air                             */
    uint32_t* hit_word;   /* w*h, packed material word (0 on air)            */
    float*    hit_t;      /* w*h, distance                                   */
    uint32_t* steps;      /* w*h, loop iterations summed over all rays        */
    uint8_t*  rays;       /* w*h, number of trace_ray calls                   */
    /* temporal accumulation (extension): running SUM of fp32 light */
    float*    accum;      /* w*h*4 (rgb + unused)                             */
    /* temporal reprojection (extension): primary.position (after the :166-180 offset)
     * of a hit, w = bit pattern of the int face code; {0,0,0,16} on air */
    float*    hit_pos;    /* w*h*4 */
} OrcBuffers;

typedef struct OrcStats {
    uint64_t rays, steps, hits, noise_fetches;
    uint64_t ub_limit_exhausted, ub_start_in_solid, ub_nan_dir, ub_coord_eq_n;
} OrcStats;

/* voxel word (GEN_MATERIALS.rs:44-51) */
uint32_t orc_pack_word(uint32_t r7, uint32_t g7, uint32_t b7, int solid);
/* CSV rule (build.rs:50-79): albedo/2 -> 7 bit; returns word for a solid material, 0 for id 0 */
uint32_t orc_material_word_from_csv(int id, int r, int g, int b);

/* pack_into (chunk.rs:125-184) for one 64^3 chunk. `words` are the per-voxel
 * packed words (solid = bit 15).  Writes materials_out (may alias words) and
 * minefield_out. */
void orc_pack_chunk(const uint32_t* words, uint32_t* materials_out, uint8_t* minefield_out);
/* Same, applied to every 64^3 chunk of an [nz][ny][nx] window, in place. */
void orc_build_minefield(uint32_t* materials, uint8_t* minefield, int nx, int ny, int nz);

/* util.rs:9-22 + pipeline.rs:195-207: fills a 192-byte RaytraceUniformData block. */
void orc_fill_frame(uint8_t block[192], const float origin[3], float heading, float pitch,
                    float sun_angle, uint32_t seed, const int32_t lr[3]);

/* pipeline.rs:213-227: after frame `prev` was submitted the host stores old_origin = origin and
 * old_transform = inverse(from_cols(right*0.4, up*0.4, forward)) for the NEXT frame.  Writes those four fields of
 * `next` from prev's origin/right/up/forward (the uniform block already holds right and up pre-scaled).
 * Matrix3::invert restated from cgmath 0.17.0 (not in the tree): cofactor rows / determinant. */
void orc_fill_old_transform(uint8_t next[192], const uint8_t prev[192]);

/* Temporal reprojection (extension; finishes the vestiges raytrace.comp:19,31,58 + pipeline.rs:213-227).
 * Per pixel, after the trace pass and before the denoise passes:
 *   L = lighting texel * 16 (RGBA16_UNORM load); n = 0, Hm = 0
 *   hit (face < 16): r = p - old_origin; v = c0*r.x + c1*r.y + c2*r.z (summed left to right);
 *     if v.z > 0: s = v.xy / v.z; f = ((s + 1) * 0.5) * (W, H); q = floor(f + 0.5);
 *     if q inside the image and the previous frame's record at q has the same face and the same floor(position):
 *       n = min(count_prev(q), MAX_SAMPLES - 1), Hm = mean_prev(q)
 *   mean = (Hm * n + L) / (n + 1); history_cur = {mean, n + 1}; lighting = store16(mean / 16);
 *   completed = store16({mean / 16, (n + 1) / MAX_SAMPLES})
 * hist_* are float4 per pixel {mean rgb, count}; hit_pos_* as OrcBuffers::hit_pos.  MAX_SAMPLES = 8 (raytrace.comp:58). */
void orc_reproject(const uint8_t frame[192], int width, int height, uint16_t* lighting, const float* hit_pos_cur,
                   const float* hit_pos_prev, const float* hist_prev, float* hist_cur, uint16_t* completed, int n_threads);

/* 256-entry LUTs for diffuse_direction (raytrace.comp:190-197): per byte b,
 * sin/cos of theta1 = 2*pi*(b/255) and sin/cos of theta2 = acos(1-2*(b/255)). */
void orc_diffuse_luts(float sin1[256], float cos1[256], float sin2[256], float cos2[256]);

/* raytrace.comp:main for rows [row0,row1). frames_accumulated: number of frames
 * already summed into out->accum (0 = none / accumulation off when accum NULL). */
int orc_trace(const OrcScene* scene, const uint8_t frame[192], int mode,
              OrcBuffers* out, int row0, int row1, OrcStats* stats, int n_threads);

/* analysis hook (tests/analysis_divergence.py): while `log` is non-NULL, orc_trace stores the march length of each of a
 * pixel's five rays (primary, sun1, dif1, sun2, dif2; 0 = not traced) at log[pixel * 5 + ray].  NULL turns it off. */
void orc_set_ray_log(uint16_t* log);
/* single ray, for known-answer tests.  out: [air, face, steps, word, voxel xyz] ; fout: [t, pos xyz] */
void orc_trace_ray(const OrcScene* scene, const int32_t lr[3], const float origin[3],
                   const float dir[3], int32_t iout[7], float fout[4]);

/* bilateral_denoise.comp for rows [row0,row1): src -> dst (RGBA16). swap_bindings
 * reproduces descriptor_sets.rs:37-39 (normal/depth exchanged on odd passes). */
void orc_denoise_pass(const uint16_t* src, uint16_t* dst, const uint16_t* depth,
                      const uint8_t* normal, int width, int height, int size,
                      int swap_bindings, int row0, int row1, int n_threads);

/* finalize.comp for G-buffer rows [row0,row1) -> out_rgba8 (full image, y flipped). */
void orc_finalize(const OrcBuffers* g, const uint16_t* lighting, const uint8_t* blue_noise,
                  uint8_t* out_rgba8, int row0, int row1, int n_threads);

/* Seeded synthetic worlds (kind 0 = heightmap, 1 = biome): fills [nz][ny][nx] words. */
void orc_generate_world(uint32_t* materials, int nx, int ny, int nz, int kind, uint32_t seed);
/* Seeded stream of n unique voxel edits {x,y,z,word} in world coordinates (BASELINE config 5). */
void orc_generate_edits(int32_t* out, int n, int nx, int ny, int nz, uint32_t seed, uint32_t frame);
/* set_block (chunk.rs:115-117) over an edit list, window addressing (w + N/2) mod N. */
void orc_apply_edits(uint32_t* materials, const int32_t* edits, int n, int nx, int ny, int nz);

/* thread->pixel map of raytrace.comp:291-294 and its inverse. */
void orc_wg_to_pixel(int wgx, int wgy, int lx, int ly, int* px, int* py);
void orc_pixel_to_wg(int px, int py, int* wgx, int* wgy);

#ifdef __cplusplus
}
#endif
#endif
