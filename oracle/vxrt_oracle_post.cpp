/*
 * vxrt_oracle_post.cpp — CPU ORACLE for shaders/glsl/bilateral_denoise.comp and
 * shaders/glsl/finalize.comp (TEST INFRASTRUCTURE ONLY, see vxrt_oracle.h).
 */
#include "vxrt_oracle.h"
#include "orc_math.h"
#ifdef _OPENMP
#include <omp.h>
#endif

namespace {
/* bilateral_denoise.comp:40-86, in source order */
struct Tap { int dx, dy; float w; };
static const Tap TAPS[36] = {
    {0, 1, 0.092566f},  {0, -1, 0.092566f},  {1, 0, 0.092566f},   {-1, 0, 0.092566f},
    {1, 1, 0.058434f},  {-1, 1, 0.058434f},  {-1, -1, 0.058434f}, {1, -1, 0.058434f},
    {2, 0, 0.023205f},  {-2, 0, 0.023205f},  {0, 2, 0.023205f},   {0, -2, 0.023205f},
    {2, 2, 0.003672f},  {-2, 2, 0.003672f},  {-2, -2, 0.003672f}, {2, -2, 0.003672f},
    {2, 1, 0.014648f},  {-2, 1, 0.014648f},  {-2, -1, 0.014648f}, {2, -1, 0.014648f},
    {1, 2, 0.014648f},  {-1, 2, 0.014648f},  {-1, -2, 0.014648f}, {1, -2, 0.014648f},
    {3, 0, 0.002289f},  {-3, 0, 0.002289f},  {0, 3, 0.002289f},   {0, -3, 0.002289f},
    {3, 1, 0.001445f},  {-3, 1, 0.001445f},  {-3, -1, 0.001445f}, {3, -1, 0.001445f},
    {1, 3, 0.001445f},  {-1, 3, 0.001445f},  {-1, -3, 0.001445f}, {1, -3, 0.001445f},
};

/* finalize.comp:21-31 */
static inline float filmic_curve(float x) {
    if (x < 0.3f) return x * x;
    else if (x < 1.13333f) return x * 0.6f - 0.09f;
    else if (x < 2.5f) return 1.0f - 0.219512195116f * (x - 2.5f) * (x - 2.5f);
    else return 1.0f;
}
}  // namespace

extern "C" {

/* bilateral_denoise.comp:35-89.  With swap_bindings the "depth" image is the R8 normal
 * buffer and the "normal" image is the R16 depth buffer (descriptor_sets.rs:37-39). */
void orc_denoise_pass(const uint16_t* src, uint16_t* dst, const uint16_t* depth,
                      const uint8_t* normal, int width, int height, int size,
                      int swap_bindings, int row0, int row1, int n_threads) {
    if (n_threads <= 0) {
#ifdef _OPENMP
        n_threads = omp_get_max_threads();
#else
        n_threads = 1;
#endif
    }
#pragma omp parallel for schedule(static) num_threads(n_threads)
    for (int y = row0; y < row1; y++) {
        for (int x = 0; x < width; x++) {
            const size_t pi = (size_t)y * width + x;
            uint32_t dv = swap_bindings ? (uint32_t)normal[pi] : (uint32_t)depth[pi];
            uint32_t nv = swap_bindings ? (uint32_t)depth[pi] : (uint32_t)normal[pi];
            float center_distance = (float)dv / 256.0f;                        /* :37 */
            uint32_t center_normal = nv;                                       /* :38 */
            if (center_normal < 16) {
                float total_weight = 0.146634f;                                /* :41 */
                float sr = ((float)src[pi * 4] / 65535.0f) * total_weight, sg = ((float)src[pi * 4 + 1] / 65535.0f) * total_weight,
                      sb = ((float)src[pi * 4 + 2] / 65535.0f) * total_weight;
                for (int t = 0; t < 36; t++) {                                 /* SAMPLE, :23-33 */
                    int qx = x + TAPS[t].dx * size, qy = y + TAPS[t].dy * size;
                    if (qx < 0) qx = 0; if (qy < 0) qy = 0;
                    if (qx >= width) qx = width - 1; if (qy >= height) qy = height - 1;
                    const size_t qi = (size_t)qy * width + qx;
                    uint32_t qd = swap_bindings ? (uint32_t)normal[qi] : (uint32_t)depth[qi];
                    uint32_t qn = swap_bindings ? (uint32_t)depth[qi] : (uint32_t)normal[qi];
                    float dist = (float)qd / 256.0f;
                    float distance_difference = 4.0f * fabsf(center_distance - dist);
                    float normal_difference = qn == center_normal ? 0.0f : 10.0f;
                    float weight = TAPS[t].w / (distance_difference + normal_difference + 1.0f);
                    total_weight += weight;
                    sr += ((float)src[qi * 4] / 65535.0f) * weight;
                    sg += ((float)src[qi * 4 + 1] / 65535.0f) * weight;
                    sb += ((float)src[qi * 4 + 2] / 65535.0f) * weight;
                }
                dst[pi * 4] = (uint16_t)unorm_store(sr / total_weight, 65535.0f);   /* :87 */
                dst[pi * 4 + 1] = (uint16_t)unorm_store(sg / total_weight, 65535.0f);
                dst[pi * 4 + 2] = (uint16_t)unorm_store(sb / total_weight, 65535.0f);
                dst[pi * 4 + 3] = 65535;
            } else {
                for (int k = 0; k < 4; k++) dst[pi * 4 + k] = src[pi * 4 + k];      /* :89 */
            }
        }
    }
}

/* finalize.comp:33-64 */
void orc_finalize(const OrcBuffers* g, const uint16_t* lighting, const uint8_t* blue_noise,
                  uint8_t* out_rgba8, int row0, int row1, int n_threads) {
    const int W = g->width, H = g->height;
    if (n_threads <= 0) {
#ifdef _OPENMP
        n_threads = omp_get_max_threads();
#else
        n_threads = 1;
#endif
    }
#pragma omp parallel for schedule(static) num_threads(n_threads)
    for (int y = row0; y < row1; y++) {
        for (int x = 0; x < W; x++) {
            const size_t pi = (size_t)y * W + x;
            float c[3];
            uint32_t depth = g->depth[pi];
            float fog_amount = (float)depth / (32.0f * 128.0f * 8.0f);             /* :47 */
            if (fog_amount > 1.0f) fog_amount = 1.0f;
            const uint8_t* bn = blue_noise + ((size_t)(y % 512) * 512 + (x % 512)) * 4;   /* :55-57 */
            for (int k = 0; k < 3; k++) {
                float albedo = (float)g->albedo[pi * 4 + k] / 255.0f;
                float emission = ((float)g->emission[pi * 4 + k] / 255.0f) * 4.0f;
                float light = ((float)lighting[pi * 4 + k] / 65535.0f) * 16.0f;
                float v = albedo * light + emission;                               /* :40 */
                if (depth < 0xFFFF) {                                              /* :44-50 */
                    float fog = ((float)g->fog[pi * 4 + k] / 255.0f) * 2.0f;
                    v = mixf(v, fog, fog_amount);
                }
                v = filmic_curve(v);
                v += ((float)bn[k] / 255.0f) / 128.0f;                             /* :58 */
                c[k] = v;
            }
            uint8_t* o = out_rgba8 + ((size_t)(H - y - 1) * W + x) * 4;            /* :60-63 */
            o[0] = (uint8_t)unorm_store(c[0], 255.0f); o[1] = (uint8_t)unorm_store(c[1], 255.0f);
            o[2] = (uint8_t)unorm_store(c[2], 255.0f); o[3] = 255;
        }
    }
}

}  // extern "C"
