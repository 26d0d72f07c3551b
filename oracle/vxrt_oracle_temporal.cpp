/* vxrt_oracle_temporal.cpp — CPU ORACLE of the temporal-reprojection extension.
 * TEST INFRASTRUCTURE ONLY (see vxrt_oracle.h).
 *
 * The reference computes the inputs of a reprojection pass every frame and never reads them: old_origin and
 * old_transform_c0..2 (src/render/pipeline/pipeline.rs:213-227, uniform block shaders/glsl/raytrace.comp:31),
 * MAX_SAMPLES = 8 (raytrace.comp:58, finalize.comp:18) and the never-written completed_buffer (raytrace.comp:19).
 * No reference behaviour exists for the pass itself, so the rule in vxrt_oracle.h (orc_reproject) is this repo's
 * definition; the reference pins only the host-side inputs, restated in orc_fill_old_transform.
 * Scalar fp32, -ffp-contract=off: one IEEE rounding per operator, the order written here. */
#include "vxrt_oracle.h"
#include "orc_math.h"
#include <cstring>
#include <omp.h>

namespace {
const float MAX_SAMPLES = 8.0f;   /* raytrace.comp:58 */

inline V3 ld3(const uint8_t* b, int off) { V3 v; memcpy(&v, b + off, 12); return v; }
inline V3 cross3(V3 a, V3 b) {   /* cgmath Vector3::cross */
    return v3(a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x);
}
}  // namespace

extern "C" {

void orc_fill_old_transform(uint8_t next[192], const uint8_t prev[192]) {
    /* pipeline.rs:214: old_origin = origin */
    memcpy(next + 80, prev + 16, 12);
    /* pipeline.rs:215-223: columns right*0.4, up*0.4, forward (the block's `right`/`up` are those products) */
    const V3 c0 = ld3(prev, 64), c1 = ld3(prev, 48), c2 = ld3(prev, 32);
    /* cgmath 0.17.0 Matrix3::determinant / invert (self[c][r]) */
    const float det = c0.x * (c1.y * c2.z - c2.y * c1.z) - c1.x * (c0.y * c2.z - c2.y * c0.z) + c2.x * (c0.y * c1.z - c1.y * c0.z);
    V3 r0 = v3(0, 0, 0), r1 = r0, r2 = r0;
    if (det != 0.0f) {   /* .expect("Screen space vectors should cover entire coordinate space.") otherwise */
        const V3 a = cross3(c1, c2), b = cross3(c2, c0), c = cross3(c0, c1);
        r0 = v3(a.x / det, a.y / det, a.z / det); r1 = v3(b.x / det, b.y / det, b.z / det); r2 = v3(c.x / det, c.y / det, c.z / det);
    }
    /* from_cols(r0, r1, r2).transpose(): column k of the inverse = (r0[k], r1[k], r2[k]) */
    const V3 i0 = v3(r0.x, r1.x, r2.x), i1 = v3(r0.y, r1.y, r2.y), i2 = v3(r0.z, r1.z, r2.z);
    memcpy(next + 96, &i0, 12); memcpy(next + 112, &i1, 12); memcpy(next + 128, &i2, 12);   /* :224-226 */
}

void orc_reproject(const uint8_t frame[192], int width, int height, uint16_t* lighting, const float* hit_pos_cur,
                   const float* hit_pos_prev, const float* hist_prev, float* hist_cur, uint16_t* completed, int n_threads) {
    const V3 old_origin = ld3(frame, 80), c0 = ld3(frame, 96), c1 = ld3(frame, 112), c2 = ld3(frame, 128);
    const float fw = (float)width, fh = (float)height;
    if (n_threads <= 0) n_threads = omp_get_max_threads();
#pragma omp parallel for num_threads(n_threads) schedule(static)
    for (int py = 0; py < height; py++)
        for (int px = 0; px < width; px++) {
            const size_t pi = (size_t)py * width + px;
            uint16_t* lt = lighting + pi * 4;
            const V3 L = v3(((float)lt[0] / 65535.0f) * 16.0f, ((float)lt[1] / 65535.0f) * 16.0f, ((float)lt[2] / 65535.0f) * 16.0f);
            const float* rec = hit_pos_cur + pi * 4;
            int32_t face; memcpy(&face, rec + 3, 4);
            float n = 0.0f; V3 Hm = v3(0, 0, 0);
            if (face < 16) {
                const V3 p = v3(rec[0], rec[1], rec[2]);
                const V3 r = sub(p, old_origin);
                const V3 v = v3((c0.x * r.x + c1.x * r.y) + c2.x * r.z, (c0.y * r.x + c1.y * r.y) + c2.y * r.z,
                                (c0.z * r.x + c1.z * r.y) + c2.z * r.z);
                if (v.z > 0.0f) {
                    const float fx = ((v.x / v.z + 1.0f) * 0.5f) * fw + 0.5f, fy = ((v.y / v.z + 1.0f) * 0.5f) * fh + 0.5f;
                    if (fx >= 0.0f && fx < fw && fy >= 0.0f && fy < fh) {     /* NaN fails */
                        const size_t qi = (size_t)(int)floorf(fy) * width + (size_t)(int)floorf(fx);
                        const float* prev = hit_pos_prev + qi * 4;
                        int32_t pface; memcpy(&pface, prev + 3, 4);
                        if (pface == face && floorf(prev[0]) == floorf(p.x) && floorf(prev[1]) == floorf(p.y) &&
                            floorf(prev[2]) == floorf(p.z)) {
                            const float* h = hist_prev + qi * 4;
                            n = fminf(h[3], MAX_SAMPLES - 1.0f);
                            Hm = v3(h[0], h[1], h[2]);
                        }
                    }
                }
            }
            const float n1 = n + 1.0f;
            const V3 mean = v3((Hm.x * n + L.x) / n1, (Hm.y * n + L.y) / n1, (Hm.z * n + L.z) / n1);
            float* hc = hist_cur + pi * 4;
            hc[0] = mean.x; hc[1] = mean.y; hc[2] = mean.z; hc[3] = n1;
            lt[0] = (uint16_t)unorm_store(mean.x / 16.0f, 65535.0f); lt[1] = (uint16_t)unorm_store(mean.y / 16.0f, 65535.0f);
            lt[2] = (uint16_t)unorm_store(mean.z / 16.0f, 65535.0f);
            if (completed) {
                uint16_t* o = completed + pi * 4;
                o[0] = lt[0]; o[1] = lt[1]; o[2] = lt[2]; o[3] = (uint16_t)unorm_store(n1 / MAX_SAMPLES, 65535.0f);
            }
        }
}

}  // extern "C"
