/*
 * vxrt_oracle.cpp — CPU ORACLE (test infrastructure, see vxrt_oracle.h).
 *
 * Hand restatement of the reference's GLSL passes in scalar fp32.  Build with
 *   g++ -O2 -fopenmp -ffp-contract=off -fno-fast-math
 * so that every +,-,*,/ and sqrt is one IEEE-754 binary32 round-to-nearest-even
 * operation in exactly the order written.  The CUDA
 * kernels execute the same sequence with __fadd_rn/__fmul_rn/__fdiv_rn/
 * __fsqrt_rn, which is what makes hit voxel / face / word bit-exact.
 *
 * Citations are file:line under /root/reference.
 */
#include "vxrt_oracle.h"
#include <cmath>
#include <cstring>
#include <cstdlib>
#include <vector>
#ifdef _OPENMP
#include <omp.h>
#endif

#include "orc_math.h"

extern "C" {

/* GEN_MATERIALS.rs:44-51 — note bit 15 (solid) overlaps red bit 1 */
uint32_t orc_pack_word(uint32_t r7, uint32_t g7, uint32_t b7, int solid) {
    uint32_t albedo = (r7 << 14) | (g7 << 7) | b7;
    return ((solid ? 1u : 0u) << 15) | albedo;
}

/* build.rs:50-79: albedo channel = csv value / 2 (integer); id 0 is air (not solid) */
uint32_t orc_material_word_from_csv(int id, int r, int g, int b) {
    return orc_pack_word((uint32_t)(r / 2), (uint32_t)(g / 2), (uint32_t)(b / 2), id != 0);
}

/* chunk.rs:125-184.  LOD k (k=1..6) holds one bool per 2^k-aligned cube: "any solid inside". */
void orc_pack_chunk(const uint32_t* words, uint32_t* materials_out, uint8_t* minefield_out) {
    const int CS = 64, CV = CS * CS * CS;
    std::vector<std::vector<uint8_t> > lods;
    for (int vol = CV / 8; vol > 0; vol /= 8) lods.push_back(std::vector<uint8_t>(vol, 0));
    std::vector<uint8_t> solid(CV);
    for (int i = 0; i < CV; i++) {
        solid[i] = (words[i] >> 15) & 1u;
        if (solid[i]) {
            int x = i % CS, y = (i / CS) % CS, z = i / (CS * CS);   /* util.rs:94-106, x fastest */
            int stride = CS / 2;
            for (size_t l = 0; l < lods.size(); l++) {
                x >>= 1; y >>= 1; z >>= 1;
                int li = (z * stride + y) * stride + x;
                if (lods[l][li]) break;
                lods[l][li] = 1;
                stride /= 2;
            }
        }
        materials_out[i] = words[i];
    }
    if (!lods[5][0]) {                       /* chunk.rs:155-161: wholly empty chunk */
        for (int i = 0; i < CV; i++) { materials_out[i] = 0; minefield_out[i] = 6; }
        return;
    }
    for (int i = 0; i < CV; i++) {
        if (solid[i]) { minefield_out[i] = 0; continue; }
        int x = i % CS, y = (i / CS) % CS, z = i / (CS * CS);
        int stride = CS / 2;
        for (size_t l = 0; l < lods.size(); l++) {
            x >>= 1; y >>= 1; z >>= 1;
            if (lods[l][(z * stride + y) * stride + x]) { minefield_out[i] = (uint8_t)(l + 1); break; }
            stride /= 2;
        }
    }
}

void orc_build_minefield(uint32_t* materials, uint8_t* minefield, int nx, int ny, int nz) {
    const int CS = 64, CV = CS * CS * CS;
    int cxn = nx / CS, cyn = ny / CS, czn = nz / CS;
#pragma omp parallel
    {
        std::vector<uint32_t> w(CV), m(CV);
        std::vector<uint8_t> f(CV);
#pragma omp for collapse(2) schedule(dynamic)
        for (int cz = 0; cz < czn; cz++)
            for (int cy = 0; cy < cyn; cy++)
                for (int cx = 0; cx < cxn; cx++) {
                    for (int z = 0; z < CS; z++)
                        for (int y = 0; y < CS; y++)
                            memcpy(&w[(z * CS + y) * CS],
                                   &materials[((size_t)(cz * CS + z) * ny + (cy * CS + y)) * nx + cx * CS],
                                   CS * sizeof(uint32_t));
                    orc_pack_chunk(w.data(), m.data(), f.data());
                    for (int z = 0; z < CS; z++)
                        for (int y = 0; y < CS; y++) {
                            size_t o = ((size_t)(cz * CS + z) * ny + (cy * CS + y)) * nx + cx * CS;
                            memcpy(&materials[o], &m[(z * CS + y) * CS], CS * sizeof(uint32_t));
                            memcpy(&minefield[o], &f[(z * CS + y) * CS], CS);
                        }
                }
    }
}

/* util.rs:9-22 (camera basis) + pipeline.rs:195-207 (uniform fill); layout structs.rs:3-31 */
void orc_fill_frame(uint8_t block[192], const float origin[3], float heading, float pitch,
                    float sun_angle, uint32_t seed, const int32_t lr[3]) {
    const float HALF_PI = 1.57079632679489661923f;
    float fwd[3] = {cosf(heading) * cosf(pitch), sinf(heading) * cosf(pitch), sinf(pitch)};
    float up[3] = {cosf(heading) * cosf(pitch + HALF_PI), sinf(heading) * cosf(pitch + HALF_PI),
                   sinf(pitch + HALF_PI)};
    /* cgmath cross: (a.y*b.z - a.z*b.y, a.z*b.x - a.x*b.z, a.x*b.y - a.y*b.x) */
    float right[3] = {fwd[1] * up[2] - fwd[2] * up[1], fwd[2] * up[0] - fwd[0] * up[2],
                      fwd[0] * up[1] - fwd[1] * up[0]};
    float up4[3] = {up[0] * 0.4f, up[1] * 0.4f, up[2] * 0.4f};
    float right4[3] = {right[0] * 0.4f, right[1] * 0.4f, right[2] * 0.4f};
    memset(block, 0, 192);
    memcpy(block + 0, &sun_angle, 4);
    memcpy(block + 4, &seed, 4);
    memcpy(block + 16, origin, 12);
    memcpy(block + 32, fwd, 12);
    memcpy(block + 48, up4, 12);
    memcpy(block + 64, right4, 12);
    memcpy(block + 160, lr, 12);   /* rotation  (shader: lr)  */
    memcpy(block + 176, lr, 12);   /* space_offset (shader: lso, unread) */
}

/* raytrace.comp:190-197.  theta is formed in fp32 as the shader does; sin/cos/acos are
 * evaluated in double and rounded once so that every libm gives the same table. */
void orc_diffuse_luts(float sin1[256], float cos1[256], float sin2[256], float cos2[256]) {
    const float PI = 3.1415926535897932384626433832795f;
    for (int b = 0; b < 256; b++) {
        float n = (float)b / 255.0f;
        float theta1 = PI * 2.0f * n;
        float theta2 = (float)acos((double)(1.0f - 2.0f * n));
        sin1[b] = (float)sin((double)theta1);
        cos1[b] = (float)cos((double)theta1);
        sin2[b] = (float)sin((double)theta2);
        cos2[b] = (float)cos((double)theta2);
    }
}

/* raytrace.comp:291-294 */
void orc_wg_to_pixel(int wgx, int wgy, int lx, int ly, int* px, int* py) {
    *px = (wgx - wgx % 16) * 8 + wgx % 16 + lx * 16;
    *py = (wgy - wgy % 16) * 8 + wgy % 16 + ly * 16;
}
void orc_pixel_to_wg(int px, int py, int* wgx, int* wgy) {
    *wgx = (px / 128) * 16 + px % 16;
    *wgy = (py / 128) * 16 + py % 16;
}

}  // extern "C"
