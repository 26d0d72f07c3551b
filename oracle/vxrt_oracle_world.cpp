/*
 * vxrt_oracle_world.cpp — CPU ORACLE for the seeded SYNTHETIC world generators and for voxel edits
 * (TEST INFRASTRUCTURE ONLY, see vxrt_oracle.h).
 *
 * These are not reference algorithms: the reference generator (src/world/generate.rs:31-85) draws
 * from thread_rng and is not reproducible, so the benchmark worlds are defined by
 * DESIGN.md §5 — integer-only value noise + Voronoi biomes + a height-band material rule shaped
 * after generate.rs:31-51 — and this file is the CPU statement of that spec.  The CUDA generator
 * (raytrace_b200/csrc/world.cu) must reproduce it bit for bit.
 */
#include "vxrt_oracle.h"
#include <cstdlib>
#include <cstring>
#include <algorithm>

namespace {

inline uint32_t mix32(uint32_t a) {
    a ^= a >> 16; a *= 0x7feb352du; a ^= a >> 15; a *= 0x846ca68bu; a ^= a >> 16; return a;
}
inline uint32_t hash3(uint32_t x, uint32_t y, uint32_t z, uint32_t s) {
    return mix32(x * 0x9E3779B1u ^ mix32(y * 0x85EBCA77u ^ mix32(z * 0xC2B2AE3Du ^ s)));
}
/* smoothstep-interpolated lattice noise, 16-bit lattice values, exact integer arithmetic */
inline uint32_t vnoise(uint32_t x, uint32_t y, uint32_t o, uint32_t S, uint32_t seed) {
    const uint32_t i = x / S, fx = x % S, j = y / S, fy = y % S;
    const uint64_t S3 = (uint64_t)S * S * S;
    const uint64_t wx = (uint64_t)fx * fx * (3 * S - 2 * fx), wy = (uint64_t)fy * fy * (3 * S - 2 * fy);
    const uint64_t v00 = hash3(i, j, o, seed) >> 16, v10 = hash3(i + 1, j, o, seed) >> 16,
                   v01 = hash3(i, j + 1, o, seed) >> 16, v11 = hash3(i + 1, j + 1, o, seed) >> 16;
    const uint64_t a = (v00 * (S3 - wx) + v10 * wx) / S3, b = (v01 * (S3 - wx) + v11 * wx) / S3;
    return (uint32_t)((a * (S3 - wy) + b * wy) / S3);
}

struct Gen { int nx, ny, nz, kind; uint32_t seed; int sx[64], sy[64], stype[64], samp[64]; };

/* voxel words of misc/materials.csv: CSV rule build.rs:50-79 + pack GEN_MATERIALS.rs:44-51 */
const uint32_t WORDS[7] = {0x0u, 0x1FC07Fu, 0x9F73Du, 0xCD319u, 0xCD9B3u, 0xF8D96u, 0x1BBA73u};

inline void biome_of(const Gen& G, int x, int y, int* type, int* amp8) {
    if (G.kind == 0) { *type = 0; *amp8 = 8; return; }
    long long best = -1; int bk = 0;
    for (int k = 0; k < 64; k++) {
        const long long dx = x - G.sx[k], dy = y - G.sy[k], d = dx * dx + dy * dy;
        if (best < 0 || d < best) { best = d; bk = k; }
    }
    *type = G.stype[bk]; *amp8 = G.samp[bk];
}
inline int column_height(const Gen& G, int x, int y, int* type) {
    int amp8; biome_of(G, x, y, type, &amp8);
    const uint32_t S0 = (uint32_t)std::min(128, G.nx / 2);
    const uint32_t amp = (uint32_t)(G.nz * 3 / 8);
    uint32_t h = (uint32_t)(G.nz / 16);
    uint32_t sum = 0;
    for (uint32_t o = 0; o < 4; o++) sum += ((amp >> o) * vnoise((uint32_t)x, (uint32_t)y, o, S0 >> o, G.seed)) >> 16;
    h += (sum * (uint32_t)amp8) >> 3;
    return std::min((int)h, G.nz * 3 / 4);
}
inline int band_material(int z, uint32_t r, int type) {
    const int lo = (type == 1) ? 4 : (type == 2) ? 5 : 2;
    const int mid = (type >= 2) ? 4 : 5;
    const int hi = 6;
    if (z < 20) return lo;
    if (z < 80) return (int)(r % 60u) < z - 20 ? mid : lo;
    if (z < 160) return (int)(r % 80u) < z - 80 ? hi : mid;
    return hi;
}

}  // namespace

extern "C" {

void orc_generate_world(uint32_t* materials, int nx, int ny, int nz, int kind, uint32_t seed) {
    Gen G; G.nx = nx; G.ny = ny; G.nz = nz; G.kind = kind; G.seed = seed;
    for (uint32_t k = 0; k < 64; k++) {
        G.sx[k] = (int)(hash3(k, 0u, 1u, seed) % (uint32_t)nx); G.sy[k] = (int)(hash3(k, 1u, 1u, seed) % (uint32_t)ny);
        G.stype[k] = (int)(hash3(k, 2u, 1u, seed) % 4u);        G.samp[k] = 4 + (int)(hash3(k, 3u, 1u, seed) % 8u);
    }
#pragma omp parallel for schedule(dynamic, 4)
    for (int y = 0; y < ny; y++) {
        for (int x = 0; x < nx; x++) {
            int type; const int ht = column_height(G, x, y, &type);
            const uint32_t cx = (uint32_t)x >> 4, cy = (uint32_t)y >> 4;
            const uint32_t th = hash3(cx, cy, 7u, seed);
            int trunk_lo = 0, trunk_hi = 0, can_lo = 0, can_hi = 0;
            if ((th & 3u) == 0u) {      /* one tree in a quarter of the 16x16 cells */
                const int tx = (int)(cx * 16 + 3 + ((th >> 2) & 7u)), ty = (int)(cy * 16 + 3 + ((th >> 5) & 7u));
                const int tall = 5 + (int)((th >> 8) & 7u);
                int ttype; const int tb = column_height(G, tx, ty, &ttype);
                if (x == tx && y == ty) { trunk_lo = tb; trunk_hi = tb + tall; }
                if (std::abs(x - tx) <= 2 && std::abs(y - ty) <= 2) { can_lo = tb + tall; can_hi = tb + tall + 3; }
            }
            for (int z = 0; z < nz; z++) {
                uint32_t w = 0;
                if (z < ht) w = WORDS[band_material(z, hash3((uint32_t)x, (uint32_t)y, (uint32_t)z, seed ^ 0xA5A5A5A5u), type)];
                else if (z >= trunk_lo && z < trunk_hi) w = WORDS[3];
                else if (z >= can_lo && z < can_hi) w = WORDS[2];
                materials[((size_t)z * ny + y) * nx + x] = w;
            }
        }
    }
}

/* Seeded edit stream (BASELINE config 5): n UNIQUE voxels (n <= 2^18 / ... see below), the first half
 * inside a 64^3 hot cube and the rest scattered, alternating set-solid (material 4) / clear.
 * World coordinates (texel = (w + N/2) mod N).  out: n * {x, y, z, word}.
 * Uniqueness by construction: edit i takes the i-th element of an affine permutation (odd multiplier
 * modulo a power of two) of the hot cube's 2^18 voxels, or of the whole window; a scattered edit that
 * lands inside the hot cube is skipped (its slot repeats the previous scattered voxel's successor). */
void orc_generate_edits(int32_t* out, int n, int nx, int ny, int nz, uint32_t seed, uint32_t frame) {
    const uint32_t hbx = hash3(frame, 0u, 11u, seed) % (uint32_t)(nx - 63), hby = hash3(frame, 1u, 11u, seed) % (uint32_t)(ny - 63);
    const uint32_t hbz = (uint32_t)(nz / 4) <= (uint32_t)(nz - 64) ? (uint32_t)(nz / 4) : 0u;
    const uint64_t V = (uint64_t)nx * ny * nz;   /* power of two */
    const uint64_t mul_hot = (hash3(frame, 2u, 11u, seed) | 1u), add_hot = hash3(frame, 3u, 11u, seed);
    const uint64_t mul_all = ((uint64_t)hash3(frame, 4u, 11u, seed) << 1) | 1u, add_all = hash3(frame, 5u, 11u, seed);
    const int n_hot = std::min(n / 2, 1 << 18);
    uint64_t j = 0;                              /* index into the whole-window permutation */
    for (int i = 0; i < n; i++) {
        uint32_t x, y, z;
        if (i < n_hot) {
            const uint32_t v = (uint32_t)((mul_hot * (uint64_t)i + add_hot) & 0x3FFFFu);
            x = hbx + (v & 63u); y = hby + ((v >> 6) & 63u); z = hbz + (v >> 12);
        } else {
            for (;;) {
                const uint64_t v = (mul_all * j + add_all) & (V - 1);
                j++;
                x = (uint32_t)(v % (uint64_t)nx); y = (uint32_t)((v / (uint64_t)nx) % (uint64_t)ny); z = (uint32_t)(v / ((uint64_t)nx * ny));
                const bool in_hot = x >= hbx && x < hbx + 64 && y >= hby && y < hby + 64 && z >= hbz && z < hbz + 64;
                if (!in_hot) break;
            }
        }
        out[i * 4 + 0] = (int32_t)x - nx / 2; out[i * 4 + 1] = (int32_t)y - ny / 2; out[i * 4 + 2] = (int32_t)z - nz / 2;
        out[i * 4 + 3] = (int32_t)((i & 1) ? 0u : WORDS[4]);
    }
}

/* set_block (chunk.rs:115-117) in window addressing, applied in list order */
void orc_apply_edits(uint32_t* materials, const int32_t* edits, int n, int nx, int ny, int nz) {
    for (int i = 0; i < n; i++) {
        const uint32_t x = (uint32_t)(edits[i * 4] + nx / 2) & (uint32_t)(nx - 1), y = (uint32_t)(edits[i * 4 + 1] + ny / 2) & (uint32_t)(ny - 1),
                       z = (uint32_t)(edits[i * 4 + 2] + nz / 2) & (uint32_t)(nz - 1);
        materials[((size_t)z * ny + y) * nx + x] = (uint32_t)edits[i * 4 + 3];
    }
}

}  // extern "C"
