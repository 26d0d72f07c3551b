/*
 * vxrt_oracle_trace.cpp — CPU ORACLE for shaders/glsl/raytrace.comp
 * (TEST INFRASTRUCTURE ONLY, see vxrt_oracle.h).  Citations: /root/reference.
 */
#include "vxrt_oracle.h"
#include "orc_math.h"
#include <cstring>
#ifdef _OPENMP
#include <omp.h>
#endif

namespace {

struct Frame {                     /* structs.rs:3-31; only the fields the shader reads */
    float sun_angle; uint32_t seed; V3 origin, forward, up, right; int32_t lr[3];
};
static Frame decode_frame(const uint8_t* b) {
    Frame f;
    memcpy(&f.sun_angle, b + 0, 4); memcpy(&f.seed, b + 4, 4);
    memcpy(&f.origin, b + 16, 12);  memcpy(&f.forward, b + 32, 12);
    memcpy(&f.up, b + 48, 12);      memcpy(&f.right, b + 64, 12);
    memcpy(f.lr, b + 160, 12);
    return f;
}

struct Hit {                       /* raytrace.comp:62-69 + the record Appendix A.7 exposes */
    V3 albedo; bool air; float distance; uint32_t normal; V3 position;
    int32_t voxel[3]; uint32_t word; uint32_t steps;
};

struct Ctx {
    const OrcScene* s; V3 lr; V3 half; V3 dims; OrcStats* st;
};

/* Appendix A.6: nearest, texel = floor(c); outside [0,N) or NaN -> border 0 */
static inline bool texel_of(float c, int n, int* i) {
    if (!(c >= 0.0f && c < (float)n)) return false;
    *i = (int)floorf(c);
    return true;
}
/* get_step, raytrace.comp:78-80 — unnormalised coords mod(position + H, N) (:106,:137) */
static inline uint32_t fetch_minefield(const Ctx& c, V3 p) {
    float tx = gmod(p.x + c.half.x, c.dims.x), ty = gmod(p.y + c.half.y, c.dims.y), tz = gmod(p.z + c.half.z, c.dims.z);
    int ix, iy, iz;
    bool ok = texel_of(tx, c.s->nx, &ix) & texel_of(ty, c.s->ny, &iy) & texel_of(tz, c.s->nz, &iz);
    if (!ok) {
        if (c.st && (tx == c.dims.x || ty == c.dims.y || tz == c.dims.z)) c.st->ub_coord_eq_n++;
        return 0;
    }
    return c.s->minefield[((size_t)iz * c.s->ny + iy) * c.s->nx + ix];
}
/* raytrace.comp:150-154 — normalised coords mod((position + H)/N, 1), texel = floor(u*N) */
static inline uint32_t fetch_world(const Ctx& c, V3 p) {
    float ux = gmod((p.x + c.half.x) / c.dims.x, 1.0f), uy = gmod((p.y + c.half.y) / c.dims.y, 1.0f),
          uz = gmod((p.z + c.half.z) / c.dims.z, 1.0f);
    int ix, iy, iz;
    bool ok = texel_of(ux * c.dims.x, c.s->nx, &ix) & texel_of(uy * c.dims.y, c.s->ny, &iy) &
              texel_of(uz * c.dims.z, c.s->nz, &iz);
    if (!ok) return 0;
    return c.s->materials[((size_t)iz * c.s->ny + iy) * c.s->nx + ix];
}

/* raytrace.comp:82-183 */
static uint16_t* g_ray_log = nullptr;   /* orc_set_ray_log: [w*h][5] march steps of primary, sun1, dif1, sun2, dif2 (0 = not traced) */
static Hit trace_ray(const Ctx& c, V3 origin, V3 direction) {
    direction = normalize3(direction);                                  /* :83 */
    Hit r; memset(&r, 0, sizeof r);
    r.position = origin;                                                /* :85 */
    V3 lpa = v3(1.0f / fabsf(direction.x), 1.0f / fabsf(direction.y), 1.0f / fabsf(direction.z)); /* :88 */
    uint32_t nrm[3] = {direction.x > 0 ? 1u : 0u, direction.y > 0 ? 3u : 2u, direction.z > 0 ? 5u : 4u}; /* :89-93 */
    V3 ms = v3(direction.x > 0 ? -1.0f : 1.0f, direction.y > 0 ? -1.0f : 1.0f, direction.z > 0 ? -1.0f : 1.0f); /* :94-98 */
    if (c.st) {
        c.st->rays++;
        if (direction.x != direction.x || direction.y != direction.y || direction.z != direction.z) c.st->ub_nan_dir++;
    }
    uint32_t cur = fetch_minefield(c, r.position);                      /* :106 */
    float step = (float)((1u << cur) / 2u);                             /* :107 */
    if (c.st && cur == 0) c.st->ub_start_in_solid++;
    r.air = true;       /* DEFINED here: loop-limit exhaustion reads as a miss (reference leaves it uninitialised, :84) */
    r.normal = 0;       /* DEFINED: reference leaves result.normal uninitialised if the loop never runs */
    bool finished = false;
    int limit = c.s->max_steps;                                         /* :109 */
    for (; limit > 0; limit--) {                                        /* :113 */
        r.steps++;
        V3 q = mul(add(r.position, c.half), ms);
        V3 l = v3((0.0001f + gmod(q.x, step)) * lpa.x, (0.0001f + gmod(q.y, step)) * lpa.y,
                  (0.0001f + gmod(q.z, step)) * lpa.z);                 /* :119 */
        float t; uint32_t n;
        if (l.x < l.y) {                                                /* :120-136 */
            if (l.x < l.z) { t = l.x; n = nrm[0]; } else { t = l.z; n = nrm[2]; }
        } else {
            if (l.y < l.z) { t = l.y; n = nrm[1]; } else { t = l.z; n = nrm[2]; }
        }
        r.position = add(r.position, muls(direction, t));
        r.normal = n;
        cur = fetch_minefield(c, r.position);                           /* :137 */
        if (fabsf(r.position.x - c.lr.x) >= c.half.x || fabsf(r.position.y - c.lr.y) >= c.half.y ||
            fabsf(r.position.z - c.lr.z) >= c.half.z) {                 /* :138-145 */
            r.air = true; finished = true; break;
        } else if (cur <= 0) {                                          /* :146-159 */
            r.air = false; finished = true;
            r.word = fetch_world(c, r.position);
            r.albedo = v3((float)((r.word >> 14) & 0x7F) / 127.0f, (float)((r.word >> 7) & 0x7F) / 127.0f,
                          (float)(r.word & 0x7F) / 127.0f);
            break;
        }
        step = (float)((1u << cur) / 2u);                               /* :161 */
    }
    if (!finished && c.st) c.st->ub_limit_exhausted++;
    if (c.st) { c.st->steps += r.steps; if (!r.air) c.st->hits++; }
    r.voxel[0] = floor_to_i32(r.position.x); r.voxel[1] = floor_to_i32(r.position.y); r.voxel[2] = floor_to_i32(r.position.z);
    r.distance = length3(sub(origin, r.position));                      /* :164 */
    const float off = 0.001f;                                           /* :166-180 */
    switch (r.normal) {
        case 0: r.position.x += off; break; case 1: r.position.x -= off; break;
        case 2: r.position.y += off; break; case 3: r.position.y -= off; break;
        case 4: r.position.z += off; break; case 5: r.position.z -= off; break;
    }
    return r;
}

struct Luts { float s1[256], c1[256], s2[256], c2[256]; };

/* raytrace.comp:189-212 with the per-byte tables of orc_diffuse_luts */
static V3 diffuse_direction(const Luts& L, const uint8_t* noise, uint32_t normal) {
    int r = noise[0], g = noise[1];
    V3 d = v3(L.s1[r] * L.s2[g], L.c1[r] * L.s2[g], L.c2[g]);
    switch (normal) {
        case 0: d.x += 1.0f; break; case 1: d.x -= 1.0f; break;
        case 2: d.y += 1.0f; break; case 3: d.y -= 1.0f; break;
        case 4: d.z += 1.0f; break; case 5: d.z -= 1.0f; break;
    }
    return normalize3(d);
}

/* raytrace.comp:259-269 */
static V3 sun_color(V3 sd) {
    float horizon = sqrtf(sd.x * sd.x + sd.y * sd.y);
    float sun_amount = fminf(1.0f - horizon, 0.02f) * 50.0f;
    V3 main_color = v3(0.9647f * 2.0f, 0.7843f * 2.0f, 0.8824f * 2.0f);
    V3 sunset_color = v3(0.7412f * 2.0f, 0.2157f * 2.0f, 0.1686f * 2.0f);
    if (sd.z >= 0.0f) return mix3(sunset_color, main_color, sun_amount);
    return mix3(sunset_color, v3(0, 0, 0), sun_amount * 2.0f);
}

/* raytrace.comp:271-288 (colour only: compared under tolerance, powf may differ by ulps) */
static V3 sample_sky(V3 d, V3 sd, V3 sunlight, bool include_sun) {
    V3 bright = v3(0.5294f, 0.8275f, 0.9647f), dark = v3(0.0863f, 0.1294f, 0.2196f);
    float sl = clampf(((sunlight.x + sunlight.y) + sunlight.z) * 0.2f - 0.02f, 0.0f, 1.0f);
    float horizon = powf(sqrtf(d.x * d.x + d.y * d.y), mixf(40.0f, 10.0f, sl));
    float sun_amount = 1.0f - 0.5f * length3(sub(sd, d));
    float halo = powf(sun_amount, mixf(5.0f, 1.0f, sl));
    float bright_amount = fminf(horizon + halo * 0.5f, 1.0f);
    V3 color = mix3(dark, bright, bright_amount * fmaxf(sl, 0.1f));
    float p5 = powf(sun_amount, 5.0f) * 0.5f;
    color = add(color, muls(sunlight, p5));
    if (sun_amount > 0.98f && include_sun) color = add(color, sunlight);
    return color;
}

static inline const uint8_t* noise_texel(const uint8_t* bn, float cx, float cy) {
    /* nearest, unnormalised, clamp-to-edge (render_data.rs:110-128) */
    int ix = (int)floorf(cx), iy = (int)floorf(cy);
    if (ix < 0) ix = 0; if (ix > 511) ix = 511; if (iy < 0) iy = 0; if (iy > 511) iy = 511;
    return bn + ((size_t)iy * 512 + ix) * 4;
}

}  // namespace

extern "C" {

void orc_set_ray_log(uint16_t* log) { g_ray_log = log; }

void orc_trace_ray(const OrcScene* scene, const int32_t lr[3], const float origin[3],
                   const float dir[3], int32_t iout[7], float fout[4]) {
    Ctx c; c.s = scene; c.st = 0;
    c.lr = v3((float)lr[0], (float)lr[1], (float)lr[2]);
    c.dims = v3((float)scene->nx, (float)scene->ny, (float)scene->nz);
    c.half = v3((float)(scene->nx / 2), (float)(scene->ny / 2), (float)(scene->nz / 2));
    Hit h = trace_ray(c, v3(origin[0], origin[1], origin[2]), v3(dir[0], dir[1], dir[2]));
    iout[0] = h.air; iout[1] = (int32_t)h.normal; iout[2] = (int32_t)h.steps; iout[3] = (int32_t)h.word;
    iout[4] = h.voxel[0]; iout[5] = h.voxel[1]; iout[6] = h.voxel[2];
    fout[0] = h.distance; fout[1] = h.position.x; fout[2] = h.position.y; fout[3] = h.position.z;
}

/* raytrace.comp:290-385 */
int orc_trace(const OrcScene* scene, const uint8_t frame[192], int mode, OrcBuffers* out,
              int row0, int row1, OrcStats* stats, int n_threads) {
    const Frame f = decode_frame(frame);
    const int W = out->width, H = out->height;
    Luts L; orc_diffuse_luts(L.s1, L.c1, L.s2, L.c2);
    /* :317-318 — per-frame constants; cos/sin in double then rounded (same rule as the CUDA host side) */
    float ca = (float)cos((double)f.sun_angle), sa = (float)sin((double)f.sun_angle);
    V3 sunangle = normalize3(v3(ca * 0.5f + (f.sun_angle - 0.5f) * 0.5f, sa, ca));
    V3 sunlight = sun_color(sunangle);
    /* :298-303 */
    const uint8_t* seed_px = noise_texel(scene->blue_noise, (float)(f.seed % 512u), (float)(f.seed / 512u));
    const float noise_base_x = ((float)seed_px[0] / 255.0f) * 255.0f, noise_base_y = ((float)seed_px[1] / 255.0f) * 255.0f;
    if (n_threads <= 0) {
#ifdef _OPENMP
        n_threads = omp_get_max_threads();
#else
        n_threads = 1;
#endif
    }
    OrcStats total; memset(&total, 0, sizeof total);
#pragma omp parallel num_threads(n_threads)
    {
        OrcStats st; memset(&st, 0, sizeof st);
        Ctx c; c.s = scene; c.st = &st;
        c.lr = v3((float)f.lr[0], (float)f.lr[1], (float)f.lr[2]);
        c.dims = v3((float)scene->nx, (float)scene->ny, (float)scene->nz);
        c.half = v3((float)(scene->nx / 2), (float)(scene->ny / 2), (float)(scene->nz / 2));
#pragma omp for schedule(dynamic, 4)
        for (int py = row0; py < row1; py++) {
            for (int px = 0; px < W; px++) {
                const size_t pi = (size_t)py * W + px;
                uint64_t steps0 = st.steps, rays0 = st.rays;
                V3 screen = v3((float)px / (float)W * 2.0f - 1.0f, (float)py / (float)H * 2.0f - 1.0f, 0); /* :296-297 */
                int wgx, wgy; orc_pixel_to_wg(px, py, &wgx, &wgy);
                float nox = noise_base_x + (float)(wgx * 8), noy = noise_base_y + (float)(wgy * 8);       /* :304 */
                st.noise_fetches += 2;
                V3 ray_start = f.origin;
                V3 ray_dir = normalize3(add(add(f.forward, muls(f.right, screen.x)), muls(f.up, screen.y))); /* :307-311 */
                if (-ray_start.y > c.half.y) {                                                             /* :312-315 */
                    float space = -ray_start.y - c.half.y;
                    ray_start = add(ray_start, muls(ray_dir, space / ray_dir.y + 0.0001f));
                }
                V3 light = v3(0, 0, 0);
                Hit primary = trace_ray(c, ray_start, ray_dir);                                            /* :320 */
                uint16_t* rl = g_ray_log ? g_ray_log + pi * 5 : nullptr;   /* analysis hook: march length of each of the five rays */
                if (rl) { rl[0] = (uint16_t)primary.steps; rl[1] = rl[2] = rl[3] = rl[4] = 0; }
                if (primary.air) {
                    light = sample_sky(ray_dir, sunangle, sunlight, true);                                 /* :322 */
                } else if (mode != ORC_MODE_PRIMARY) {
                    const uint8_t* nv = noise_texel(scene->blue_noise, gmod(nox, 512.0f), gmod(noy, 512.0f)); /* :324 */
                    st.noise_fetches++;
                    V3 sun_dir = normalize3(add(sunangle, v3(((float)nv[0] / 255.0f) * 0.05f, ((float)nv[1] / 255.0f) * 0.05f, 0.0f * 0.05f))); /* :186 */
                    Hit sun1 = trace_ray(c, primary.position, sun_dir);                                    /* :325 */
                    if (rl) rl[1] = (uint16_t)sun1.steps;
                    if (sun1.air) light = add(light, sunlight);
                    if (mode == ORC_MODE_FULL) {
                        V3 d1 = diffuse_direction(L, nv, primary.normal);                                  /* :329 */
                        Hit dif1 = trace_ray(c, primary.position, d1);                                     /* :330 */
                        if (rl) rl[2] = (uint16_t)dif1.steps;
                        if (dif1.air) {
                            light = add(light, sample_sky(d1, sunangle, sunlight, true));
                        } else {
                            V3 light2 = v3(0, 0, 0);
                            /* :336 — +2/512 on unnormalised coords lands in the same texel */
                            const uint8_t* nv2 = noise_texel(scene->blue_noise, gmod(nox + 2.0f / 512.0f, 512.0f), gmod(noy + 2.0f / 512.0f, 512.0f));
                            st.noise_fetches++;
                            V3 sun_dir2 = normalize3(add(sunangle, v3(((float)nv2[0] / 255.0f) * 0.05f, ((float)nv2[1] / 255.0f) * 0.05f, 0.0f * 0.05f)));
                            Hit sun2 = trace_ray(c, dif1.position, sun_dir2);                              /* :337 */
                            if (rl) rl[3] = (uint16_t)sun2.steps;
                            if (sun2.air) light2 = add(light2, sunlight);
                            V3 d2 = diffuse_direction(L, nv2, dif1.normal);                                /* :341 */
                            Hit dif2 = trace_ray(c, dif1.position, d2);                                    /* :342 */
                            if (rl) rl[4] = (uint16_t)dif2.steps;
                            if (dif2.air) light2 = add(light2, sample_sky(d2, sunangle, sunlight, true));
                            light2 = mul(light2, dif1.albedo);                                             /* :346 */
                            /* + dif1.emission == 0 (:155, :347) */
                            light = add(light, light2);
                        }
                    }
                }
                /* temporal accumulation (extension): running sum, mean is what gets stored */
                if (out->accum) {
                    float* a = out->accum + pi * 4;
                    a[0] += light.x; a[1] += light.y; a[2] += light.z; a[3] += 1.0f;
                    light = v3(a[0] / a[3], a[1] / a[3], a[2] / a[3]);
                }
                if (out->lighting) {                                                                       /* :352-356 */
                    uint16_t* o = out->lighting + pi * 4;
                    o[0] = (uint16_t)unorm_store(light.x / 16.0f, 65535.0f); o[1] = (uint16_t)unorm_store(light.y / 16.0f, 65535.0f);
                    o[2] = (uint16_t)unorm_store(light.z / 16.0f, 65535.0f); o[3] = (uint16_t)unorm_store(1.0f / 16.0f, 65535.0f);
                }
                if (out->depth) {                                                                          /* :357-365 */
                    uint32_t d = 0xFFFF;
                    if (!primary.air) d = f2u_trunc(length3(sub(f.origin, primary.position)) * 32.0f);
                    out->depth[pi] = (uint16_t)d;
                }
                if (out->normal) out->normal[pi] = primary.air ? 16 : (uint8_t)primary.normal;             /* :366-370 */
                if (out->albedo) {                                                                         /* :371-375 */
                    uint8_t* o = out->albedo + pi * 4;
                    V3 a = primary.air ? v3(1, 1, 1) : primary.albedo;
                    o[0] = (uint8_t)unorm_store(a.x, 255.0f); o[1] = (uint8_t)unorm_store(a.y, 255.0f);
                    o[2] = (uint8_t)unorm_store(a.z, 255.0f); o[3] = 255;
                }
                if (out->emission) {                                                                       /* :376-380; emission == 0 */
                    uint8_t* o = out->emission + pi * 4;
                    o[0] = o[1] = o[2] = 0; o[3] = primary.air ? 0 : 255;
                }
                if (out->fog) {                                                                            /* :381-385 */
                    uint8_t* o = out->fog + pi * 4;
                    V3 s = sample_sky(ray_dir, sunangle, sunlight, false);
                    o[0] = (uint8_t)unorm_store(s.x / 2.0f, 255.0f); o[1] = (uint8_t)unorm_store(s.y / 2.0f, 255.0f);
                    o[2] = (uint8_t)unorm_store(s.z / 2.0f, 255.0f); o[3] = 255;
                }
                if (out->hit_voxel) { out->hit_voxel[pi * 3] = primary.voxel[0]; out->hit_voxel[pi * 3 + 1] = primary.voxel[1]; out->hit_voxel[pi * 3 + 2] = primary.voxel[2]; }
                if (out->hit_face) out->hit_face[pi] = primary.air ? 16 : (uint8_t)primary.normal;
                if (out->hit_pos) {
                    float* o = out->hit_pos + pi * 4;
                    const int32_t fc = primary.air ? 16 : (int32_t)primary.normal;
                    o[0] = primary.air ? 0.0f : primary.position.x; o[1] = primary.air ? 0.0f : primary.position.y;
                    o[2] = primary.air ? 0.0f : primary.position.z; memcpy(o + 3, &fc, 4);
                }
                if (out->hit_word) out->hit_word[pi] = primary.air ? 0u : primary.word;
                if (out->hit_t) out->hit_t[pi] = primary.distance;
                if (out->steps) out->steps[pi] = (uint32_t)(st.steps - steps0);
                if (out->rays) out->rays[pi] = (uint8_t)(st.rays - rays0);
            }
        }
#pragma omp critical
        {
            total.rays += st.rays; total.steps += st.steps; total.hits += st.hits; total.noise_fetches += st.noise_fetches;
            total.ub_limit_exhausted += st.ub_limit_exhausted; total.ub_start_in_solid += st.ub_start_in_solid;
            total.ub_nan_dir += st.ub_nan_dir; total.ub_coord_eq_n += st.ub_coord_eq_n;
        }
    }
    if (stats) *stats = total;
    return 0;
}

}  // extern "C"
