// trace_wave_dev.cuh — device helpers of the wavefront form of the FULL-lighting tree (trace_wave.cu), shared with the primary
// stage inside k_trace_fast (trace_fast.cu): starting the two rays of an entry (set-up, the first steps in place, the queue
// push), filing a ray's result, and the stores that end a pixel.
#pragma once
#include "trace_fast_dev.cuh"

namespace {

enum { WAVE_AIR = 1u, WAVE_FACE_SHIFT = 1, WAVE_E4 = 16u, WAVE_SUN1 = 1u << 24 };
// march steps a ray takes where it is spawned, before it is queued.  Half of the shadow and bounce rays end within two steps
// (p50 = 1-2 steps, p90 = 25-31): they never cost a queue slot, a refill or a lane of the march kernel.
#ifndef VXRT_WAVE_PEEL
#define VXRT_WAVE_PEEL 2
#endif

// The result of ray r of a level, from the state trace_ray's loop ends in (:138-183): sky or hit, the face code of the last step
// (E7), the stop position offset along it (:166-180).  Returns the ray's share of the frame counters in k_trace_fast's packing.
__device__ __forceinline__ uint32_t wave_file_result(const TraceParams& P, int level, uint32_t r, float x, float y, float z, float qx, float qy, float qz,
                                                     float dx, float dy, float dz, float lx, float ly, float lz, uint32_t cur, int left, bool sky,
                                                     uint32_t mxm, uint32_t mym, uint32_t mzm, uint32_t nxu, uint32_t nyu) {
    bool air = true;
    uint32_t idx = 0, add = C_RAY + (uint32_t)(P.max_steps - left);
    if (!sky) {                                                                        // sky first, :138-145
        if (cur == 0) {                                                                // :146-159
            air = false;
            const float M1 = 12582912.0f;
            const uint32_t ix = (uint32_t)__float_as_int(__fadd_rd(qx, M1)) & mxm, iy = (uint32_t)__float_as_int(__fadd_rd(qy, M1)) & mym,
                           iz = (uint32_t)__float_as_int(__fadd_rd(qz, M1)) & mzm;
            idx = (iz * nyu + iy) * nxu + ix;
            add += C_HIT;
        } else add += C_LIMIT;                                                         // loop limit exhausted: reads as a miss
    }
    const bool dpx = dx > 0.0f, dpy = dy > 0.0f, dpz = dz > 0.0f;                      // :89-93
    uint32_t normal;
    if (lx < ly) normal = (lx < lz) ? (dpx ? 1u : 0u) : (dpz ? 5u : 4u);               // (E7) :120-136 of the last step
    else normal = (ly < lz) ? (dpy ? 3u : 2u) : (dpz ? 5u : 4u);
    const float off = (normal & 1u) ? -0.001f : 0.001f;                                // :166-180
    if ((normal >> 1) == 0) x = fadd(x, off);
    else if ((normal >> 1) == 1) y = fadd(y, off);
    else z = fadd(z, off);
    P.wave.res_f[level][r] = (air ? (uint32_t)WAVE_AIR : 0u) | (normal << WAVE_FACE_SHIFT);
    if ((r & 1u) && !air) P.wave.res_p[level][r >> 1] = make_float4(x, y, z, __uint_as_float(idx));
    return add;
}

// A ray about to be queued: what k_wave_march keeps per lane.
struct WaveRay { float4 p, d, l; bool live; };

// trace_ray (:82-162) for ray r up to VXRT_WAVE_PEEL iterations of its loop.  `d` is the direction the lighting tree hands over
// (normalised once by sun_ray_dir / diffuse_direction); :83 normalises it again.  A ray that ends here has its result filed;
// `add` receives its share of the frame counters (a ray that goes on is counted where it ends).
__device__ __forceinline__ WaveRay wave_start_ray(const TraceParams& P, int level, uint32_t r, F3 o, F3 d, uint32_t& add) {
    WaveRay w;
    w.live = false;
    const F3 dir = normalize3_shared(d);                                               // :83
    const F3 lpa = rcp_abs3(dir);                                                      // :88
    float qx = fadd(o.x, P.hx), qy = fadd(o.y, P.hy), qz = fadd(o.z, P.hz);
    const bool bad = !(qx == qx) | !(qy == qy) | !(qz == qz) | !(dir.x == dir.x) | !(dir.y == dir.y) | !(dir.z == dir.z);
    const float M1 = 12582912.0f;                                                      // (E2)+(E5)
    const uint32_t mxm = (uint32_t)P.nx - 1u, mym = (uint32_t)P.ny - 1u, mzm = (uint32_t)P.nz - 1u, nxu = (uint32_t)P.nx, nyu = (uint32_t)P.ny;
    uint32_t cur = 0;
    if (!bad) {
        const uint32_t ix = (uint32_t)__float_as_int(__fadd_rd(qx, M1)) & mxm, iy = (uint32_t)__float_as_int(__fadd_rd(qy, M1)) & mym,
                       iz = (uint32_t)__float_as_int(__fadd_rd(qz, M1)) & mzm;
        cur = __ldg(P.minefield + (iz * nyu + iy) * nxu + ix);                         // :106
    }
    if (cur == 0) {                            // (E4) start inside a solid / NaN: a "hit" of word 0 after one iteration (trace.cu)
        P.wave.res_f[level][r] = ((dir.z > 0.0f ? 5u : 4u) << WAVE_FACE_SHIFT) | WAVE_E4;
        if (r & 1u) { const float qnan = CUDART_NAN_F; P.wave.res_p[level][r >> 1] = make_float4(qnan, qnan, qnan, 0.0f); }
        add += C_RAY + 1u + C_HIT;
        return w;
    }
    F3 pos = o;
    int left = P.max_steps;
#pragma unroll
    for (int s = 0; s < VXRT_WAVE_PEEL; s++) {                                         // :113, the scalar form of k_trace_fast's loop
        const float Ms = __int_as_float((int)(cur << 23) + 0x4AC00000);                // 1.5 * 2^(22+cur), :161
        const float ax = flip_if_positive(qx, dir.x), ay = flip_if_positive(qy, dir.y), az = flip_if_positive(qz, dir.z);
        const float mx = fsub(ax, fsub(__fadd_rd(ax, Ms), Ms));                        // (E5)
        const float my = fsub(ay, fsub(__fadd_rd(ay, Ms), Ms));
        const float mz = fsub(az, fsub(__fadd_rd(az, Ms), Ms));
        const float lx = fmul(fadd(0.0001f, mx), lpa.x), ly = fmul(fadd(0.0001f, my), lpa.y), lz = fmul(fadd(0.0001f, mz), lpa.z);   // :119
        const float t = fminf(fminf(lx, ly), lz);                                      // (E7)
        pos.x = fadd(pos.x, fmul(dir.x, t));
        pos.y = fadd(pos.y, fmul(dir.y, t));
        pos.z = fadd(pos.z, fmul(dir.z, t));
        qx = fadd(pos.x, P.hx); qy = fadd(pos.y, P.hy); qz = fadd(pos.z, P.hz);
        const uint32_t ix = (uint32_t)__float_as_int(__fadd_rd(qx, M1)) & mxm, iy = (uint32_t)__float_as_int(__fadd_rd(qy, M1)) & mym,
                       iz = (uint32_t)__float_as_int(__fadd_rd(qz, M1)) & mzm;
        cur = __ldg(P.minefield + (iz * nyu + iy) * nxu + ix);                         // :137
        left--;
        const bool sky = (pos.x >= P.sky_hi[0]) | (pos.x <= P.sky_lo[0]) | (pos.y >= P.sky_hi[1]) | (pos.y <= P.sky_lo[1]) |
                         (pos.z >= P.sky_hi[2]) | (pos.z <= P.sky_lo[2]);              // :138-145, (E8)
        if (sky | (cur == 0) | (left == 0)) {                                          // :138-146, :113
            add += wave_file_result(P, level, r, pos.x, pos.y, pos.z, qx, qy, qz, dir.x, dir.y, dir.z, lx, ly, lz, cur, left, sky, mxm, mym, mzm, nxu, nyu);
            return w;
        }
    }
    w.live = true;
    w.p = make_float4(pos.x, pos.y, pos.z, __uint_as_float(r));
    w.d = make_float4(dir.x, dir.y, dir.z, __uint_as_float(cur));
    w.l = make_float4(lpa.x, lpa.y, lpa.z, __int_as_float(left));
    return w;
}

// the lanes of `m` whose ray goes on take consecutive slots of the level's march queue (one atomic for the group)
__device__ __forceinline__ void wave_push_ray(const TraceParams& P, int level, uint32_t m, const WaveRay& w) {
    const uint32_t live = __ballot_sync(m, w.live);
    if (live == 0) return;
    const int lane = threadIdx.x & 31, leader = __ffs(live) - 1;
    uint32_t base = 0;
    if (lane == leader) base = atomicAdd(P.wave.counts + 2 + level, (uint32_t)__popc(live));
    base = __shfl_sync(m, base, leader);
    if (w.live) {
        const uint32_t j = base + (uint32_t)__popc(live & ((1u << lane) - 1u));
        P.wave.rq_p[level][j] = w.p; P.wave.rq_d[level][j] = w.d; P.wave.rq_l[level][j] = w.l;
    }
}

// A pixel reaches a level of the tree (:324-330 after the primary hit, :336-342 after the first-bounce hit): one entry and its
// sun and bounce rays.  Called by whichever lanes of a warp got here together.  The shadow rays of the group are queued next to
// each other (parallel rays from neighbouring pixels), then the bounce rays.
__device__ __forceinline__ uint32_t wave_spawn(const TraceParams& P, int level, uint32_t pix, F3 hp, uint32_t face, uchar4 nv, uint32_t meta_hi, uint32_t word1) {
    const uint32_t m = __activemask();
    const int lane = threadIdx.x & 31, leader = __ffs(m) - 1;
    uint32_t base = 0;
    if (lane == leader) base = atomicAdd(P.wave.counts + level, (uint32_t)__popc(m));
    base = __shfl_sync(m, base, leader);
    const uint32_t e = base + (uint32_t)__popc(m & ((1u << lane) - 1u));
    P.wave.ent_o[level][e] = make_float4(hp.x, hp.y, hp.z, __uint_as_float(pix));
    P.wave.ent_m[level][e] = face | ((uint32_t)nv.x << 8) | ((uint32_t)nv.y << 16) | meta_hi;
    if (level == 1) P.wave.ent_w[e] = word1;
    uint32_t add = 0;
    const WaveRay s = wave_start_ray(P, level, 2 * e, hp, sun_ray_dir(P, nv), add);                              // :325 / :337
    wave_push_ray(P, level, m, s);
    const WaveRay b = wave_start_ray(P, level, 2 * e + 1, hp, diffuse_direction_shared(P, nv.x, nv.y, face), add);   // :329-330 / :341-342
    wave_push_ray(P, level, m, b);
    return add;
}

// the stores that end a pixel (k_trace_fast's finish_pixel without the per-pixel counters): temporal accumulation, lighting
__device__ __forceinline__ void wave_finish_pixel(const TraceParams& P, size_t pi, F3 lt) {
    if (P.accum) {
        float4 a = P.accum[pi];
        a.x = fadd(a.x, lt.x); a.y = fadd(a.y, lt.y); a.z = fadd(a.z, lt.z); a.w = fadd(a.w, 1.0f);
        P.accum[pi] = a;
        lt = f3(fdiv(a.x, a.w), fdiv(a.y, a.w), fdiv(a.z, a.w));
    }
    P.lighting[pi] = make_ushort4((unsigned short)unorm_store(fmul(lt.x, 0.0625f), 65535.0f), (unsigned short)unorm_store(fmul(lt.y, 0.0625f), 65535.0f),
                                  (unsigned short)unorm_store(fmul(lt.z, 0.0625f), 65535.0f), (unsigned short)unorm_store(0.0625f, 65535.0f));
}

}  // namespace
