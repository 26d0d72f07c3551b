// trace_wave.cu — the FULL-lighting trace pass (shaders/glsl/raytrace.comp:290-385) as a wavefront.
//
// k_trace_fast keeps a warp's 32 pixels together through the five rays of the lighting tree (:320-350).  A round lasts as long as
// its longest ray and bounce rays are short on the median and long in the tail, so its march loop runs at 13.9 of 32 lanes on C3
// (DESIGN.md 4.1); k_trace_pool (trace_pool.cu) refills lanes inside one kernel and loses more to the latency of its in-warp
// service than it gains.  Here the tree is cut at its rays:
//
//   k_trace_fast<.., WAVE>   primary ray, G-buffer, fog; a pixel whose primary ray hits pushes ENTRY e of level 0 and starts its two
//                            rays 2e (sun, :325) and 2e + 1 (bounce, :329-330): set-up and the first two steps in place - half of
//                            them end there - the others go to the level's march queue (position, direction, 1/|direction|, texel)
//   k_wave_march (level 0)   persistent warps; a lane marches whatever ray it holds, and when a third of the warp has finished the
//                            finished lanes file their results and take the next rays of the queue
//   k_wave_resolve<0>        one thread per entry: bounce ray in the sky -> the pixel's light (:331-332); else the entry of level 1
//                            (second noise texel :336, sun2 :337, bounce2 :341-342)
//   k_wave_march (level 1)
//   k_wave_resolve<1>        light2, albedo, the pixel's light (:338-348)
//
// The march kernel holds nothing but rays (no shading code, 32 registers, 64 warps per SM), and everything with a chain of
// dependent fetches (noise -> tables -> first texel, the voxel word of a hit) runs one thread per entry on full warps.  Every ray
// is the ray of the sequential form - same origin, direction, steps, the same fp32 operations - and a pixel's light is summed
// in the order of :320-350, so images, running sum and frame counters equal k_trace_fast's bit for bit
// (tests/test_gpu_parity.py::test_wavefront_is_invisible).
#include "trace_wave_dev.cuh"
#include <stdlib.h>

namespace {

constexpr int MW = 4;                          // warps per CTA of the march kernel (each works alone)
#ifndef VXRT_WAVE_MINB
#define VXRT_WAVE_MINB 16                      // CTAs per SM: 32 registers, every warp slot (the march is latency-bound)
#endif
#ifndef VXRT_WAVE_IDLE
#define VXRT_WAVE_IDLE 12                      // finished lanes that end a march phase while the level still has rays to hand out
#endif
#ifndef VXRT_WAVE_CHUNK
#define VXRT_WAVE_CHUNK 64                     // rays a warp reserves per atomic on the level's counter
#endif
#ifndef VXRT_WAVE_VOTE
#define VXRT_WAVE_VOTE 4                       // march steps between two looks at the number of marching lanes
#endif

__device__ __forceinline__ uint32_t wave_tex3d(unsigned long long tex, uint32_t x, uint32_t y, uint32_t z) {
    uint32_t r0, r1, r2, r3;
    asm("tex.3d.v4.u32.s32 {%0,%1,%2,%3}, [%4, {%5,%6,%7,%7}];" : "=r"(r0), "=r"(r1), "=r"(r2), "=r"(r3) : "l"(tex), "r"(x), "r"(y), "r"(z));
    return r0;
}

// trace_ray's loop (:113-162) for the rays of one level.  LR0 / DIMS / FIELD as in k_trace_fast; the occupancy field must not
// cross a 4 GiB line ((E15), checked by vxrt_trace_wave_applies).
template <int LR0, int DIMS, int FIELD>
__global__ void __launch_bounds__(32 * MW, VXRT_WAVE_MINB) k_wave_march(const __grid_constant__ TraceParams P, int level) {
    constexpr bool SD = DIMS != 0;
    constexpr int LX = DIMS & 0xFF, LY = (DIMS >> 8) & 0xFF, LZ = (DIMS >> 16) & 0xFF;
    const uint32_t FULL = 0xffffffffu;
    const int lane = threadIdx.x & 31;
    const uint32_t below = (1u << lane) - 1u;
    // per-frame march constants pinned in registers through volatile shared memory, as in k_trace_fast (from the constant bank
    // ptxas reloads them inside the loop)
    __shared__ MarchConsts s_mc;
    if (threadIdx.x == 0) {
        s_mc.hx = P.hx; s_mc.hy = P.hy; s_mc.hz = P.hz;
        s_mc.mxm = (uint32_t)P.nx - 1u; s_mc.mym = (uint32_t)P.ny - 1u; s_mc.mzm = (uint32_t)P.nz - 1u;
        s_mc.nxu = (uint32_t)P.nx; s_mc.nyu = (uint32_t)P.ny;
        s_mc.base = (unsigned long long)P.minefield; s_mc.one2 = P.one2;
    }
    __syncthreads();
    const volatile MarchConsts* mc = &s_mc;
    const float hx = SD ? (float)(1 << (LX - 1)) : mc->hx, hy = SD ? (float)(1 << (LY - 1)) : mc->hy, hz = SD ? (float)(1 << (LZ - 1)) : mc->hz;
    const uint32_t mxm = SD ? (1u << LX) - 1u : mc->mxm, mym = SD ? (1u << LY) - 1u : mc->mym, mzm = SD ? (1u << LZ) - 1u : mc->mzm,
                   nxu = mc->nxu, nyu = mc->nyu;
    const unsigned long long base64 = mc->base;
    const uint32_t blo = (uint32_t)base64, bhi = (uint32_t)(base64 >> 32);              // (E15)
    const float M1 = 12582912.0f;              // 1.5 * 2^23
    f2 one2; one2 = mc->one2;
    const f2 hxy = pk(hx, hy);
    const float4* const rq_p = P.wave.rq_p[level];
    const float4* const rq_d = P.wave.rq_d[level];
    const float4* const rq_l = P.wave.rq_l[level];
    uint32_t* const next = P.wave.counts + 4 + level;
    const uint32_t n_rays = __ldcg(P.wave.counts + 2 + level);

    auto sky_test = [&](float x, float y, float z) -> bool {                             // :138-145, (E8)
        if (LR0 == 2) return fmaxf(fmaxf(fabsf(x), fabsf(y)), fabsf(z)) >= hx;
        if (LR0) return (fabsf(x) >= hx) | (fabsf(y) >= hy) | (fabsf(z) >= hz);
        return (x >= P.sky_hi[0]) | (x <= P.sky_lo[0]) | (y >= P.sky_hi[1]) | (y <= P.sky_lo[1]) | (z >= P.sky_hi[2]) | (z <= P.sky_lo[2]);
    };

    // the ray this lane holds.  has: it holds one; left: iterations of :113's limit to go; done: the march ended
    bool has = false, done = false;
    uint32_t r = 0, cur = 0;
    int left = 0;
    f2 pxy = pk(0.f, 0.f), dxy = pk(0.f, 0.f), lpxy = pk(0.f, 0.f), qxy = pk(0.f, 0.f);
    float pz = 0.f, dz = 0.f, lpz = 0.f, qz = 0.f, lx = 0.f, ly = 0.f, lz = 0.f;
    uint32_t c_steps = 0, c_rh = 0, c_limit = 0;   // frame counters of this lane: march steps; rays | hits << 16; limit exhausted
    uint32_t q_next = 0, q_end = 0;            // warp-uniform: the reserved rays not handed out yet
    bool exhausted = n_rays == 0;

    for (;;) {
        // ---- finished lanes file their result -----------------------------------------------------------------------------
        if (has && done) {
            const float x = lo2(pxy), y = hi2(pxy);
            const uint32_t add = wave_file_result(P, level, r, x, y, pz, lo2(qxy), hi2(qxy), qz, lo2(dxy), hi2(dxy), dz, lx, ly, lz, cur, left,
                                                  sky_test(x, y, pz), mxm, mym, mzm, nxu, nyu);
            c_steps += add & 0xFFFFFu; c_rh += ((add >> 20) & 7u) | (((add >> 23) & 7u) << 16); c_limit += add >> 29;
            has = false;
        }
        __syncwarp();
        // ---- idle lanes take the next rays of the level -------------------------------------------------------------------
        const uint32_t idle = __ballot_sync(FULL, !has);
        if (idle != 0 && !exhausted) {
            if (q_next == q_end) {
                uint32_t nb = 0;
                if (lane == 0) nb = atomicAdd(next, (uint32_t)VXRT_WAVE_CHUNK);
                nb = __shfl_sync(FULL, nb, 0);
                if (nb >= n_rays) exhausted = true;
                else { q_next = nb; q_end = min(nb + (uint32_t)VXRT_WAVE_CHUNK, n_rays); }
            }
            if (!exhausted) {
                const uint32_t take = min((uint32_t)__popc(idle), q_end - q_next), rank = (uint32_t)__popc(idle & below);
                if (!has && rank < take) {
                    const uint32_t j = q_next + rank;
                    const float4 p4 = __ldcg(rq_p + j), d4 = __ldcg(rq_d + j), l4 = __ldcg(rq_l + j);
                    pxy = pk(p4.x, p4.y); pz = p4.z; r = __float_as_uint(p4.w);
                    dxy = pk(d4.x, d4.y); dz = d4.z; cur = __float_as_uint(d4.w);
                    lpxy = pk(l4.x, l4.y); lpz = l4.z; left = __float_as_int(l4.w);
                    qxy = add2(pxy, hxy); qz = fadd(pz, hz);
                    has = true; done = false;
                }
                q_next += take;
            }
        }
        if (!__any_sync(FULL, has)) break;
        // ---- march --------------------------------------------------------------------------------------------------------
        if (has && !done) {
            float Ms = __int_as_float((int)(cur << 23) + 0x4AC00000);                  // 1.5 * 2^(22+cur)
            done = true;                       // the loop is left at the end of the ray, or (rarely) by the vote below
            for (;;) {                                                                 // :113, (E10)
                bool ended = false;
#pragma unroll
                for (int u = 0; u < VXRT_WAVE_VOTE; u++) {
                    const float dx = lo2(dxy), dy = hi2(dxy);
                    const float ax = flip_if_positive(lo2(qxy), dx), ay = flip_if_positive(hi2(qxy), dy), az = flip_if_positive(qz, dz);
                    const f2 axy = pk(ax, ay), Ms2 = pk(Ms, Ms);
                    const f2 mxy = sub2(axy, sub2(add2_rd(axy, Ms2), Ms2));            // (E5)
                    const float mz = fsub(az, fsub(__fadd_rd(az, Ms), Ms));
                    const f2 lxy = mul2(add2(mxy, pk(0.0001f, 0.0001f)), lpxy);        // :119
                    lz = fmul(fadd(0.0001f, mz), lpz);
                    lx = lo2(lxy); ly = hi2(lxy);
                    const float t = fminf(fminf(lx, ly), lz);                          // (E7)
                    const f2 sxy = mul2(dxy, pk(t, t));
                    pxy = fma2(sxy, one2, pxy);                                        // (E14)
                    pz = fadd(pz, fmul(dz, t));
                    qxy = add2(pxy, hxy); qz = fadd(pz, hz);
                    const f2 fxy = add2_rd(qxy, pk(M1, M1));
                    const uint32_t ix = (uint32_t)__float_as_int(lo2(fxy)) & mxm, iy = (uint32_t)__float_as_int(hi2(fxy)) & mym,
                                   iz = (uint32_t)__float_as_int(__fadd_rd(qz, M1)) & mzm;
                    if (FIELD == 1) cur = wave_tex3d(P.field_tex, ix, iy, iz);
                    else cur = ldg_u8_halves((iz * nyu + iy) * nxu + (ix | blo), bhi); // :137, (E15)
                    left--;
                    if (sky_test(lo2(pxy), hi2(pxy), pz) | (cur == 0) | (left == 0)) { ended = true; break; }   // :138-146, :113
                    Ms = __int_as_float((int)(cur << 23) + 0x4AC00000);                // :161
                }
                if (ended) break;
                // while the level has rays left, the phase ends when enough lanes have finished to be worth a refill
                if (!exhausted && __popc(__activemask()) <= 32 - VXRT_WAVE_IDLE) { done = false; break; }
            }
        }
        __syncwarp();
    }
    // frame counters: rays, steps, hits, limit exhausted
    const uint32_t st = __reduce_add_sync(FULL, c_steps), ry = __reduce_add_sync(FULL, c_rh & 0xFFFFu), ht = __reduce_add_sync(FULL, c_rh >> 16),
                   lm = __reduce_add_sync(FULL, c_limit);
    if (lane == 0) {
        if (ry) atomicAdd(P.counters + 0, (unsigned long long)ry);
        if (st) atomicAdd(P.counters + 1, (unsigned long long)st);
        if (ht) atomicAdd(P.counters + 2, (unsigned long long)ht);
        if (lm) atomicAdd(P.counters + 4, (unsigned long long)lm);
    }
}

// One thread per entry of a level, after the level's rays have marched.
//   LEVEL 0 (:326-336): bounce ray in the sky -> light = sun1 + sky, the pixel is done; else the entry of level 1.
//   LEVEL 1 (:338-348): light2 = sun2 + sky2, times the first-bounce albedo, plus sun1.
#ifndef VXRT_WAVE_RESOLVE_MINB
#define VXRT_WAVE_RESOLVE_MINB 5              // 51 registers: the pass is a chain of dependent fetches per entry and lives on occupancy
#endif
template <int LEVEL>
__global__ void __launch_bounds__(256, VXRT_WAVE_RESOLVE_MINB) k_wave_resolve(const __grid_constant__ TraceParams P) {
    const uint32_t n = __ldcg(P.wave.counts + LEVEL);
    const F3 sunlight = f3(P.sl_r, P.sl_g, P.sl_b), zero = f3(0, 0, 0);
    uint32_t noise_fetches = 0, c_steps = 0, c_rays = 0, c_hits = 0, c_limit = 0;   // level 0: the rays of level 1 that end where they are spawned
    for (uint32_t e = blockIdx.x * blockDim.x + threadIdx.x; e < n; e += gridDim.x * blockDim.x) {
        const float4 o4 = __ldcg(P.wave.ent_o[LEVEL] + e);
        const uint32_t meta = __ldcg(P.wave.ent_m[LEVEL] + e);
        const uint32_t fs = __ldcg(P.wave.res_f[LEVEL] + 2 * e), fd = __ldcg(P.wave.res_f[LEVEL] + 2 * e + 1);
        const uint32_t pix = __float_as_uint(o4.w);
        const int px = (int)(pix & 0xFFFFu), py = (int)(pix >> 16);
        const size_t pi = (size_t)py * P.width + px;
        const bool sun_air = (fs & WAVE_AIR) != 0, dif_air = (fd & WAVE_AIR) != 0;
        if (LEVEL == 0) {
            if (dif_air) {
                const F3 sky = sky_with_sun(P, diffuse_direction_shared(P, (meta >> 8) & 0xFFu, (meta >> 16) & 0xFFu, meta & 7u));
                wave_finish_pixel(P, pi, add3(sun_air ? sunlight : zero, sky));        // :331-332
            } else {
                const float4 p4 = __ldcg(P.wave.res_p[0] + e);
                const uint32_t word1 = (fd & WAVE_E4) ? 0u : __ldg(P.materials + __float_as_uint(p4.w));   // (E3); (E4): word 0
                float nox, noy;
                noise_offset(P, px, py, nox, noy);
                const float two_over = fdiv(2.0f, 512.0f);
                const uchar4 nv = noise_texel(P.noise, gmod512(fadd(nox, two_over)), gmod512(fadd(noy, two_over)));   // :336
                noise_fetches++;
                const uint32_t add = wave_spawn(P, 1, pix, f3(p4.x, p4.y, p4.z), (fd >> WAVE_FACE_SHIFT) & 7u, nv, sun_air ? (uint32_t)WAVE_SUN1 : 0u, word1);
                c_steps += add & 0xFFFFFu; c_rays += (add >> 20) & 7u; c_hits += (add >> 23) & 7u; c_limit += add >> 29;
            }
        } else {
            F3 light2 = sun_air ? sunlight : zero;                                     // 0 + sunlight
            if (dif_air) light2 = add3(light2, sky_with_sun(P, diffuse_direction_shared(P, (meta >> 8) & 0xFFu, (meta >> 16) & 0xFFu, meta & 7u)));   // :343-345
            light2 = mul3(light2, decode_albedo(P, __ldcg(P.wave.ent_w + e)));         // :346 (+ emission 0, :347)
            wave_finish_pixel(P, pi, add3((meta & WAVE_SUN1) ? sunlight : zero, light2));   // :348
        }
    }
    if (LEVEL == 0) {
        const uint32_t v[5] = {c_rays, c_steps, c_hits, noise_fetches, c_limit};
#pragma unroll
        for (int k = 0; k < 5; k++) {
            const uint32_t s = __reduce_add_sync(0xffffffffu, v[k]);
            if ((threadIdx.x & 31) == 0 && s) atomicAdd(P.counters + k, (unsigned long long)s);
        }
    }
}

#define VXRT_DIMS(lx, ly, lz) ((lx) | ((ly) << 8) | ((lz) << 16))
template <int DIMS, int FIELD>
const void* march_kernel(int form) {
    if (form == 2) return (const void*)k_wave_march<2, DIMS, FIELD>;
    if (form == 1) return (const void*)k_wave_march<1, DIMS, FIELD>;
    if (FIELD != 0) return nullptr;
    return (const void*)k_wave_march<0, DIMS, 0>;
}
const void* march_kernel_for(const TraceParams& p) {
    const bool lr0 = p.lrx == 0.0f && p.lry == 0.0f && p.lrz == 0.0f;
    const int form = !lr0 ? 0 : (p.nx == p.ny && p.ny == p.nz) ? 2 : 1;
    const int dims = VXRT_DIMS(p.lx, p.ly, p.lz);
    if (p.field == 1 && form != 0) {
        if (dims == VXRT_DIMS(9, 9, 9)) return march_kernel<VXRT_DIMS(9, 9, 9), 1>(form);
        if (dims == VXRT_DIMS(11, 11, 8)) return march_kernel<VXRT_DIMS(11, 11, 8), 1>(form);
    }
    if (dims == VXRT_DIMS(9, 9, 9)) return march_kernel<VXRT_DIMS(9, 9, 9), 0>(form);
    if (dims == VXRT_DIMS(11, 11, 8)) return march_kernel<VXRT_DIMS(11, 11, 8), 0>(form);
    return march_kernel<0, 0>(form);
}

}  // namespace

cudaError_t vxrt_preload_trace_wave() {
    cudaFuncAttributes a;
    cudaError_t e;
    for (int form = 0; form < 3; form++) {
        const void* ks[5] = {march_kernel<0, 0>(form), march_kernel<VXRT_DIMS(9, 9, 9), 0>(form), march_kernel<VXRT_DIMS(11, 11, 8), 0>(form),
                             march_kernel<VXRT_DIMS(9, 9, 9), 1>(form), march_kernel<VXRT_DIMS(11, 11, 8), 1>(form)};
        for (const void* k : ks) if (k && (e = cudaFuncGetAttributes(&a, k)) != cudaSuccess) return e;
    }
    if ((e = cudaFuncGetAttributes(&a, k_wave_resolve<0>)) != cudaSuccess) return e;
    return cudaFuncGetAttributes(&a, k_wave_resolve<1>);
}

// bytes of one handle's wavefront buffers for `entries` pixels, and their carving (16-byte arrays first)
size_t vxrt_trace_wave_bytes(size_t entries) {
    const size_t n = (entries + 63) & ~(size_t)63;
    return 256 + 2 * (n * 16 /*ent_o*/ + 3 * 2 * n * 16 /*rq_p, rq_d, rq_l*/ + n * 16 /*res_p*/ + n * 4 /*ent_m*/ + 2 * n * 4 /*res_f*/) + n * 4 /*ent_w*/;
}
void vxrt_trace_wave_carve(WaveBufs* w, void* base, size_t entries) {
    const size_t n = (entries + 63) & ~(size_t)63;
    char* p = (char*)base;
    w->counts = (uint32_t*)p; p += 256;
    for (int l = 0; l < 2; l++) {
        w->ent_o[l] = (float4*)p; p += n * 16;
        w->rq_p[l] = (float4*)p; p += 2 * n * 16;
        w->rq_d[l] = (float4*)p; p += 2 * n * 16;
        w->rq_l[l] = (float4*)p; p += 2 * n * 16;
        w->res_p[l] = (float4*)p; p += n * 16;
    }
    for (int l = 0; l < 2; l++) {
        w->ent_m[l] = (uint32_t*)p; p += n * 4;
        w->res_f[l] = (uint32_t*)p; p += 2 * n * 4;
    }
    w->ent_w = (uint32_t*)p;
    w->cap = (uint32_t)entries;
}

// The wavefront form runs FULL lighting on the lean march's preconditions, whole bands only, without the per-pixel parity
// buffers (their step and ray counts are summed per pixel by k_trace_fast) and without the reprojection records.
bool vxrt_trace_wave_applies(const TraceParams& p) {
    // opt-in (vxrt_set_trace_form, or VXRT_WAVE=1 for unchanged callers): measured on the B200 it wins where rays are long (C5, 40
    // steps per ray: 0.946 -> 0.846 ms) and loses where they are short (C3 0.438 -> 0.536 ms, C4 2.57 -> 2.81 ms) - see DESIGN.md 4.1
    static const bool env_on = [] { const char* e = getenv("VXRT_WAVE"); return e && atoi(e) != 0; }();
    if (!(p.trace_form == VXRT_TRACE_WAVEFRONT || (p.trace_form == VXRT_TRACE_AUTO && env_on)) || !p.fast || !p.wave.counts || p.mode != VXRT_MODE_FULL || p.hit_voxel || p.hit_pos) return false;
    if ((p.flags & VXRT_FLAG_RAY_HANDOFF) || p.own_row0 > p.row0 || p.own_row1 < p.row1 || p.width > 65535 || p.height > 65535) return false;
    if ((size_t)(p.row1 - p.row0) * p.width > p.wave.cap || p.field > 1) return false;
    const unsigned long long base = (unsigned long long)p.minefield, bytes = (unsigned long long)p.nx * p.ny * p.nz;
    if ((base & (unsigned long long)(p.nx - 1)) != 0 || (base & 0xFFFFFFFFull) + bytes > 0x100000000ull) return false;   // (E15)
    dim3 g, b;
    return vxrt_trace_fast_wave_kernel(p, &g, &b) != nullptr && march_kernel_for(p) != nullptr;
}

cudaError_t vxrt_launch_trace_wave(const TraceParams& p, cudaStream_t s) {
    if (p.row1 - p.row0 <= 0) return cudaSuccess;
    cudaError_t e = cudaMemsetAsync(p.wave.counts, 0, 32, s);
    if (e != cudaSuccess) return e;
    dim3 grid, block;
    const void* k1 = vxrt_trace_fast_wave_kernel(p, &grid, &block);
    void* a1[1] = {(void*)&p};
    if ((e = cudaLaunchKernel(k1, grid, block, a1, 0, s)) != cudaSuccess) return e;
    static const int sms = [] { int d = 0, n = 148; cudaGetDevice(&d); cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, d); return n; }();
    const void* km = march_kernel_for(p);
    for (int level = 0; level < 2; level++) {
        void* am[2] = {(void*)&p, (void*)&level};
        if ((e = cudaLaunchKernel(km, dim3((unsigned)(sms * VXRT_WAVE_MINB)), dim3(32 * MW), am, 0, s)) != cudaSuccess) return e;
        if (level == 0) k_wave_resolve<0><<<sms * VXRT_WAVE_RESOLVE_MINB, 256, 0, s>>>(p);
        else k_wave_resolve<1><<<sms * VXRT_WAVE_RESOLVE_MINB, 256, 0, s>>>(p);
        if ((e = cudaGetLastError()) != cudaSuccess) return e;
    }
    return cudaSuccess;
}
