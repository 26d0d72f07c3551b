// trace.cu — k_trace: per-pixel ray generation, occupancy-LOD march, shadow/diffuse rays, shading
// and G-buffer stores.  Replaces shaders/glsl/raytrace.comp (reference, 400 lines of GLSL).
//
// Numeric contract: the march advances `position += direction * len` in fp32 and the
// hit voxel / face depend on the accumulated rounding, so every rounding operation of the reference
// expression tree is issued here as an explicit round-to-nearest intrinsic (__fadd_rn, __fmul_rn,
// __fdiv_rn, __fsqrt_rn): nvcc never contracts those into FMAs.  FMAs and integer shortcuts are used
// only where the result is provably the same real number rounded once (noted inline).
#include "trace_common.cuh"

namespace {

// raytrace.comp:82-183.  See DESIGN.md §3 for the equivalence notes marked (E1)..(E4).
__device__ __forceinline__ Hit trace_ray(const TraceParams& P, F3 origin, F3 direction, RayStats& st) {
    direction = normalize3(direction);                                         // :83
    Hit r;
    F3 pos = origin;                                                           // :85
    const float lpax = fdiv(1.0f, fabsf(direction.x)), lpay = fdiv(1.0f, fabsf(direction.y)),
                lpaz = fdiv(1.0f, fabsf(direction.z));                         // :88
    const bool px = direction.x > 0.0f, py = direction.y > 0.0f, pz = direction.z > 0.0f;
    const uint32_t nrmx = px ? 1u : 0u, nrmy = py ? 3u : 2u, nrmz = pz ? 5u : 4u;   // :89-93
    const float msx = px ? -1.0f : 1.0f, msy = py ? -1.0f : 1.0f, msz = pz ? -1.0f : 1.0f;   // :94-98
    st.rays++;
    const uint32_t mxm = (uint32_t)P.nx - 1u, mym = (uint32_t)P.ny - 1u, mzm = (uint32_t)P.nz - 1u;

    // (E2) texel = floor(mod(p + H, N)) == floor(p + H) & (N-1) for every finite p reachable here:
    // mod() of a power of two is exact and can never round up to N.
    float qx = fadd(pos.x, P.hx), qy = fadd(pos.y, P.hy), qz = fadd(pos.z, P.hz);
    const bool bad = !(qx == qx) | !(qy == qy) | !(qz == qz) | !(direction.x == direction.x) |
                     !(direction.y == direction.y) | !(direction.z == direction.z);
    uint32_t cur = 0;
    if (!bad) {
        uint32_t ix = (uint32_t)__float2int_rd(qx) & mxm, iy = (uint32_t)__float2int_rd(qy) & mym,
                 iz = (uint32_t)__float2int_rd(qz) & mzm;
        cur = __ldg(P.minefield + ((((size_t)iz << P.ly) + iy) << P.lx) + ix);                // :106
    }
    r.air = true; r.normal = 0; r.word = 0; r.steps = 0; r.albedo = f3(0, 0, 0);
    if (bad | (cur == 0)) {
        // (E4) start inside a solid, NaN origin or NaN direction: step_size = 0 -> mod(x, 0) = NaN ->
        // every comparison false -> z branch -> position = NaN -> border texel 0 -> "hit" of word 0
        // after exactly one iteration.  IEEE fixes this outcome, so it is produced directly.
        const float qnan = CUDART_NAN_F;
        r.air = false; r.normal = nrmz; r.steps = 1; r.position = f3(qnan, qnan, qnan);
        r.vx = r.vy = r.vz = INT32_MIN; r.distance = qnan;
        st.steps += 1; st.hits += 1;
        return r;
    }
    // step = 2^(cur-1), inv = 2^(1-cur), built from exponent bits (cur in 1..31)
    float step = __int_as_float((126 + (int)cur) << 23), inv = __int_as_float((128 - (int)cur) << 23);
    bool finished = false;
    int limit = P.max_steps;
    uint32_t n = 0;
    for (; limit > 0; limit--) {                                               // :113
        // :119  len = (0.0001 + mod((pos + H) * muls, step)) * length_per_axis
        // (E1) mod(a, s) = a - s*floor(a/s): a/s == a*inv exactly (power of two), s*floor() is exact,
        // so fma(-s, floor(a*inv), a) is the same single rounding as the reference's subtraction.
        float ax = fmul(qx, msx), ay = fmul(qy, msy), az = fmul(qz, msz);
        float mx = __fmaf_rn(-step, floorf(fmul(ax, inv)), ax);
        float my = __fmaf_rn(-step, floorf(fmul(ay, inv)), ay);
        float mz = __fmaf_rn(-step, floorf(fmul(az, inv)), az);
        float lx = fmul(fadd(0.0001f, mx), lpax), ly = fmul(fadd(0.0001f, my), lpay), lz = fmul(fadd(0.0001f, mz), lpaz);
        float t;
        if (lx < ly) {                                                         // :120-136
            if (lx < lz) { t = lx; n = nrmx; } else { t = lz; n = nrmz; }
        } else {
            if (ly < lz) { t = ly; n = nrmy; } else { t = lz; n = nrmz; }
        }
        pos.x = fadd(pos.x, fmul(direction.x, t));
        pos.y = fadd(pos.y, fmul(direction.y, t));
        pos.z = fadd(pos.z, fmul(direction.z, t));
        qx = fadd(pos.x, P.hx); qy = fadd(pos.y, P.hy); qz = fadd(pos.z, P.hz);
        uint32_t ix = (uint32_t)__float2int_rd(qx) & mxm, iy = (uint32_t)__float2int_rd(qy) & mym,
                 iz = (uint32_t)__float2int_rd(qz) & mzm;
        const size_t idx = ((((size_t)iz << P.ly) + iy) << P.lx) + ix;
        cur = __ldg(P.minefield + idx);                                        // :137
        if (fabsf(fsub(pos.x, P.lrx)) >= P.hx || fabsf(fsub(pos.y, P.lry)) >= P.hy ||
            fabsf(fsub(pos.z, P.lrz)) >= P.hz) {                               // :138-145 sky first
            r.air = true; finished = true; limit--; break;
        } else if (cur == 0) {                                                 // :146-159
            r.air = false; finished = true; limit--;
            // (E3) the normalised-coordinate fetch mod((p+H)/N, 1)*N addresses the same texel
            r.word = __ldg(P.materials + idx);
            r.albedo = f3(fdiv((float)((r.word >> 14) & 0x7F), 127.0f), fdiv((float)((r.word >> 7) & 0x7F), 127.0f),
                          fdiv((float)(r.word & 0x7F), 127.0f));
            break;
        }
        step = __int_as_float((126 + (int)cur) << 23); inv = __int_as_float((128 - (int)cur) << 23);   // :161
    }
    r.steps = (uint32_t)(P.max_steps - limit);
    r.normal = n;
    if (!finished) st.limit++;
    st.steps += r.steps; if (!r.air) st.hits++;
    r.vx = floor_to_i32(pos.x); r.vy = floor_to_i32(pos.y); r.vz = floor_to_i32(pos.z);
    r.distance = length3(sub3(origin, pos));                                   // :164
    const float off = 0.001f;                                                  // :166-180
    switch (r.normal) {
        case 0: pos.x = fadd(pos.x, off); break; case 1: pos.x = fsub(pos.x, off); break;
        case 2: pos.y = fadd(pos.y, off); break; case 3: pos.y = fsub(pos.y, off); break;
        case 4: pos.z = fadd(pos.z, off); break; case 5: pos.z = fsub(pos.z, off); break;
    }
    r.position = pos;
    return r;
}

// One thread per pixel; a warp covers an 8x4 pixel tile, a block 16x8 pixels (4 warps).
__global__ void __launch_bounds__(128) k_trace_general(const __grid_constant__ TraceParams P) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int px = blockIdx.x * 16 + (warp & 1) * 8 + (lane & 7);
    const int py = P.row0 + blockIdx.y * 8 + (warp >> 1) * 4 + (lane >> 3);
    RayStats st = {0, 0, 0, 0};
    uint32_t noise_fetches = 0;
    const bool active = px < P.width && py < P.row1;
    if (active) {
        const size_t pi = (size_t)py * P.width + px;
        const F3 origin = f3(P.ox, P.oy, P.oz), sunangle = f3(P.sun_x, P.sun_y, P.sun_z), sunlight = f3(P.sl_r, P.sl_g, P.sl_b);
        const float sx = fsub(fmul(fdiv((float)px, (float)P.width), 2.0f), 1.0f);       // :296-297
        const float sy = fsub(fmul(fdiv((float)py, (float)P.height), 2.0f), 1.0f);
        // :298-304 — gl_WorkGroupID recovered from the pixel (inverse of the 16-stride interleave, :291-294)
        const int wgx = (px >> 7) * 16 + (px & 15), wgy = (py >> 7) * 16 + (py & 15);
        const float nox = fadd(P.noise_base_x, (float)(wgx * 8)), noy = fadd(P.noise_base_y, (float)(wgy * 8));
        noise_fetches = 2;
        F3 ray_start = origin;
        const F3 fwd = f3(P.fx, P.fy, P.fz), right = f3(P.rx, P.ry, P.rz), up = f3(P.ux, P.uy, P.uz);
        const F3 ray_dir = normalize3(add3(add3(fwd, scale3(right, sx)), scale3(up, sy)));   // :307-311
        if (-ray_start.y > P.hy) {                                                           // :312-315
            float space = fsub(-ray_start.y, P.hy);
            ray_start = add3(ray_start, scale3(ray_dir, fadd(fdiv(space, ray_dir.y), 0.0001f)));
        }
        F3 light = f3(0, 0, 0);
        const Hit primary = trace_ray(P, ray_start, ray_dir, st);                            // :320
        if (primary.air) {
            light = sample_sky(ray_dir, sunangle, sunlight, true);                           // :322
        } else if (P.mode != VXRT_MODE_PRIMARY) {
            const uchar4 nv = noise_texel(P.noise, gmod512(nox), gmod512(noy));              // :324
            noise_fetches++;
            const F3 jitter = f3(fmul(fdiv((float)nv.x, 255.0f), 0.05f), fmul(fdiv((float)nv.y, 255.0f), 0.05f), 0.0f);
            const Hit sun1 = trace_ray(P, primary.position, normalize3(add3(sunangle, jitter)), st);   // :325, :186
            if (sun1.air) light = add3(light, sunlight);
            if (P.mode == VXRT_MODE_FULL) {
                const F3 d1 = diffuse_direction(P, nv.x, nv.y, primary.normal);              // :329
                const Hit dif1 = trace_ray(P, primary.position, d1, st);                     // :330
                if (dif1.air) {
                    light = add3(light, sample_sky(d1, sunangle, sunlight, true));
                } else {
                    F3 light2 = f3(0, 0, 0);
                    const float two_over = fdiv(2.0f, 512.0f);
                    const uchar4 nv2 = noise_texel(P.noise, gmod512(fadd(nox, two_over)), gmod512(fadd(noy, two_over)));   // :336
                    noise_fetches++;
                    const F3 jitter2 = f3(fmul(fdiv((float)nv2.x, 255.0f), 0.05f), fmul(fdiv((float)nv2.y, 255.0f), 0.05f), 0.0f);
                    const Hit sun2 = trace_ray(P, dif1.position, normalize3(add3(sunangle, jitter2)), st);   // :337
                    if (sun2.air) light2 = add3(light2, sunlight);
                    const F3 d2 = diffuse_direction(P, nv2.x, nv2.y, dif1.normal);           // :341
                    const Hit dif2 = trace_ray(P, dif1.position, d2, st);                    // :342
                    if (dif2.air) light2 = add3(light2, sample_sky(d2, sunangle, sunlight, true));
                    light2 = mul3(light2, dif1.albedo);                                      // :346 (+ emission 0, :347)
                    light = add3(light, light2);
                }
            }
        }
        if (P.accum) {   // temporal accumulation (extension): running sum, the mean is stored
            float4 a = P.accum[pi];
            a.x = fadd(a.x, light.x); a.y = fadd(a.y, light.y); a.z = fadd(a.z, light.z); a.w = fadd(a.w, 1.0f);
            P.accum[pi] = a;
            light = f3(fdiv(a.x, a.w), fdiv(a.y, a.w), fdiv(a.z, a.w));
        }
        // G-buffer stores, :352-385 (formats render_data.rs:166-189)
        P.lighting[pi] = make_ushort4((unsigned short)unorm_store(fmul(light.x, 0.0625f), 65535.0f),
                                      (unsigned short)unorm_store(fmul(light.y, 0.0625f), 65535.0f),
                                      (unsigned short)unorm_store(fmul(light.z, 0.0625f), 65535.0f),
                                      (unsigned short)unorm_store(0.0625f, 65535.0f));
        uint32_t d = 0xFFFF;
        if (!primary.air) d = __float2uint_rz(fmul(length3(sub3(origin, primary.position)), 32.0f));
        P.depth[pi] = (uint16_t)d;
        P.normal[pi] = primary.air ? 16 : (uint8_t)primary.normal;
        const F3 alb = primary.air ? f3(1, 1, 1) : primary.albedo;
        P.albedo[pi] = make_uchar4((unsigned char)unorm_store(alb.x, 255.0f), (unsigned char)unorm_store(alb.y, 255.0f),
                                   (unsigned char)unorm_store(alb.z, 255.0f), 255);
        P.emission[pi] = make_uchar4(0, 0, 0, primary.air ? 0 : 255);
        const F3 fogc = sample_sky(ray_dir, sunangle, sunlight, false);
        P.fog[pi] = make_uchar4((unsigned char)unorm_store(fmul(fogc.x, 0.5f), 255.0f), (unsigned char)unorm_store(fmul(fogc.y, 0.5f), 255.0f),
                                (unsigned char)unorm_store(fmul(fogc.z, 0.5f), 255.0f), 255);
        if (P.hit_pos)                                                                       // record for k_reproject (temporal.cu)
            P.hit_pos[pi] = primary.air ? make_float4(0.0f, 0.0f, 0.0f, __int_as_float(16))
                                        : make_float4(primary.position.x, primary.position.y, primary.position.z, __int_as_float((int)primary.normal));
        if (P.hit_voxel) {
            P.hit_voxel[pi * 3] = primary.vx; P.hit_voxel[pi * 3 + 1] = primary.vy; P.hit_voxel[pi * 3 + 2] = primary.vz;
            P.hit_face[pi] = primary.air ? 16 : (uint8_t)primary.normal;
            P.hit_word[pi] = primary.air ? 0u : primary.word;
            P.hit_t[pi] = primary.distance;
            P.steps[pi] = st.steps;
            P.rays[pi] = (uint8_t)st.rays;
        }
    }
    // frame counters: one atomic per warp and counter
    if (py < P.own_row0 || py >= P.own_row1) { st.rays = st.steps = st.hits = st.limit = 0; noise_fetches = 0; }   // halo rows: traced, not counted
    uint32_t v[5] = {st.rays, st.steps, st.hits, noise_fetches, st.limit};
#pragma unroll
    for (int k = 0; k < 5; k++) {
        uint32_t s = __reduce_add_sync(0xffffffffu, v[k]);
        if (lane == 0 && s) atomicAdd(P.counters + k, (unsigned long long)s);
    }
}

}  // namespace

// see peer.cu: every kernel is loaded when a handle is created (a lazy load cannot complete under a spinning kernel)
#define VXRT_PRELOAD(k) do { cudaFuncAttributes a_; cudaError_t e_ = cudaFuncGetAttributes(&a_, k); if (e_ != cudaSuccess) return e_; } while (0)
cudaError_t vxrt_preload_trace() { VXRT_PRELOAD(k_trace_general); return cudaSuccess; }

const void* vxrt_trace_kernel(const TraceParams& p, dim3* grid, dim3* block, size_t* smem) {
    *smem = 0;
    if (p.fast) {
        if (const void* k = vxrt_trace_pool_kernel(p, grid, block, smem)) return k;     // FULL lighting: rays pooled per warp (trace_pool.cu)
        *smem = 0;
        return vxrt_trace_fast_kernel(p, grid, block);
    }
    *grid = dim3((p.width + 15) / 16, (p.row1 - p.row0 + 7) / 8);
    *block = dim3(128);
    return (const void*)k_trace_general;
}

cudaError_t vxrt_launch_trace(const TraceParams& p, cudaStream_t s) {
    if (p.row1 - p.row0 <= 0) return cudaSuccess;
    if (vxrt_trace_wave_applies(p)) return vxrt_launch_trace_wave(p, s);                // FULL lighting as a wavefront (trace_wave.cu)
    dim3 grid, block;
    size_t smem = 0;
    const void* k = vxrt_trace_kernel(p, &grid, &block, &smem);
    void* args[1] = {(void*)&p};
    return cudaLaunchKernel(k, grid, block, args, smem, s);
}
