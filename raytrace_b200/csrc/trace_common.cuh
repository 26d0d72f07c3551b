// trace_common.cuh — fp32 helpers shared by the two forms of the trace kernel (trace.cu: general,
// trace_fast.cu: instruction-lean).  Every rounding of the reference expression tree is an explicit
// round-to-nearest intrinsic, which nvcc never contracts into an FMA.
#pragma once
#include "vxrt_internal.h"
#include <math_constants.h>

namespace {

struct F3 { float x, y, z; };
__device__ __forceinline__ F3 f3(float x, float y, float z) { F3 r; r.x = x; r.y = y; r.z = z; return r; }
__device__ __forceinline__ float fadd(float a, float b) { return __fadd_rn(a, b); }
__device__ __forceinline__ float fsub(float a, float b) { return __fsub_rn(a, b); }
__device__ __forceinline__ float fmul(float a, float b) { return __fmul_rn(a, b); }
__device__ __forceinline__ float fdiv(float a, float b) { return __fdiv_rn(a, b); }
__device__ __forceinline__ F3 add3(F3 a, F3 b) { return f3(fadd(a.x, b.x), fadd(a.y, b.y), fadd(a.z, b.z)); }
__device__ __forceinline__ F3 sub3(F3 a, F3 b) { return f3(fsub(a.x, b.x), fsub(a.y, b.y), fsub(a.z, b.z)); }
__device__ __forceinline__ F3 mul3(F3 a, F3 b) { return f3(fmul(a.x, b.x), fmul(a.y, b.y), fmul(a.z, b.z)); }
__device__ __forceinline__ F3 scale3(F3 a, float s) { return f3(fmul(a.x, s), fmul(a.y, s), fmul(a.z, s)); }
// dot summed x -> y -> z
__device__ __forceinline__ float dot3(F3 a, F3 b) { return fadd(fadd(fmul(a.x, b.x), fmul(a.y, b.y)), fmul(a.z, b.z)); }
__device__ __forceinline__ float length3(F3 a) { return __fsqrt_rn(dot3(a, a)); }
__device__ __forceinline__ F3 normalize3(F3 a) { float l = length3(a); return f3(fdiv(a.x, l), fdiv(a.y, l), fdiv(a.z, l)); }
__device__ __forceinline__ float mixf(float a, float b, float t) { return fadd(fmul(a, fsub(1.0f, t)), fmul(b, t)); }
__device__ __forceinline__ F3 mix3(F3 a, F3 b, float t) { return f3(mixf(a.x, b.x, t), mixf(a.y, b.y, t), mixf(a.z, b.z, t)); }
__device__ __forceinline__ float clampf(float x, float lo, float hi) { return fminf(fmaxf(x, lo), hi); }
__device__ __forceinline__ uint32_t unorm_store(float x, float maxv) { return __float2uint_rn(fmul(clampf(x, 0.0f, 1.0f), maxv)); }
__device__ __forceinline__ int32_t floor_to_i32(float f) {
    float fl = floorf(f);
    if (!(fl >= -2147483648.0f && fl < 2147483648.0f)) return INT32_MIN;
    return (int32_t)fl;
}

struct Hit { F3 albedo; F3 position; float distance; uint32_t normal, word, steps; int32_t vx, vy, vz; bool air; };

struct RayStats { uint32_t rays, steps, hits, limit; };

// raytrace.comp:189-212 through the 256-entry tables (noise channels are 8-bit)
__device__ __forceinline__ F3 diffuse_direction(const TraceParams& P, uint32_t nr, uint32_t ng, uint32_t normal) {
    const float s1 = __ldg(P.lut + nr), c1 = __ldg(P.lut + 256 + nr), s2 = __ldg(P.lut + 512 + ng), c2 = __ldg(P.lut + 768 + ng);
    F3 d = f3(fmul(s1, s2), fmul(c1, s2), c2);
    switch (normal) {
        case 0: d.x = fadd(d.x, 1.0f); break; case 1: d.x = fsub(d.x, 1.0f); break;
        case 2: d.y = fadd(d.y, 1.0f); break; case 3: d.y = fsub(d.y, 1.0f); break;
        case 4: d.z = fadd(d.z, 1.0f); break; case 5: d.z = fsub(d.z, 1.0f); break;
    }
    return normalize3(d);
}

// raytrace.comp:271-288 (colour only: tolerance-compared; powf is CUDA's accurate powf)
__device__ __noinline__ F3 sample_sky(F3 d, F3 sd, F3 sunlight, bool include_sun) {
    const F3 bright = f3(0.5294f, 0.8275f, 0.9647f), dark = f3(0.0863f, 0.1294f, 0.2196f);
    float sl = clampf(fsub(fmul(fadd(fadd(sunlight.x, sunlight.y), sunlight.z), 0.2f), 0.02f), 0.0f, 1.0f);
    float horizon = powf(__fsqrt_rn(fadd(fmul(d.x, d.x), fmul(d.y, d.y))), mixf(40.0f, 10.0f, sl));
    float sun_amount = fsub(1.0f, fmul(0.5f, length3(sub3(sd, d))));
    float halo = powf(sun_amount, mixf(5.0f, 1.0f, sl));
    float bright_amount = fminf(fadd(horizon, fmul(halo, 0.5f)), 1.0f);
    F3 color = mix3(dark, bright, fmul(bright_amount, fmaxf(sl, 0.1f)));
    float p5 = fmul(powf(sun_amount, 5.0f), 0.5f);
    color = add3(color, scale3(sunlight, p5));
    if (sun_amount > 0.98f && include_sun) color = add3(color, sunlight);
    return color;
}

__device__ __forceinline__ uchar4 noise_texel(const uint8_t* bn, float cx, float cy) {
    int ix = (int)floorf(cx), iy = (int)floorf(cy);
    ix = min(max(ix, 0), VXRT_NOISE_SIZE - 1); iy = min(max(iy, 0), VXRT_NOISE_SIZE - 1);
    return __ldg(reinterpret_cast<const uchar4*>(bn) + iy * VXRT_NOISE_SIZE + ix);
}
// (E11) x / 2^k and x * 2^-k are the correctly rounded value of the same real number (denormal results included), so a
// division by 2, 16 or 512 is written as the multiplication: one FMUL instead of the ~12-instruction IEEE division.
__device__ __forceinline__ float gmod512(float x) { return fsub(x, fmul(512.0f, floorf(fmul(x, 0.001953125f)))); }

}  // namespace
