/* orc_math.h — scalar fp32 GLSL built-ins as the oracle defines them.
 * TEST INFRASTRUCTURE ONLY (see vxrt_oracle.h). */
#ifndef ORC_MATH_H
#define ORC_MATH_H
#include <cmath>
#include <cstdint>
namespace {

struct V3 { float x, y, z; };
static inline V3 v3(float x, float y, float z) { V3 r = {x, y, z}; return r; }
static inline V3 add(V3 a, V3 b) { return v3(a.x + b.x, a.y + b.y, a.z + b.z); }
static inline V3 sub(V3 a, V3 b) { return v3(a.x - b.x, a.y - b.y, a.z - b.z); }
static inline V3 mul(V3 a, V3 b) { return v3(a.x * b.x, a.y * b.y, a.z * b.z); }
static inline V3 muls(V3 a, float s) { return v3(a.x * s, a.y * s, a.z * s); }
/* dot summed x -> y -> z (Appendix A) */
static inline float dot3(V3 a, V3 b) { return (a.x * b.x + a.y * b.y) + a.z * b.z; }
static inline float length3(V3 a) { return sqrtf(dot3(a, a)); }
static inline V3 normalize3(V3 a) { float l = length3(a); return v3(a.x / l, a.y / l, a.z / l); }
/* GLSL mod(x,y) = x - y*floor(x/y) */
static inline float gmod(float x, float y) { return x - y * floorf(x / y); }
static inline float mixf(float a, float b, float t) { return a * (1.0f - t) + b * t; }
static inline V3 mix3(V3 a, V3 b, float t) { return v3(mixf(a.x, b.x, t), mixf(a.y, b.y, t), mixf(a.z, b.z, t)); }
/* clamp as min(max(x,lo),hi); NaN -> lo (fmaxf returns the non-NaN operand) */
static inline float clampf(float x, float lo, float hi) { return fminf(fmaxf(x, lo), hi); }

/* float -> int used for "floor(position)" in the hit record; NaN/inf/out of range -> INT32_MIN */
static inline int32_t floor_to_i32(float f) {
    float fl = floorf(f);
    if (!(fl >= -2147483648.0f && fl < 2147483648.0f)) return INT32_MIN;
    return (int32_t)fl;
}
/* GLSL uint(f): truncation; defined here for NaN/negative -> 0, >= 2^32 -> 0xFFFFFFFF */
static inline uint32_t f2u_trunc(float f) {
    if (!(f > 0.0f)) return 0u;
    if (f >= 4294967296.0f) return 0xFFFFFFFFu;
    return (uint32_t)f;
}
/* UNORMn store: rint(clamp(x,0,1)*(2^n-1)), round half to even (Appendix A) */
static inline uint32_t unorm_store(float x, float maxv) { return (uint32_t)nearbyintf(clampf(x, 0.0f, 1.0f) * maxv); }

}  // namespace
#endif
