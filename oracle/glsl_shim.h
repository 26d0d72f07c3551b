/*
 * glsl_shim.h — the subset of GLSL 450 that the reference's compute shaders use, as plain C++.
 *
 * TEST INFRASTRUCTURE ONLY.  oracle/make_ref.py wraps the reference's own shader text
 * (/root/reference/shaders/glsl/{raytrace,bilateral_denoise,finalize}.comp), where it lies, into a C++
 * struct that derives from nothing and only relies on the names below; the result (oracle/_ref/, git-
 * ignored, never committed) is the "reference run here" that pins the hand-written oracle.
 *
 * Built-ins follow SURVEY.md Appendix A exactly as oracle/orc_math.h does (fp32, one IEEE rounding per
 * operator, no contraction): mod(x,y) = x - y*floor(x/y); normalize(v) = v / sqrt(dot); dot summed
 * x -> y -> z; mix(a,b,t) = a*(1-t) + b*t; sin/cos/acos evaluated in double and rounded once; pow = powf;
 * UNORM store = rint(clamp(x,0,1)*(2^n-1)); samplers: nearest, texel = floor(coord), border -> 0 for the
 * 3-D integer textures, clamp-to-edge for blue noise (render_data.rs:54-133).
 */
#ifndef GLSL_SHIM_H
#define GLSL_SHIM_H
#include <cmath>
#include <cstdint>
#include <type_traits>

namespace glsl {

typedef unsigned int uint;
#define GLSL_ARITH(T) typename std::enable_if<std::is_arithmetic<T>::value, int>::type = 0

struct ivec2 {
    union { struct { int x, y; }; struct { int r, g; }; };
    ivec2() : x(0), y(0) {}
    template <class T, GLSL_ARITH(T)> explicit ivec2(T s) : x((int)s), y((int)s) {}
    template <class A, class B, GLSL_ARITH(A), GLSL_ARITH(B)> ivec2(A a, B b) : x((int)a), y((int)b) {}
    ivec2& operator+=(const ivec2& o) { x += o.x; y += o.y; return *this; }
    ivec2& operator*=(const ivec2& o) { x *= o.x; y *= o.y; return *this; }
};
typedef ivec2 uvec2;
inline ivec2 operator+(ivec2 a, ivec2 b) { return ivec2(a.x + b.x, a.y + b.y); }
inline ivec2 operator-(ivec2 a, ivec2 b) { return ivec2(a.x - b.x, a.y - b.y); }
inline ivec2 operator*(ivec2 a, ivec2 b) { return ivec2(a.x * b.x, a.y * b.y); }
inline ivec2 operator%(ivec2 a, ivec2 b) { return ivec2(a.x % b.x, a.y % b.y); }
template <class T, GLSL_ARITH(T)> inline ivec2 operator*(ivec2 a, T s) { return ivec2(a.x * (int)s, a.y * (int)s); }

struct ivec3 {
    union { struct { int x, y, z; }; struct { int r, g, b; }; };
    ivec3() : x(0), y(0), z(0) {}
    template <class T, GLSL_ARITH(T)> explicit ivec3(T s) : x((int)s), y((int)s), z((int)s) {}
    template <class A, class B, class C, GLSL_ARITH(A), GLSL_ARITH(B), GLSL_ARITH(C)>
    ivec3(A a, B b, C c) : x((int)a), y((int)b), z((int)c) {}
    ivec2 xy() const { return ivec2(x, y); }
};
typedef ivec3 uvec3;

struct uvec4 {
    union { struct { uint x, y, z, w; }; struct { uint r, g, b, a; }; };
    uvec4() : x(0), y(0), z(0), w(0) {}
    template <class T, GLSL_ARITH(T)> explicit uvec4(T s) : x((uint)s), y((uint)s), z((uint)s), w((uint)s) {}
    template <class A, class B, class C, class D, GLSL_ARITH(A), GLSL_ARITH(B), GLSL_ARITH(C), GLSL_ARITH(D)>
    uvec4(A a, B b, C c, D d) : x((uint)a), y((uint)b), z((uint)c), w((uint)d) {}
};

struct vec2 {
    union { struct { float x, y; }; struct { float r, g; }; };
    vec2() : x(0), y(0) {}
    template <class T, GLSL_ARITH(T)> explicit vec2(T s) : x((float)s), y((float)s) {}
    template <class A, class B, GLSL_ARITH(A), GLSL_ARITH(B)> vec2(A a, B b) : x((float)a), y((float)b) {}
    vec2(const ivec2& v) : x((float)v.x), y((float)v.y) {}            /* GLSL implicit int -> float vector */
    vec2& operator+=(const vec2& o) { x += o.x; y += o.y; return *this; }
};
inline vec2 operator+(vec2 a, vec2 b) { return vec2(a.x + b.x, a.y + b.y); }
inline vec2 operator-(vec2 a, vec2 b) { return vec2(a.x - b.x, a.y - b.y); }
inline vec2 operator*(vec2 a, vec2 b) { return vec2(a.x * b.x, a.y * b.y); }
inline vec2 operator/(vec2 a, vec2 b) { return vec2(a.x / b.x, a.y / b.y); }
template <class T, GLSL_ARITH(T)> inline vec2 operator*(vec2 a, T s) { return vec2(a.x * (float)s, a.y * (float)s); }
template <class T, GLSL_ARITH(T)> inline vec2 operator*(T s, vec2 a) { return vec2((float)s * a.x, (float)s * a.y); }
template <class T, GLSL_ARITH(T)> inline vec2 operator/(vec2 a, T s) { return vec2(a.x / (float)s, a.y / (float)s); }

struct vec3 {
    union { struct { float x, y, z; }; struct { float r, g, b; }; };
    vec3() : x(0), y(0), z(0) {}
    template <class T, GLSL_ARITH(T)> explicit vec3(T s) : x((float)s), y((float)s), z((float)s) {}
    template <class A, class B, class C, GLSL_ARITH(A), GLSL_ARITH(B), GLSL_ARITH(C)>
    vec3(A a, B b, C c) : x((float)a), y((float)b), z((float)c) {}
    template <class T, GLSL_ARITH(T)> vec3(const vec2& v, T c) : x(v.x), y(v.y), z((float)c) {}
    vec3(const ivec3& v) : x((float)v.x), y((float)v.y), z((float)v.z) {}
    vec2 xy() const { return vec2(x, y); }
    vec2 rg() const { return vec2(x, y); }
    vec3& operator+=(const vec3& o) { x += o.x; y += o.y; z += o.z; return *this; }
    vec3& operator-=(const vec3& o) { x -= o.x; y -= o.y; z -= o.z; return *this; }
    vec3& operator*=(const vec3& o) { x *= o.x; y *= o.y; z *= o.z; return *this; }
    template <class T, GLSL_ARITH(T)> vec3& operator*=(T s) { x *= (float)s; y *= (float)s; z *= (float)s; return *this; }
};
inline vec3 operator+(vec3 a, vec3 b) { return vec3(a.x + b.x, a.y + b.y, a.z + b.z); }
inline vec3 operator-(vec3 a, vec3 b) { return vec3(a.x - b.x, a.y - b.y, a.z - b.z); }
inline vec3 operator*(vec3 a, vec3 b) { return vec3(a.x * b.x, a.y * b.y, a.z * b.z); }
inline vec3 operator/(vec3 a, vec3 b) { return vec3(a.x / b.x, a.y / b.y, a.z / b.z); }
template <class T, GLSL_ARITH(T)> inline vec3 operator*(vec3 a, T s) { return vec3(a.x * (float)s, a.y * (float)s, a.z * (float)s); }
template <class T, GLSL_ARITH(T)> inline vec3 operator*(T s, vec3 a) { return vec3((float)s * a.x, (float)s * a.y, (float)s * a.z); }
template <class T, GLSL_ARITH(T)> inline vec3 operator/(vec3 a, T s) { return vec3(a.x / (float)s, a.y / (float)s, a.z / (float)s); }

struct vec4 {
    union { struct { float x, y, z, w; }; struct { float r, g, b, a; }; };
    vec4() : x(0), y(0), z(0), w(0) {}
    template <class T, GLSL_ARITH(T)> explicit vec4(T s) : x((float)s), y((float)s), z((float)s), w((float)s) {}
    template <class A, class B, class C, class D, GLSL_ARITH(A), GLSL_ARITH(B), GLSL_ARITH(C), GLSL_ARITH(D)>
    vec4(A a_, B b_, C c_, D d_) : x((float)a_), y((float)b_), z((float)c_), w((float)d_) {}
    template <class T, GLSL_ARITH(T)> vec4(const vec3& v, T d_) : x(v.x), y(v.y), z(v.z), w((float)d_) {}
    vec3 rgb() const { return vec3(x, y, z); }
    vec2 rg() const { return vec2(x, y); }
    vec2 xy() const { return vec2(x, y); }
};
template <class T, GLSL_ARITH(T)> inline vec4 operator/(vec4 a, T s) { return vec4(a.x / (float)s, a.y / (float)s, a.z / (float)s, a.w / (float)s); }
template <class T, GLSL_ARITH(T)> inline vec4 operator*(vec4 a, T s) { return vec4(a.x * (float)s, a.y * (float)s, a.z * (float)s, a.w * (float)s); }

/* ---- built-ins (Appendix A definitions) ------------------------------------------------------------- */
inline float floor(float x) { return ::floorf(x); }
inline float abs(float x) { return ::fabsf(x); }
inline int abs(int x) { return x < 0 ? -x : x; }
inline float sqrt(float x) { return ::sqrtf(x); }
inline float sin(float x) { return (float)::sin((double)x); }
inline float cos(float x) { return (float)::cos((double)x); }
inline float acos(float x) { return (float)::acos((double)x); }
inline float pow(float x, float y) { return ::powf(x, y); }
inline float min(float a, float b) { return ::fminf(a, b); }
inline float max(float a, float b) { return ::fmaxf(a, b); }
inline float clamp(float x, float lo, float hi) { return ::fminf(::fmaxf(x, lo), hi); }
inline float mix(float a, float b, float t) { return a * (1.0f - t) + b * t; }
inline vec3 mix(vec3 a, vec3 b, float t) { return vec3(mix(a.x, b.x, t), mix(a.y, b.y, t), mix(a.z, b.z, t)); }
inline float mod(float x, float y) { return x - y * ::floorf(x / y); }
inline vec2 mod(vec2 a, vec2 b) { return vec2(mod(a.x, b.x), mod(a.y, b.y)); }
inline vec3 mod(vec3 a, vec3 b) { return vec3(mod(a.x, b.x), mod(a.y, b.y), mod(a.z, b.z)); }
template <class T, GLSL_ARITH(T)> inline vec3 mod(vec3 a, T s) { return vec3(mod(a.x, (float)s), mod(a.y, (float)s), mod(a.z, (float)s)); }
inline vec3 abs(vec3 a) { return vec3(::fabsf(a.x), ::fabsf(a.y), ::fabsf(a.z)); }
inline float dot(vec3 a, vec3 b) { return (a.x * b.x + a.y * b.y) + a.z * b.z; }
inline float length(vec2 a) { return ::sqrtf(a.x * a.x + a.y * a.y); }
inline float length(vec3 a) { return ::sqrtf(dot(a, a)); }
inline vec3 normalize(vec3 a) { float l = length(a); return vec3(a.x / l, a.y / l, a.z / l); }
/* uint(f): truncation; NaN / negative -> 0, >= 2^32 -> 0xFFFFFFFF (same definition as orc_math.h f2u_trunc) */
inline uint to_uint(float f) { if (!(f > 0.0f)) return 0u; if (f >= 4294967296.0f) return 0xFFFFFFFFu; return (uint)f; }
inline uint to_uint(uint v) { return v; }
inline uint to_uint(int v) { return (uint)v; }

/* ---- samplers (render_data.rs:54-133) ----------------------------------------------------------------- */
struct usampler3D {           /* nearest, CLAMP_TO_BORDER, INT_OPAQUE_BLACK; R8_UINT or R32_UINT */
    const void* data = nullptr; int nx = 0, ny = 0, nz = 0; int bytes = 1; bool unnormalized = false;
};
inline bool texel_of(float c, int n, int* i) { if (!(c >= 0.0f && c < (float)n)) return false; *i = (int)::floorf(c); return true; }
inline uvec4 texture(const usampler3D& s, vec3 p) {
    float cx = p.x, cy = p.y, cz = p.z;
    if (!s.unnormalized) { cx = cx * (float)s.nx; cy = cy * (float)s.ny; cz = cz * (float)s.nz; }
    int ix, iy, iz;
    const bool ok = texel_of(cx, s.nx, &ix) & texel_of(cy, s.ny, &iy) & texel_of(cz, s.nz, &iz);
    if (!ok) return uvec4(0, 0, 0, 1);
    const size_t i = ((size_t)iz * s.ny + iy) * s.nx + ix;
    const uint v = s.bytes == 1 ? (uint)((const uint8_t*)s.data)[i] : ((const uint32_t*)s.data)[i];
    return uvec4(v, 0, 0, 1);
}
inline uvec4 textureLod(const usampler3D& s, vec3 p, float) { return texture(s, p); }

struct sampler2D {            /* blue noise: RGBA8_UNORM, nearest, unnormalised coords, CLAMP_TO_EDGE */
    const uint8_t* data = nullptr; int w = 0, h = 0;
};
inline vec4 texture(const sampler2D& s, vec2 p) {
    int ix = (int)::floorf(p.x), iy = (int)::floorf(p.y);
    if (ix < 0) ix = 0; if (ix > s.w - 1) ix = s.w - 1; if (iy < 0) iy = 0; if (iy > s.h - 1) iy = s.h - 1;
    const uint8_t* t = s.data + ((size_t)iy * s.w + ix) * 4;
    return vec4((float)t[0] / 255.0f, (float)t[1] / 255.0f, (float)t[2] / 255.0f, (float)t[3] / 255.0f);
}

/* ---- storage images (formats render_data.rs:166-189) ------------------------------------------------------ */
enum ImageFormat { FMT_none = 0, FMT_rgba8, FMT_rgba16, FMT_r8ui, FMT_r16ui };
struct image2D {              /* UNORM formats */
    ImageFormat fmt; void* data = nullptr; int w = 0, h = 0;
    explicit image2D(ImageFormat f) : fmt(f) {}
};
struct uimage2D {             /* UINT formats */
    ImageFormat fmt; void* data = nullptr; int w = 0, h = 0;
    explicit uimage2D(ImageFormat f) : fmt(f) {}
};
inline ivec2 imageSize(const image2D& im) { return ivec2(im.w, im.h); }
inline ivec2 imageSize(const uimage2D& im) { return ivec2(im.w, im.h); }
inline uint unorm_enc(float x, float maxv) { return (uint)::nearbyintf(clamp(x, 0.0f, 1.0f) * maxv); }
inline void imageStore(image2D& im, ivec2 p, vec4 v) {
    if (!im.data || p.x < 0 || p.y < 0 || p.x >= im.w || p.y >= im.h) return;     /* out-of-range stores are discarded */
    const size_t i = ((size_t)p.y * im.w + p.x) * 4;
    if (im.fmt == FMT_rgba8) {
        uint8_t* d = (uint8_t*)im.data + i;
        d[0] = (uint8_t)unorm_enc(v.x, 255.0f); d[1] = (uint8_t)unorm_enc(v.y, 255.0f);
        d[2] = (uint8_t)unorm_enc(v.z, 255.0f); d[3] = (uint8_t)unorm_enc(v.w, 255.0f);
    } else {
        uint16_t* d = (uint16_t*)im.data + i;
        d[0] = (uint16_t)unorm_enc(v.x, 65535.0f); d[1] = (uint16_t)unorm_enc(v.y, 65535.0f);
        d[2] = (uint16_t)unorm_enc(v.z, 65535.0f); d[3] = (uint16_t)unorm_enc(v.w, 65535.0f);
    }
}
inline vec4 imageLoad(const image2D& im, ivec2 p) {
    if (!im.data || p.x < 0 || p.y < 0 || p.x >= im.w || p.y >= im.h) return vec4(0.0f);
    const size_t i = ((size_t)p.y * im.w + p.x) * 4;
    if (im.fmt == FMT_rgba8) {
        const uint8_t* d = (const uint8_t*)im.data + i;
        return vec4((float)d[0] / 255.0f, (float)d[1] / 255.0f, (float)d[2] / 255.0f, (float)d[3] / 255.0f);
    }
    const uint16_t* d = (const uint16_t*)im.data + i;
    return vec4((float)d[0] / 65535.0f, (float)d[1] / 65535.0f, (float)d[2] / 65535.0f, (float)d[3] / 65535.0f);
}
inline void imageStore(uimage2D& im, ivec2 p, uvec4 v) {
    if (!im.data || p.x < 0 || p.y < 0 || p.x >= im.w || p.y >= im.h) return;
    const size_t i = (size_t)p.y * im.w + p.x;
    if (im.fmt == FMT_r8ui) ((uint8_t*)im.data)[i] = (uint8_t)v.x; else ((uint16_t*)im.data)[i] = (uint16_t)v.x;
}
inline uvec4 imageLoad(const uimage2D& im, ivec2 p) {
    if (!im.data || p.x < 0 || p.y < 0 || p.x >= im.w || p.y >= im.h) return uvec4(0);
    const size_t i = (size_t)p.y * im.w + p.x;
    return uvec4(im.fmt == FMT_r8ui ? (uint)((const uint8_t*)im.data)[i] : (uint)((const uint16_t*)im.data)[i], 0, 0, 0);
}

/* compute-shader built-in variables: members of every wrapped shader struct */
struct ComputeBuiltins {
    uvec3 gl_WorkGroupID, gl_LocalInvocationID, gl_GlobalInvocationID;
    uvec3 gl_WorkGroupSize = uvec3(8, 8, 1);
};

}  // namespace glsl
#endif
