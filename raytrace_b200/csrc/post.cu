// post.cu — k_denoise (shaders/glsl/bilateral_denoise.comp) and k_finalize (shaders/glsl/finalize.comp).
// Same explicit-rounding rule as trace.cu so the 16-bit / 8-bit images match the oracle bit for bit.
#include "vxrt_internal.h"

namespace {

__device__ __forceinline__ float fadd(float a, float b) { return __fadd_rn(a, b); }
__device__ __forceinline__ float fsub(float a, float b) { return __fsub_rn(a, b); }
__device__ __forceinline__ float fmul(float a, float b) { return __fmul_rn(a, b); }
__device__ __forceinline__ float fdiv(float a, float b) { return __fdiv_rn(a, b); }
__device__ __forceinline__ uint32_t unorm_store(float x, float maxv) {
    return __float2uint_rn(fmul(fminf(fmaxf(x, 0.0f), 1.0f), maxv));
}

// bilateral_denoise.comp:40-86, source order
__constant__ int8_t c_tap_dx[36] = {0, 0, 1, -1, 1, -1, -1, 1, 2, -2, 0, 0, 2, -2, -2, 2, 2, -2, -2, 2, 1, -1, -1, 1,
                                    3, -3, 0, 0, 3, -3, -3, 3, 1, -1, -1, 1};
__constant__ int8_t c_tap_dy[36] = {1, -1, 0, 0, 1, 1, -1, -1, 0, 0, 2, -2, 2, 2, -2, -2, 1, 1, -1, -1, 2, 2, -2, -2,
                                    0, 0, 3, -3, 1, 1, -1, -1, 3, 3, -3, -3};
__constant__ float c_tap_w[36] = {0.092566f, 0.092566f, 0.092566f, 0.092566f, 0.058434f, 0.058434f, 0.058434f, 0.058434f,
                                  0.023205f, 0.023205f, 0.023205f, 0.023205f, 0.003672f, 0.003672f, 0.003672f, 0.003672f,
                                  0.014648f, 0.014648f, 0.014648f, 0.014648f, 0.014648f, 0.014648f, 0.014648f, 0.014648f,
                                  0.002289f, 0.002289f, 0.002289f, 0.002289f, 0.001445f, 0.001445f, 0.001445f, 0.001445f,
                                  0.001445f, 0.001445f, 0.001445f, 0.001445f};

// bilateral_denoise.comp:35-89.  SWAP reproduces the pong descriptor set (descriptor_sets.rs:37-39).
template <bool SWAP>
__global__ void __launch_bounds__(256) k_denoise(const ushort4* __restrict__ src, ushort4* __restrict__ dst,
                                                 const uint16_t* __restrict__ depth, const uint8_t* __restrict__ normal,
                                                 int width, int height, int size, int row0, int row1) {
    const int x = blockIdx.x * 32 + (threadIdx.x & 31), y = row0 + blockIdx.y * 8 + (threadIdx.x >> 5);
    if (x >= width || y >= row1) return;
    const size_t pi = (size_t)y * width + x;
    const uint32_t dv = SWAP ? (uint32_t)normal[pi] : (uint32_t)depth[pi];
    const uint32_t center_normal = SWAP ? (uint32_t)depth[pi] : (uint32_t)normal[pi];
    const ushort4 c = src[pi];
    if (center_normal >= 16) { dst[pi] = c; return; }                          // :88-89
    const float center_distance = fdiv((float)dv, 256.0f);                     // :37
    float total_weight = 0.146634f;                                            // :41
    float sr = fmul(fdiv((float)c.x, 65535.0f), total_weight), sg = fmul(fdiv((float)c.y, 65535.0f), total_weight),
          sb = fmul(fdiv((float)c.z, 65535.0f), total_weight);
#pragma unroll 4
    for (int t = 0; t < 36; t++) {                                             // SAMPLE, :23-33
        int qx = x + c_tap_dx[t] * size, qy = y + c_tap_dy[t] * size;
        qx = min(max(qx, 0), width - 1); qy = min(max(qy, 0), height - 1);     // sampleAt, :14-21
        const size_t qi = (size_t)qy * width + qx;
        const uint32_t qd = SWAP ? (uint32_t)normal[qi] : (uint32_t)depth[qi];
        const uint32_t qn = SWAP ? (uint32_t)depth[qi] : (uint32_t)normal[qi];
        const float dist = fdiv((float)qd, 256.0f);
        const float dd = fmul(4.0f, fabsf(fsub(center_distance, dist)));
        const float nd = qn == center_normal ? 0.0f : 10.0f;
        const float weight = fdiv(c_tap_w[t], fadd(fadd(dd, nd), 1.0f));
        total_weight = fadd(total_weight, weight);
        const ushort4 q = src[qi];
        sr = fadd(sr, fmul(fdiv((float)q.x, 65535.0f), weight));
        sg = fadd(sg, fmul(fdiv((float)q.y, 65535.0f), weight));
        sb = fadd(sb, fmul(fdiv((float)q.z, 65535.0f), weight));
    }
    dst[pi] = make_ushort4((unsigned short)unorm_store(fdiv(sr, total_weight), 65535.0f),
                           (unsigned short)unorm_store(fdiv(sg, total_weight), 65535.0f),
                           (unsigned short)unorm_store(fdiv(sb, total_weight), 65535.0f), 65535);   // :87
}

// finalize.comp:21-31
__device__ __forceinline__ float filmic_curve(float x) {
    if (x < 0.3f) return fmul(x, x);
    else if (x < 1.13333f) return fsub(fmul(x, 0.6f), 0.09f);
    else if (x < 2.5f) return fsub(1.0f, fmul(fmul(0.219512195116f, fsub(x, 2.5f)), fsub(x, 2.5f)));
    else return 1.0f;
}

// finalize.comp:33-64
__global__ void __launch_bounds__(256) k_finalize(const uchar4* __restrict__ albedo, const uchar4* __restrict__ emission,
                                                  const uchar4* __restrict__ fog, const ushort4* __restrict__ lighting,
                                                  const uint16_t* __restrict__ depth, const uint8_t* __restrict__ noise,
                                                  uchar4* __restrict__ out, int width, int height, int row0, int row1) {
    const int x = blockIdx.x * 32 + (threadIdx.x & 31), y = row0 + blockIdx.y * 8 + (threadIdx.x >> 5);
    if (x >= width || y >= row1) return;
    const size_t pi = (size_t)y * width + x;
    const uchar4 a = albedo[pi], e = emission[pi], f = fog[pi];
    const ushort4 l = lighting[pi];
    const uint32_t d = depth[pi];
    const uchar4 bn = reinterpret_cast<const uchar4*>(noise)[(y & 511) * VXRT_NOISE_SIZE + (x & 511)];   // :55-57
    float fog_amount = fdiv((float)d, 32.0f * 128.0f * 8.0f);                  // :47
    if (fog_amount > 1.0f) fog_amount = 1.0f;
    const float av[3] = {(float)a.x, (float)a.y, (float)a.z}, ev[3] = {(float)e.x, (float)e.y, (float)e.z},
                fv[3] = {(float)f.x, (float)f.y, (float)f.z}, lv[3] = {(float)l.x, (float)l.y, (float)l.z},
                nv[3] = {(float)bn.x, (float)bn.y, (float)bn.z};
    uint32_t o[3];
#pragma unroll
    for (int k = 0; k < 3; k++) {
        float v = fadd(fmul(fdiv(av[k], 255.0f), fmul(fdiv(lv[k], 65535.0f), 16.0f)), fmul(fdiv(ev[k], 255.0f), 4.0f));   // :36-40
        if (d < 0xFFFF) {                                                      // :44-50
            const float fg = fmul(fdiv(fv[k], 255.0f), 2.0f);
            v = fadd(fmul(v, fsub(1.0f, fog_amount)), fmul(fg, fog_amount));
        }
        v = filmic_curve(v);
        v = fadd(v, fdiv(fdiv(nv[k], 255.0f), 128.0f));                        // :58
        o[k] = unorm_store(v, 255.0f);
    }
    out[(size_t)(height - y - 1) * width + x] = make_uchar4((unsigned char)o[0], (unsigned char)o[1], (unsigned char)o[2], 255);   // :60-63
}

}  // namespace

cudaError_t vxrt_launch_denoise(const ushort4* src, ushort4* dst, const uint16_t* depth, const uint8_t* normal,
                                int width, int height, int size, int swap, int row0, int row1, cudaStream_t s) {
    if (row1 <= row0) return cudaSuccess;
    dim3 grid((width + 31) / 32, (row1 - row0 + 7) / 8);
    if (swap) k_denoise<true><<<grid, 256, 0, s>>>(src, dst, depth, normal, width, height, size, row0, row1);
    else k_denoise<false><<<grid, 256, 0, s>>>(src, dst, depth, normal, width, height, size, row0, row1);
    return cudaGetLastError();
}

cudaError_t vxrt_launch_finalize(const uchar4* albedo, const uchar4* emission, const uchar4* fog, const ushort4* lighting,
                                 const uint16_t* depth, const uint8_t* noise, uchar4* out, int width, int height,
                                 int row0, int row1, cudaStream_t s) {
    if (row1 <= row0) return cudaSuccess;
    dim3 grid((width + 31) / 32, (row1 - row0 + 7) / 8);
    k_finalize<<<grid, 256, 0, s>>>(albedo, emission, fog, lighting, depth, noise, out, width, height, row0, row1);
    return cudaGetLastError();
}
