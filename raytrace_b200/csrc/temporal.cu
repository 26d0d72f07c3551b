// temporal.cu — k_reproject: temporal reprojection of the lighting image under camera motion (extension, SURVEY §8f
// row 4).  The reference fills the inputs of such a pass every frame and never reads them: old_origin and
// old_transform_c0..2 (src/render/pipeline/pipeline.rs:213-227; uniform block shaders/glsl/raytrace.comp:31),
// MAX_SAMPLES = 8 (raytrace.comp:58, finalize.comp:18) and the bound, never written completed_buffer
// (raytrace.comp:19).  The rule (include/vxrt.h, VXRT_FLAG_REPROJECT) is this repo's; the oracle states it in
// oracle/vxrt_oracle_temporal.cpp and this kernel issues the same fp32 operations in the same order with explicit
// round-to-nearest intrinsics, so history, counts and images match it bit for bit on equal inputs.
//
// One thread per pixel, runs between the trace pass and the first denoise pass.  HBM-bound: per pixel it reads the
// lighting texel (8 B) and its hit record (16 B), for a hit whose reprojection lands inside the frame the previous
// frame's record and history at that pixel (16 + 16 B, gathered: neighbouring pixels land on neighbouring texels
// unless the camera rolls), and writes history (16 B), lighting (8 B) and completed (8 B): 88 B per pixel.
#include "trace_common.cuh"

namespace {

__global__ void __launch_bounds__(256) k_reproject(const __grid_constant__ ReprojectParams P) {
    const int px = blockIdx.x * 32 + (threadIdx.x & 31), py = blockIdx.y * 8 + (threadIdx.x >> 5);
    if (px >= P.width || py >= P.height) return;
    const size_t pi = (size_t)py * P.width + px;
    const ushort4 lt = P.lighting[pi];
    const F3 L = f3(fmul(fdiv((float)lt.x, 65535.0f), 16.0f), fmul(fdiv((float)lt.y, 65535.0f), 16.0f), fmul(fdiv((float)lt.z, 65535.0f), 16.0f));
    const float4 rec = __ldg(P.hit_pos_cur + pi);
    const int face = __float_as_int(rec.w);
    float n = 0.0f;
    F3 Hm = f3(0, 0, 0);
    if (face < 16) {
        const F3 r = sub3(f3(rec.x, rec.y, rec.z), f3(P.oox, P.ooy, P.ooz));
        const float vx = fadd(fadd(fmul(P.c0x, r.x), fmul(P.c1x, r.y)), fmul(P.c2x, r.z));
        const float vy = fadd(fadd(fmul(P.c0y, r.x), fmul(P.c1y, r.y)), fmul(P.c2y, r.z));
        const float vz = fadd(fadd(fmul(P.c0z, r.x), fmul(P.c1z, r.y)), fmul(P.c2z, r.z));
        if (vz > 0.0f) {
            const float fw = (float)P.width, fh = (float)P.height;
            const float fx = fadd(fmul(fmul(fadd(fdiv(vx, vz), 1.0f), 0.5f), fw), 0.5f);
            const float fy = fadd(fmul(fmul(fadd(fdiv(vy, vz), 1.0f), 0.5f), fh), 0.5f);
            if (fx >= 0.0f && fx < fw && fy >= 0.0f && fy < fh) {          // NaN fails
                const size_t qi = (size_t)__float2int_rd(fy) * P.width + (size_t)__float2int_rd(fx);
                const float4 prev = __ldg(P.hit_pos_prev + qi);
                if (__float_as_int(prev.w) == face && floorf(prev.x) == floorf(rec.x) && floorf(prev.y) == floorf(rec.y) &&
                    floorf(prev.z) == floorf(rec.z)) {
                    const float4 h = __ldg(P.hist_prev + qi);
                    n = fminf(h.w, 7.0f);                                    // MAX_SAMPLES - 1
                    Hm = f3(h.x, h.y, h.z);
                }
            }
        }
    }
    const float n1 = fadd(n, 1.0f);
    const F3 mean = f3(fdiv(fadd(fmul(Hm.x, n), L.x), n1), fdiv(fadd(fmul(Hm.y, n), L.y), n1), fdiv(fadd(fmul(Hm.z, n), L.z), n1));
    P.hist_cur[pi] = make_float4(mean.x, mean.y, mean.z, n1);
    const ushort4 q = make_ushort4((unsigned short)unorm_store(fmul(mean.x, 0.0625f), 65535.0f), (unsigned short)unorm_store(fmul(mean.y, 0.0625f), 65535.0f),
                                   (unsigned short)unorm_store(fmul(mean.z, 0.0625f), 65535.0f), lt.w);
    P.lighting[pi] = q;
    P.completed[pi] = make_ushort4(q.x, q.y, q.z, (unsigned short)unorm_store(fmul(n1, 0.125f), 65535.0f));
}

}  // namespace

#define VXRT_PRELOAD(k) do { cudaFuncAttributes a_; cudaError_t e_ = cudaFuncGetAttributes(&a_, k); if (e_ != cudaSuccess) return e_; } while (0)
cudaError_t vxrt_preload_temporal() { VXRT_PRELOAD(k_reproject); return cudaSuccess; }

cudaError_t vxrt_launch_reproject(const ReprojectParams& p, cudaStream_t s) {
    dim3 grid((p.width + 31) / 32, (p.height + 7) / 8);
    k_reproject<<<grid, 256, 0, s>>>(p);
    return cudaGetLastError();
}
