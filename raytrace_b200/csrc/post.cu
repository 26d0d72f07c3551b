// post.cu — k_denoise (shaders/glsl/bilateral_denoise.comp) and k_finalize (shaders/glsl/finalize.comp).
// Same explicit-rounding rule as trace.cu so the 16-bit / 8-bit images match the oracle bit for bit.
#include "vxrt_internal.h"
#include "peer_dev.cuh"
#include <string.h>

namespace {

__device__ __forceinline__ float fadd(float a, float b) { return __fadd_rn(a, b); }
__device__ __forceinline__ float fsub(float a, float b) { return __fsub_rn(a, b); }
__device__ __forceinline__ float fmul(float a, float b) { return __fmul_rn(a, b); }
__device__ __forceinline__ float fdiv(float a, float b) { return __fdiv_rn(a, b); }
__device__ __forceinline__ uint32_t unorm_store(float x, float maxv) {
    return __float2uint_rn(fmul(fminf(fmaxf(x, 0.0f), 1.0f), maxv));
}

// bilateral_denoise.comp:35-89 as a shared-memory tiled, register-sliding kernel.  SWAP reproduces the pong
// descriptor set (descriptor_sets.rs:37-39: odd passes read the R8 normal image as "depth" and the R16 depth
// image as "normal").
//
// The filter is colour work: its contract is the stated tolerance (<= 1/255 per 8-bit channel after
// finalize), not bit equality, so this kernel uses FMA contraction, a reciprocal-multiply for the UNORM
// decode, MUFU.RCP for the 36 weight divisions and its own summation order.  The 16-bit quantisation point
// after every pass is kept.
//
// Tiling: a pass with dilation S only couples pixels whose y differ by multiples of S, so
// a CTA owns 128 contiguous x (coalesced rows) x L rows of ONE y-residue class and stages
// (128 + 6S) x (L + 6) texels in shared memory, decoded once to {r,g,b,dist} floats + the normal code.
// A thread computes G = 4 vertically adjacent outputs of its column at once: the 10 staged rows they touch are
// loaded ONCE into registers (58 texels for 4 x 37 taps, i.e. 14.5 shared-memory texel loads per output
// instead of 37), each texel feeding every output it is a tap of.  The earlier one-output-per-thread form
// was bound by shared-memory bandwidth (LDS.128 = 4 wavefronts per tap); this one by issue / MUFU.RCP.
// MUFU.RCP alone (1 ulp-ish): __frcp_rn expands to a 12-instruction IEEE sequence with a slow-path call
__device__ __forceinline__ float rcp_fast(float x) { float r; asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x)); return r; }
// 16-bit shared load into a 32-bit register: keeps the normal-code comparison a plain 32-bit ISETP (the
// compiler otherwise narrows it to 16 bits and pays two PRMTs per tap)
__device__ __forceinline__ uint32_t lds_u16(const uint16_t* p) {
    uint32_t r;
    asm volatile("ld.shared.u16 %0, [%1];" : "=r"(r) : "r"((uint32_t)__cvta_generic_to_shared(p)));
    return r;
}
// Packed fp32 accumulation (sm_100a FFMA2 = fma.rn.f32x2, one issue slot for two lanes): VXRT_DN_PACK 1 packs the
// (r, g) sums (1976 -> 1840 SASS instructions per four outputs); 3 also packs (b, weight sum) against texels staged as
// {r, g, b, 1.0} + {dist, normal} (1664 instructions, 24 B of shared memory per texel instead of 18); 2 builds the
// {b, 1.0} pair in registers (the extra MOVs cancel the gain).  Measured: the passes are bound by their
// tails, not by issue slots, so all of them are within 1 % of each other in time.
#ifndef VXRT_DN_PACK
#define VXRT_DN_PACK 1
#endif
typedef unsigned long long dn_f2;
__device__ __forceinline__ dn_f2 dn_pk(float lo, float hi) { dn_f2 r; asm("mov.b64 %0, {%1,%2};" : "=l"(r) : "f"(lo), "f"(hi)); return r; }
__device__ __forceinline__ void dn_unpk(dn_f2 v, float& lo, float& hi) { asm("mov.b64 {%0,%1}, %2;" : "=f"(lo), "=f"(hi) : "l"(v)); }
__device__ __forceinline__ dn_f2 dn_fma2(dn_f2 a, dn_f2 b, dn_f2 c) { dn_f2 r; asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r) : "l"(a), "l"(b), "l"(c)); return r; }
constexpr int DN_TX = 128;                     // pixels per CTA row
#ifndef VXRT_DN_TY
#define VXRT_DN_TY 2
#endif
constexpr int DN_TY = VXRT_DN_TY;              // thread rows per CTA
#ifndef VXRT_DN_MINB
#define VXRT_DN_MINB 5                         // 48 registers: 5 CTAs/SM where shared memory allows (S <= 8)
#endif
constexpr int DN_G = 4;                        // vertically adjacent outputs per thread
// decimated rows a thread walks per tile: dilations 1,2,4 / dilations 8,16 (wider x halo -> shorter tile)
#ifndef VXRT_DN_UNROLL
#define VXRT_DN_UNROLL 2
#endif
constexpr int DN_UNROLL = VXRT_DN_UNROLL;      // staging loop unroll (global loads in flight per thread)
#ifndef VXRT_DN_L0
#define VXRT_DN_L0 8
#endif
#ifndef VXRT_DN_L1
#define VXRT_DN_L1 8
#endif
template <int S, int L> struct DnTile {
    static constexpr int W = DN_TX + 6 * S, R = L + 6, N = W * R;
    // PACK 3: texel = {r, g, b, 1.0} + {dist, normal code} (24 B) so (b, 1.0) is a register pair straight from LDS.128
    static constexpr size_t BYTES = (size_t)N * (sizeof(float4) + (VXRT_DN_PACK == 3 ? sizeof(float2) : sizeof(uint16_t)));
};
// SAMPLE list of bilateral_denoise.comp:43-86 as a function of (|dx|, |dy|); 0 = not a tap
__host__ __device__ constexpr float dn_weight(int ax, int ay) {
    const int lo = ax < ay ? ax : ay, hi = ax < ay ? ay : ax;
    return hi == 0 ? 0.146634f
         : hi == 1 ? (lo == 0 ? 0.092566f : 0.058434f)
         : hi == 2 ? (lo == 0 ? 0.023205f : lo == 1 ? 0.014648f : 0.003672f)
         : (lo == 0 ? 0.002289f : lo == 1 ? 0.001445f : 0.0f);
}

// finalize.comp:21-31
__device__ __forceinline__ float filmic_curve(float x) {
    if (x < 0.3f) return x * x;
    else if (x < 1.13333f) return fmaf(x, 0.6f, -0.09f);
    else if (x < 2.5f) return fmaf(-0.219512195116f * (x - 2.5f), x - 2.5f, 1.0f);
    else return 1.0f;
}

// finalize.comp:33-64 for one pixel whose lighting texel is already in registers.  Colour work under the tolerance contract
// (<= 1/255 per channel): UNORM decodes are reciprocal multiplies folded into the constants, FMA contraction allowed.
struct FinalizeImages { const uchar4* albedo; const uchar4* emission; const uchar4* fog; const uint16_t* depth; const uint8_t* noise; uchar4* out; int row0, row1; };
__device__ __forceinline__ void finalize_pixel(const FinalizeImages& F, int width, int height, int x, int y, ushort4 l) {
    const size_t pi = (size_t)y * width + x;
    const uchar4 a = __ldg(F.albedo + pi), e = __ldg(F.emission + pi), f = __ldg(F.fog + pi);
    const uint32_t d = __ldg(F.depth + pi);
    const uchar4 bn = __ldg(reinterpret_cast<const uchar4*>(F.noise) + (y & 511) * VXRT_NOISE_SIZE + (x & 511));   // :55-57
    const float fog_amount = fminf((float)d * (1.0f / (32.0f * 128.0f * 8.0f)), 1.0f);   // :47-48
    const float av[3] = {(float)a.x, (float)a.y, (float)a.z}, ev[3] = {(float)e.x, (float)e.y, (float)e.z},
                fv[3] = {(float)f.x, (float)f.y, (float)f.z}, lv[3] = {(float)l.x, (float)l.y, (float)l.z},
                nv[3] = {(float)bn.x, (float)bn.y, (float)bn.z};
    uint32_t o[3];
#pragma unroll
    for (int k = 0; k < 3; k++) {
        float v = fmaf(av[k] * lv[k], 16.0f / (255.0f * 65535.0f), ev[k] * (4.0f / 255.0f));   // :36-40
        if (d < 0xFFFF) v = fmaf(fv[k] * (2.0f / 255.0f) - v, fog_amount, v);                    // :44-50 mix(v, fog*2, amount)
        v = fmaf(nv[k], 1.0f / (255.0f * 128.0f), filmic_curve(v));                               // :52-58
        o[k] = __float2uint_rn(__saturatef(v) * 255.0f);
    }
    F.out[(size_t)(height - y - 1) * width + x] = make_uchar4((unsigned char)o[0], (unsigned char)o[1], (unsigned char)o[2], 255);   // :60-63
}

// FIN: the last pass also runs finalize.comp on every pixel it stores (pipeline.rs:117-123 folded into pipeline.rs:98-115's last
// dispatch): the lighting texel goes to the lighting image as before AND, from registers, through finalize_pixel - one launch,
// one grid tail and 8 B/pixel of reads less per frame.
template <int S, int L, bool SWAP, bool FIN>
__global__ void __launch_bounds__(DN_TX * DN_TY, VXRT_DN_MINB) k_denoise(const ushort4* __restrict__ src, ushort4* __restrict__ dst,
                                                   const uint16_t* __restrict__ depth, const uint8_t* __restrict__ normal,
                                                   int width, int height, int row0, int row1, int groups, const __grid_constant__ FinalizeImages F) {
    using T = DnTile<S, L>;
    auto put = [&](int y, size_t pi, ushort4 v) {
        dst[pi] = v;
        if (FIN && y >= F.row0 && y < F.row1) finalize_pixel(F, width, height, (int)(pi - (size_t)y * width), y, v);
    };
    extern __shared__ float4 dn_tile[];
    uint16_t* dn_nrm = reinterpret_cast<uint16_t*>(dn_tile + T::N);
    float2* dn_aux = reinterpret_cast<float2*>(dn_tile + T::N);       // VXRT_DN_PACK == 3
    const int tid = threadIdx.y * DN_TX + threadIdx.x;
    const int bx = blockIdx.x * DN_TX;
    const int res = blockIdx.y / groups, grp = blockIdx.y - res * groups;   // y-residue class, group of L decimated rows
    const int ybase = row0 + res + grp * L * S;                              // G-buffer row of decimated row 0
    const int x = bx + threadIdx.x;
    // Pixels whose normal code is >= 16 (sky) are plain copies (:88-89).  A tile with no other pixel skips the
    // staging and the taps altogether; the result is the same image.
    {
        bool work = false;
        if (x < width)
            for (int o = threadIdx.y * DN_G; o < L; o += DN_TY * DN_G)
#pragma unroll
                for (int k = 0; k < DN_G; k++) {
                    const int y = ybase + (o + k) * S;
                    if (y < row1) { const size_t pi = (size_t)y * width + x; work |= (SWAP ? (uint32_t)__ldg(depth + pi) : (uint32_t)__ldg(normal + pi)) < 16u; }
                }
        // Programmatic dependent launch (passes 2..6): this grid may become resident while the previous pass drains;
        // everything above reads only the trace pass's images.  From here on the previous pass is complete and visible,
        // and the next pass may start taking the slots this one frees.  Both are no-ops in a plain launch.
        asm volatile("griddepcontrol.wait;" ::: "memory");
        asm volatile("griddepcontrol.launch_dependents;");
        if (!__syncthreads_or(work)) {
            if (x < width)
                for (int o = threadIdx.y * DN_G; o < L; o += DN_TY * DN_G)
#pragma unroll
                    for (int k = 0; k < DN_G; k++) {
                        const int y = ybase + (o + k) * S;
                        if (y < row1) { const size_t pi = (size_t)y * width + x; put(y, pi, __ldg(src + pi)); }
                    }
            return;
        }
    }
    // stage: texel (tx, ty) <- image (clamp(bx + tx - 3S), clamp(ybase + (ty - 3) * S))   (sampleAt, :14-21)
#pragma unroll DN_UNROLL
    for (int e = tid; e < T::N; e += DN_TX * DN_TY) {
        const int ty = e / T::W, tx = e - ty * T::W;
        const int gx = min(max(bx + tx - 3 * S, 0), width - 1), gy = min(max(ybase + (ty - 3) * S, 0), height - 1);
        const size_t gi = (size_t)gy * width + gx;
        const ushort4 q = __ldg(src + gi);
        const uint32_t d = __ldg(depth + gi), n = __ldg(normal + gi);
        const float k = 1.0f / 65535.0f;
#if VXRT_DN_PACK == 3
        dn_tile[e] = make_float4((float)q.x * k, (float)q.y * k, (float)q.z * k, 1.0f);
        dn_aux[e] = make_float2((float)(SWAP ? n : d) * (1.0f / 256.0f), __uint_as_float(SWAP ? d : n));   // :37
#else
        dn_tile[e] = make_float4((float)q.x * k, (float)q.y * k, (float)q.z * k, (float)(SWAP ? n : d) * (1.0f / 256.0f));   // :37
        dn_nrm[e] = (uint16_t)(SWAP ? d : n);
#endif
    }
    __syncthreads();
    if (x >= width) return;
    const int c0 = threadIdx.x + 3 * S;                    // this thread's column in the tile
#pragma unroll 1
    for (int o0 = threadIdx.y * DN_G; o0 < L; o0 += DN_TY * DN_G) {       // outputs o0 .. o0+G-1 (decimated rows)
        if (ybase + o0 * S >= row1) break;
        float tw[DN_G], sr[DN_G], sg[DN_G], sb[DN_G], cd[DN_G];
        dn_f2 srg[DN_G], sbw[DN_G];                        // VXRT_DN_PACK: (r, g) and (b, weight) sums as f32x2 pairs
        uint32_t cn[DN_G];
#pragma unroll
        for (int k = 0; k < DN_G; k++) {                   // centre of output k is staged row o0 + k + 3
#if VXRT_DN_PACK == 3
            { const float2 a = dn_aux[(o0 + k + 3) * T::W + c0]; cd[k] = a.x; cn[k] = __float_as_uint(a.y); }
#else
            cd[k] = dn_tile[(o0 + k + 3) * T::W + c0].w;
            cn[k] = lds_u16(dn_nrm + (o0 + k + 3) * T::W + c0);
#endif
            tw[k] = 0.0f; sr[k] = 0.0f; sg[k] = 0.0f; sb[k] = 0.0f;
            srg[k] = dn_pk(0.0f, 0.0f); sbw[k] = dn_pk(0.0f, 0.0f);
        }
        bool sky = true;
#pragma unroll
        for (int k = 0; k < DN_G; k++) sky &= cn[k] >= 16;
        if (__all_sync(__activemask(), sky)) {             // the warp's 32 x 4 outputs are all copies
#pragma unroll
            for (int k = 0; k < DN_G; k++) {
                const int y = ybase + (o0 + k) * S;
                if (y < row1) { const size_t pi = (size_t)y * width + x; put(y, pi, __ldg(src + pi)); }
            }
            continue;
        }
#pragma unroll
        for (int rho = 0; rho < DN_G + 6; rho++) {         // staged row o0 + rho is tap row dy = rho - 3 - k of output k
            // widest tap row any of the G outputs needs from this staged row
            int wmax = 0;
#pragma unroll
            for (int k = 0; k < DN_G; k++) {
                const int dy = rho - 3 - k, ay = dy < 0 ? -dy : dy;
                const int wd = ay <= 1 ? 3 : ay == 2 ? 2 : ay == 3 ? 1 : 0;
                wmax = wd > wmax ? wd : wmax;
            }
#pragma unroll
            for (int dx = -3; dx <= 3; dx++) {
                if (dx < -wmax || dx > wmax) continue;
                const float4 t = dn_tile[(o0 + rho) * T::W + c0 + dx * S];
#if VXRT_DN_PACK == 3
                const float2 ta = dn_aux[(o0 + rho) * T::W + c0 + dx * S];
                const float td = ta.x; const uint32_t tn = __float_as_uint(ta.y);
                const dn_f2 trg = dn_pk(t.x, t.y), tb1 = dn_pk(t.z, t.w);
#else
                const uint32_t tn = lds_u16(dn_nrm + (o0 + rho) * T::W + c0 + dx * S);
                const float td = t.w;
                const dn_f2 trg = dn_pk(t.x, t.y), tb1 = dn_pk(t.z, 1.0f);
#endif
#pragma unroll
                for (int k = 0; k < DN_G; k++) {
                    const int dy = rho - 3 - k;
                    if (dy < -3 || dy > 3) continue;
                    const float wgt = dn_weight(dx < 0 ? -dx : dx, dy < 0 ? -dy : dy);
                    if (wgt == 0.0f) continue;
                    float w;
                    if (dx == 0 && dy == 0) {
                        w = wgt;                                               // :41 centre tap, no bilateral term
                    } else {
                        // w = wgt / (4|dc - dq| + (nq == nc ? 0 : 10) + 1)   (:23-33), one reciprocal of the scaled sum
                        const float pen = (tn == cn[k]) ? (1.0f / wgt) : (11.0f / wgt);
                        w = rcp_fast(fmaf(4.0f / wgt, fabsf(cd[k] - td), pen));
                    }
#if VXRT_DN_PACK >= 2
                    const dn_f2 ww = dn_pk(w, w);
                    srg[k] = dn_fma2(trg, ww, srg[k]); sbw[k] = dn_fma2(tb1, ww, sbw[k]);
#elif VXRT_DN_PACK == 1
                    tw[k] += w; srg[k] = dn_fma2(trg, dn_pk(w, w), srg[k]); sb[k] = fmaf(t.z, w, sb[k]);
#else
                    tw[k] += w; sr[k] = fmaf(t.x, w, sr[k]); sg[k] = fmaf(t.y, w, sg[k]); sb[k] = fmaf(t.z, w, sb[k]);
#endif
                }
            }
        }
#pragma unroll
        for (int k = 0; k < DN_G; k++) {
            const int y = ybase + (o0 + k) * S;
            if (y >= row1) break;
            const size_t pi = (size_t)y * width + x;
#if VXRT_DN_PACK >= 1
            dn_unpk(srg[k], sr[k], sg[k]);
#endif
#if VXRT_DN_PACK >= 2
            dn_unpk(sbw[k], sb[k], tw[k]);
#endif
            if (cn[k] >= 16) {
                put(y, pi, __ldg(src + pi));                                   // :88-89
            } else {
                const float inv = rcp_fast(tw[k]);
                put(y, pi, make_ushort4((unsigned short)unorm_store(sr[k] * inv, 65535.0f), (unsigned short)unorm_store(sg[k] * inv, 65535.0f),
                                        (unsigned short)unorm_store(sb[k] * inv, 65535.0f), 65535));   // :87
            }
        }
    }
}

// Launch priority of the passes after the trace pass (cudaLaunchAttributePriority), for several frames in flight on one
// GPU.  Measured (misc/r1_runs/run_r.sh): at the highest priority C2 goes 0.278 -> 0.287 ms per frame and C3 0.672 -> 0.745, so
// the default is the stream's priority; VXRT_POST_PRIORITY=1 turns it on.
int post_priority() {
    static const int p = [] {
        const char* e = getenv("VXRT_POST_PRIORITY");
        if (!e || atoi(e) == 0) return 0;
        int lo = 0, hi = 0;
        return cudaDeviceGetStreamPriorityRange(&lo, &hi) == cudaSuccess ? hi : 0;   // "greatest" priority is the lowest number
    }();
    return p;
}
template <int S, int L, bool FIN>
cudaError_t launch_denoise_s(const ushort4* src, ushort4* dst, const uint16_t* depth, const uint8_t* normal, int width,
                             int height, int swap, int row0, int row1, cudaStream_t s, bool pdl, const FinalizeImages& F) {
    using T = DnTile<S, L>;
    const int rows = row1 - row0;
    const int per_res = (rows + S - 1) / S;                 // decimated rows of the fullest residue class
    const int groups = (per_res + L - 1) / L;
    dim3 grid((width + DN_TX - 1) / DN_TX, S * groups), block(DN_TX, DN_TY);
    static_assert(L % DN_G == 0, "tile height must be a multiple of the outputs per thread");
    static bool attr_set = false;                           // per template instantiation
    if (!attr_set) {
        cudaError_t e = cudaFuncSetAttribute(k_denoise<S, L, false, FIN>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)T::BYTES);
        if (e == cudaSuccess) e = cudaFuncSetAttribute(k_denoise<S, L, true, FIN>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)T::BYTES);
        if (e != cudaSuccess) return e;
        attr_set = true;
    }
    cudaLaunchConfig_t lc;
    memset(&lc, 0, sizeof lc);
    lc.gridDim = grid; lc.blockDim = block; lc.dynamicSmemBytes = T::BYTES; lc.stream = s;
    cudaLaunchAttribute at[2];
    int na = 0;
    if (post_priority() != 0) { at[na].id = cudaLaunchAttributePriority; at[na].val.priority = post_priority(); na++; }
    if (pdl) { at[na].id = cudaLaunchAttributeProgrammaticStreamSerialization; at[na].val.programmaticStreamSerializationAllowed = 1; na++; }
    lc.attrs = at; lc.numAttrs = na;
    if (swap) return cudaLaunchKernelEx(&lc, k_denoise<S, L, true, FIN>, src, dst, depth, normal, width, height, row0, row1, groups, F);
    return cudaLaunchKernelEx(&lc, k_denoise<S, L, false, FIN>, src, dst, depth, normal, width, height, row0, row1, groups, F);
}

// finalize.comp:33-64.  Colour work under the tolerance contract (<= 1/255 per channel): UNORM decodes are
// reciprocal multiplies folded into the constants, FMA contraction allowed (the IEEE-division form spent ~480
// instructions per pixel and was issue-bound; this one is bound by its 30 B/pixel of traffic).
template <bool PEER>
__global__ void __launch_bounds__(256) k_finalize(const uchar4* __restrict__ albedo, const uchar4* __restrict__ emission,
                                                  const uchar4* __restrict__ fog, const ushort4* __restrict__ lighting,
                                                  const uint16_t* __restrict__ depth, const uint8_t* __restrict__ noise,
                                                  uchar4* __restrict__ out, int width, int height, int row0, int row1,
                                                  const __grid_constant__ VxrtPeerTail tail) {
    // peer mode (peer.cu): `out` is the ROOT's image in peer memory; do not store into it before the root released it
    if (PEER && tail.wait) {
        if (threadIdx.x == 0) vx_spin_until(tail.wait, tail.value, tail.err);
        __syncthreads();
    }
    // grid-stride over 32x8-pixel tiles: a few CTAs per SM, so the peer-mode tail costs one fence + one ticket per
    // CTA, not per tile
    // (PEER only; the single-GPU form is one tile per CTA of a 2-D grid)
    const int tiles_x = (width + 31) / 32, ntiles = tiles_x * ((row1 - row0 + 7) / 8);
    for (int tile = PEER ? blockIdx.x : 0; tile < (PEER ? ntiles : 1); tile += gridDim.x) {
    const int ty = PEER ? tile / tiles_x : blockIdx.y, tx = PEER ? tile - ty * tiles_x : blockIdx.x;
    const int x = tx * 32 + (threadIdx.x & 31), y = row0 + ty * 8 + (threadIdx.x >> 5);
    if (x >= width || y >= row1) continue;
    const FinalizeImages F = {albedo, emission, fog, depth, noise, out, row0, row1};
    finalize_pixel(F, width, height, x, y, __ldg(lighting + (size_t)y * width + x));
    }
    if (PEER && tail.n_post) {                             // the last CTA to finish publishes the frame number
        __shared__ bool last;
        __syncthreads();                                   // the CTA's stores happen-before thread 0's fence
        if (threadIdx.x == 0) { __threadfence_system(); last = atomicAdd(tail.ticket, 1u) == gridDim.x - 1; }
        __syncthreads();
        if (last && threadIdx.x < tail.n_post) {
            if (threadIdx.x == 0) *tail.ticket = 0;
            __threadfence_system();
            vx_st_release_sys(tail.post[threadIdx.x], tail.value);
        }
    }
}

// Per-row work counters of the last frame for the load-balanced band plan (vxrt_plan_bands_weighted): march
// steps, trace_ray calls and pixels the denoise filters (normal code < 16).  One CTA per row.
__global__ void __launch_bounds__(256) k_row_cost(const uint32_t* __restrict__ steps, const uint8_t* __restrict__ rays,
                                                  const uint8_t* __restrict__ normal, int width, float* __restrict__ out) {
    const int y = blockIdx.x;
    unsigned long long s = 0; uint32_t r = 0, n = 0;
    for (int x = threadIdx.x; x < width; x += 256) {
        const size_t pi = (size_t)y * width + x;
        s += steps[pi]; r += rays[pi]; n += normal[pi] < 16 ? 1u : 0u;
    }
    __shared__ unsigned long long ss[256]; __shared__ uint32_t sr[256], sn[256];
    ss[threadIdx.x] = s; sr[threadIdx.x] = r; sn[threadIdx.x] = n;
    __syncthreads();
    for (int k = 128; k > 0; k >>= 1) {
        if (threadIdx.x < k) { ss[threadIdx.x] += ss[threadIdx.x + k]; sr[threadIdx.x] += sr[threadIdx.x + k]; sn[threadIdx.x] += sn[threadIdx.x + k]; }
        __syncthreads();
    }
    if (threadIdx.x == 0) { out[3 * y] = (float)ss[0]; out[3 * y + 1] = (float)sr[0]; out[3 * y + 2] = (float)sn[0]; }
}

}  // namespace


// CUDA loads a kernel's code at its first launch, and that load cannot complete while another kernel is resident
// and spinning on a flag (peer.cu).  Every kernel of this file is therefore loaded when a handle is created.
#define VXRT_PRELOAD(k) do { cudaFuncAttributes a_; cudaError_t e_ = cudaFuncGetAttributes(&a_, k); if (e_ != cudaSuccess) return e_; } while (0)
template <int S, int L, bool FIN> cudaError_t preload_denoise() { VXRT_PRELOAD((k_denoise<S, L, false, FIN>)); VXRT_PRELOAD((k_denoise<S, L, true, FIN>)); return cudaSuccess; }
cudaError_t vxrt_preload_post() {
    cudaError_t e;
    if ((e = preload_denoise<1, VXRT_DN_L0, false>()) || (e = preload_denoise<2, VXRT_DN_L0, false>()) || (e = preload_denoise<4, VXRT_DN_L0, false>()) ||
        (e = preload_denoise<8, VXRT_DN_L1, false>()) || (e = preload_denoise<16, VXRT_DN_L1, false>()) || (e = preload_denoise<16, VXRT_DN_L1, true>())) return e;
    VXRT_PRELOAD(k_finalize<false>); VXRT_PRELOAD(k_finalize<true>); VXRT_PRELOAD(k_row_cost);
    return cudaSuccess;
}

cudaError_t vxrt_launch_row_cost(const uint32_t* steps, const uint8_t* rays, const uint8_t* normal, int width, int height,
                                 float* out, cudaStream_t s) {
    k_row_cost<<<height, 256, 0, s>>>(steps, rays, normal, width, out);
    return cudaGetLastError();
}

cudaError_t vxrt_launch_denoise(const ushort4* src, ushort4* dst, const uint16_t* depth, const uint8_t* normal,
                                int width, int height, int size, int swap, int row0, int row1, cudaStream_t s, bool pdl, const VxrtFinalizeFold* fold) {
    if (row1 <= row0) return cudaSuccess;
    FinalizeImages F; memset(&F, 0, sizeof F);
    if (fold) {                                            // the last pass (size 16) with finalize.comp folded in
        if (size != 16) return cudaErrorInvalidValue;
        F.albedo = fold->albedo; F.emission = fold->emission; F.fog = fold->fog; F.depth = depth; F.noise = fold->noise; F.out = fold->out;
        F.row0 = fold->row0; F.row1 = fold->row1;
        return launch_denoise_s<16, VXRT_DN_L1, true>(src, dst, depth, normal, width, height, swap, row0, row1, s, pdl, F);
    }
    switch (size) {                                        // pipeline.rs:103: sizes 1, 2, 4, 8, 8, 16
        case 1: return launch_denoise_s<1, VXRT_DN_L0, false>(src, dst, depth, normal, width, height, swap, row0, row1, s, pdl, F);
        case 2: return launch_denoise_s<2, VXRT_DN_L0, false>(src, dst, depth, normal, width, height, swap, row0, row1, s, pdl, F);
        case 4: return launch_denoise_s<4, VXRT_DN_L0, false>(src, dst, depth, normal, width, height, swap, row0, row1, s, pdl, F);
        case 8: return launch_denoise_s<8, VXRT_DN_L1, false>(src, dst, depth, normal, width, height, swap, row0, row1, s, pdl, F);
        case 16: return launch_denoise_s<16, VXRT_DN_L1, false>(src, dst, depth, normal, width, height, swap, row0, row1, s, pdl, F);
        default: return cudaErrorInvalidValue;
    }
}

cudaError_t vxrt_launch_finalize(const uchar4* albedo, const uchar4* emission, const uchar4* fog, const ushort4* lighting,
                                 const uint16_t* depth, const uint8_t* noise, uchar4* out, int width, int height,
                                 int row0, int row1, cudaStream_t s, const VxrtPeerTail* tail) {
    if (row1 <= row0) return cudaSuccess;
    const int ntiles = ((width + 31) / 32) * ((row1 - row0 + 7) / 8);
    VxrtPeerTail none; memset(&none, 0, sizeof none);
    if (tail)       // peer tail: one fence + ticket per CTA, so keep the CTA count low
        k_finalize<true><<<ntiles > 148 * 8 ? 148 * 8 : ntiles, 256, 0, s>>>(albedo, emission, fog, lighting, depth, noise, out, width, height, row0, row1, *tail);
    else {
        cudaLaunchConfig_t lc;
        memset(&lc, 0, sizeof lc);
        lc.gridDim = dim3((width + 31) / 32, (row1 - row0 + 7) / 8); lc.blockDim = dim3(256); lc.stream = s;
        cudaLaunchAttribute at[1];
        at[0].id = cudaLaunchAttributePriority; at[0].val.priority = post_priority();
        lc.attrs = at; lc.numAttrs = post_priority() != 0 ? 1 : 0;
        return cudaLaunchKernelEx(&lc, k_finalize<false>, albedo, emission, fog, lighting, depth, noise, out, width, height, row0, row1, none);
    }
    return cudaGetLastError();
}
