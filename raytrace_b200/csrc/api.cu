// api.cu — the extern "C" boundary of libvxrt (include/vxrt.h): handle life cycle, world upload,
// frame parameters, the per-frame launch sequence and read-back.  Replaces the dispatch code of
// src/render/pipeline/{pipeline.rs,render_data.rs,terrain_upload.rs} (reference).
// No CPU fallback lives here: every compute entry point launches CUDA kernels or fails.
#include "vxrt_internal.h"
#include <math.h>
#include <stdio.h>
#include <string.h>
#include <stdlib.h>
#include <new>

namespace {

std::string g_create_error;

#define CU(call)                                                                              \
    do {                                                                                      \
        cudaError_t e_ = (call);                                                              \
        if (e_ != cudaSuccess) {                                                              \
            h->err = std::string(#call) + ": " + cudaGetErrorString(e_);                      \
            return VXRT_ERR_CUDA;                                                             \
        }                                                                                     \
    } while (0)

int fail(VxrtHandle* h, int code, const char* msg) { h->err = msg; return code; }

bool pow2(int v) { return v > 0 && (v & (v - 1)) == 0; }
int ilog2(int v) { int l = 0; while ((1 << l) < v) l++; return l; }

// pipeline.rs:103 — dilation of each of the six bilateral passes
const int kDenoiseSizes[6] = {1, 2, 4, 8, 8, 16};

// host-side fp32 helpers; this file is compiled with -ffp-contract=off so each operator is one rounding
struct H3 { float x, y, z; };
float hmix(float a, float b, float t) { return a * (1.0f - t) + b * t; }

// raytrace.comp:317-318: per-frame constants evaluated once (cos/sin in double, rounded once — the
// same rule the oracle uses, oracle/vxrt_oracle_trace.cpp orc_trace)
void sun_constants(float sun_angle, H3* dir, H3* light) {
    const float ca = (float)cos((double)sun_angle), sa = (float)sin((double)sun_angle);
    H3 v = {ca * 0.5f + (sun_angle - 0.5f) * 0.5f, sa, ca};
    const float l = sqrtf((v.x * v.x + v.y * v.y) + v.z * v.z);
    dir->x = v.x / l; dir->y = v.y / l; dir->z = v.z / l;
    // sun_color, raytrace.comp:259-269
    const float horizon = sqrtf(dir->x * dir->x + dir->y * dir->y);
    const float sun_amount = fminf(1.0f - horizon, 0.02f) * 50.0f;
    const H3 main_color = {0.9647f * 2.0f, 0.7843f * 2.0f, 0.8824f * 2.0f};
    const H3 sunset = {0.7412f * 2.0f, 0.2157f * 2.0f, 0.1686f * 2.0f};
    if (dir->z >= 0.0f) {
        light->x = hmix(sunset.x, main_color.x, sun_amount); light->y = hmix(sunset.y, main_color.y, sun_amount);
        light->z = hmix(sunset.z, main_color.z, sun_amount);
    } else {
        const float t = sun_amount * 2.0f;
        light->x = hmix(sunset.x, 0.0f, t); light->y = hmix(sunset.y, 0.0f, t); light->z = hmix(sunset.z, 0.0f, t);
    }
}

// raytrace.comp:190-197 as 256-entry tables (noise channels are 8 bit); same rule as orc_diffuse_luts
void diffuse_luts(float* lut) {
    const float PI = 3.1415926535897932384626433832795f;
    for (int b = 0; b < 256; b++) {
        const float n = (float)b / 255.0f;
        const float theta1 = PI * 2.0f * n;
        const float theta2 = (float)acos((double)(1.0f - 2.0f * n));
        lut[b] = (float)sin((double)theta1);
        lut[256 + b] = (float)cos((double)theta1);
        lut[512 + b] = (float)sin((double)theta2);
        lut[768 + b] = (float)cos((double)theta2);
    }
}

// (E12) trace_fast.cu: values the shader computes from small integers, tabulated with the shader's own sequence of
// IEEE single-precision operations (volatile keeps the host compiler from contracting or widening them)
void value_tables(float* lut, int width, int height) {
    for (int n = 0; n < 256; n++) { volatile float q = (float)n / 255.0f; volatile float j = q * 0.05f; lut[VXRT_LUT_JITTER + n] = j; }
    for (int n = 0; n < 128; n++) { volatile float q = (float)n / 127.0f; lut[VXRT_LUT_ALBEDO + n] = q; }
    for (int a = 0; a < 2; a++) {
        const int n = a ? height : width;
        float* t = lut + VXRT_LUT_SCREEN + (a ? width : 0);
        for (int i = 0; i < n; i++) { volatile float q = (float)i / (float)n; volatile float m = q * 2.0f; volatile float r = m - 1.0f; t[i] = r; }
    }
}

// trace_fast.cu writes the albedo image as 2n + (n >> 6) per 7-bit field: the same byte as UNORM8(n / 127) with this host's
// fp32 operations (the sequence of raytrace.comp:156-158 + the RGBA8 store), for every n
bool albedo_bytes_ok() {
    for (int n = 0; n < 128; n++) {
        volatile float q = (float)n / 127.0f; volatile float c = q < 0.0f ? 0.0f : (q > 1.0f ? 1.0f : q); volatile float v = c * 255.0f;
        if ((int)nearbyintf(v) != 2 * n + (n >> 6)) return false;
    }
    return true;
}

// (E8) trace_fast.cu: g(p) = fl(p - lr) is monotone in p, so {p : |g(p)| >= H} = {p <= lo} U {p >= hi}.
// lo/hi are found by bisection over the floats in their numeric order.
uint32_t float_key(float f) { uint32_t u; memcpy(&u, &f, 4); return (u & 0x80000000u) ? ~u : (u | 0x80000000u); }
float key_float(uint32_t k) { uint32_t u = (k & 0x80000000u) ? (k & 0x7FFFFFFFu) : ~k; float f; memcpy(&f, &u, 4); return f; }
void sky_thresholds(float lr, float half, float* lo, float* hi) {
    auto g = [lr](float p) { volatile float r = p - lr; return (float)r; };
    uint32_t a = float_key(-3.0e38f), b = float_key(3.0e38f);      // g(a) < H <= g(b)
    while (b - a > 1) { const uint32_t m = a + (b - a) / 2; if (g(key_float(m)) >= half) b = m; else a = m; }
    *hi = key_float(b);
    a = float_key(-3.0e38f); b = float_key(3.0e38f);               // g(a) <= -H < g(b)
    while (b - a > 1) { const uint32_t m = a + (b - a) / 2; if (g(key_float(m)) <= -half) a = m; else b = m; }
    *lo = key_float(a);
}

size_t pixels(const VxrtHandle* h) { return (size_t)h->cfg.width * h->cfg.height; }
size_t voxels(const VxrtHandle* h) { return (size_t)h->cfg.nx * h->cfg.ny * h->cfg.nz; }

struct BufDesc { void* ptr; size_t bytes; };
BufDesc buffer_of(const VxrtHandle* h, int id) {
    const size_t P = pixels(h);
    switch (id) {
        case VXRT_BUF_OUTPUT: return {h->d_output, P * 4};
        case VXRT_BUF_LIGHTING: return {h->d_lighting, P * 8};
        case VXRT_BUF_DEPTH: return {h->d_depth, P * 2};
        case VXRT_BUF_NORMAL: return {h->d_normal, P};
        case VXRT_BUF_ALBEDO: return {h->d_albedo, P * 4};
        case VXRT_BUF_EMISSION: return {h->d_emission, P * 4};
        case VXRT_BUF_FOG: return {h->d_fog, P * 4};
        case VXRT_BUF_HIT_VOXEL: return {h->d_hit_voxel, h->d_hit_voxel ? P * 12 : 0};
        case VXRT_BUF_HIT_FACE: return {h->d_hit_face, h->d_hit_face ? P : 0};
        case VXRT_BUF_HIT_WORD: return {h->d_hit_word, h->d_hit_word ? P * 4 : 0};
        case VXRT_BUF_HIT_T: return {h->d_hit_t, h->d_hit_t ? P * 4 : 0};
        case VXRT_BUF_STEPS: return {h->d_steps, h->d_steps ? P * 4 : 0};
        case VXRT_BUF_RAYS: return {h->d_rays, h->d_rays ? P : 0};
        case VXRT_BUF_ACCUM: return {h->d_accum, h->d_accum ? P * 16 : 0};
        case VXRT_BUF_LIGHTING_RAW: return {h->d_lighting_raw, h->d_lighting_raw ? P * 8 : 0};
        case VXRT_BUF_HIT_POS: return {h->d_hit_pos[h->hist_idx], h->d_hit_pos[0] ? P * 16 : 0};
        case VXRT_BUF_HISTORY: return {h->d_hist[h->hist_idx], h->d_hist[0] ? P * 16 : 0};
        case VXRT_BUF_COMPLETED: return {h->d_completed, h->d_completed ? P * 8 : 0};
        default: return {nullptr, 0};
    }
}

int alloc_zero(VxrtHandle* h, void** p, size_t bytes) {
    CU(cudaMalloc(p, bytes));
    CU(cudaMemsetAsync(*p, 0, bytes, h->stream));
    return VXRT_OK;
}

// 3-D box copy host -> window (both x fastest).  Box extents (ex,ey,ez) at texel (x0,y0,z0).
template <typename T>
int upload_box(VxrtHandle* h, T* d_base, const T* src, int x0, int y0, int z0, int ex, int ey, int ez) {
    cudaMemcpy3DParms p;
    memset(&p, 0, sizeof p);
    p.srcPtr = make_cudaPitchedPtr(const_cast<T*>(src), (size_t)ex * sizeof(T), (size_t)ex, (size_t)ey);
    p.dstPtr = make_cudaPitchedPtr(d_base, (size_t)h->cfg.nx * sizeof(T), (size_t)h->cfg.nx, (size_t)h->cfg.ny);
    p.dstPos = make_cudaPos((size_t)x0 * sizeof(T), (size_t)y0, (size_t)z0);
    p.extent = make_cudaExtent((size_t)ex * sizeof(T), (size_t)ey, (size_t)ez);
    p.kind = cudaMemcpyHostToDevice;
    CU(cudaMemcpy3DAsync(&p, h->stream));
    return VXRT_OK;
}

int mark_dirty_box(VxrtHandle* h, int x0, int y0, int z0, int ex, int ey, int ez) {
    // host-side list of chunks to rebuild, pushed as flags
    const int cxn = h->cfg.nx >> 6, cyn = h->cfg.ny >> 6, czn = h->cfg.nz >> 6;
    std::string flags((size_t)cxn * cyn * czn * 4, '\0');
    uint32_t* f = reinterpret_cast<uint32_t*>(&flags[0]);
    for (int cz = z0 >> 6; cz <= (z0 + ez - 1) >> 6; cz++)
        for (int cy = y0 >> 6; cy <= (y0 + ey - 1) >> 6; cy++)
            for (int cx = x0 >> 6; cx <= (x0 + ex - 1) >> 6; cx++) f[(cz * cyn + cy) * cxn + cx] = 1u;
    CU(cudaMemcpyAsync(h->d_dirty, f, flags.size(), cudaMemcpyHostToDevice, h->stream));
    CU(cudaStreamSynchronize(h->stream));
    return VXRT_OK;
}

// World writes on a handle whose world other handles borrow (vxrt_create_shared): the borrowers' frames in flight
// finish first; their next frame waits for the write (ev_world) on its own stream.
int world_write_begin(VxrtHandle* h, const char* who) {
    if (h->world_owner) { h->err = std::string(who) + ": this handle borrows its world (vxrt_create_shared); write through the owner"; return VXRT_ERR_STATE; }
    for (VxrtHandle* s : h->sharers) CU(cudaStreamSynchronize(s->stream));
    return VXRT_OK;
}
// layout A/B (VXRT_L2_PERSIST=<hit ratio in percent>): the occupancy field as a persisting access-policy window on the handle's
// stream (up to the device's window / set-aside limits), everything else streaming through the rest of L2.  Measured (r2_m):
// no gain on C2 (0.1354 -> 0.1378 ms per trace pass, L2 hits 68.4 -> 68.3 %) or C4 (2.789 -> 2.800 ms), so it stays a switch.
int set_l2_window(VxrtHandle* h, cudaStream_t s) {
    const char* e = getenv("VXRT_L2_PERSIST");
    if (!e || !h->d_minefield) return VXRT_OK;
    int dev = h->cfg.device, max_win = 0, max_persist = 0;
    CU(cudaDeviceGetAttribute(&max_win, cudaDevAttrMaxAccessPolicyWindowSize, dev));
    CU(cudaDeviceGetAttribute(&max_persist, cudaDevAttrMaxPersistingL2CacheSize, dev));
    if (max_win <= 0 || max_persist <= 0) return VXRT_OK;
    CU(cudaDeviceSetLimit(cudaLimitPersistingL2CacheSize, (size_t)max_persist));
    cudaStreamAttrValue v;
    memset(&v, 0, sizeof v);
    size_t bytes = voxels(h);
    if (bytes > (size_t)max_win) bytes = (size_t)max_win;
    v.accessPolicyWindow.base_ptr = h->d_minefield;
    v.accessPolicyWindow.num_bytes = bytes;
    v.accessPolicyWindow.hitRatio = (float)atoi(e) / 100.0f;
    v.accessPolicyWindow.hitProp = cudaAccessPropertyPersisting;
    v.accessPolicyWindow.missProp = cudaAccessPropertyStreaming;
    CU(cudaStreamSetAttribute(s, cudaStreamAttributeAccessPolicyWindow, &v));
    static bool said = false;
    if (!said) { fprintf(stderr, "vxrt: L2 persisting window %zu MiB (max window %d MiB, set-aside %d MiB), hit ratio %.2f\n", bytes >> 20, max_win >> 20, max_persist >> 20, v.accessPolicyWindow.hitRatio); said = true; }
    return VXRT_OK;
}

int world_write_end(VxrtHandle* h, const int* box = nullptr /* x0 y0 z0 ex ey ez of what was written, NULL = anything */) {
    if (h->field_arr) {                          // the texture's copy of the occupancy field follows every world write
        const int whole[6] = {0, 0, 0, h->cfg.nx, h->cfg.ny, h->cfg.nz};
        const int* b = box ? box : whole;
        cudaMemcpy3DParms p;
        memset(&p, 0, sizeof p);
        p.srcPtr = make_cudaPitchedPtr(h->d_minefield, (size_t)h->cfg.nx, (size_t)h->cfg.nx, (size_t)h->cfg.ny);
        p.srcPos = make_cudaPos((size_t)b[0], (size_t)b[1], (size_t)b[2]);
        p.dstArray = h->field_arr;
        p.dstPos = make_cudaPos((size_t)b[0], (size_t)b[1], (size_t)b[2]);
        p.extent = make_cudaExtent((size_t)b[3], (size_t)b[4], (size_t)b[5]);
        p.kind = cudaMemcpyDeviceToDevice;
        CU(cudaMemcpy3DAsync(&p, h->stream));
    }
    if (h->d_field_nib) CU(vxrt_launch_field_to_nibbles(h->d_minefield, h->d_field_nib, h->cfg.nx, h->cfg.ny, h->cfg.nz, box, h->stream));
    if (h->d_field_brick) CU(vxrt_launch_field_to_bricks(h->d_minefield, h->d_field_brick, h->cfg.nx, h->cfg.ny, h->cfg.nz, box, h->stream));
    if (h->sharers.empty()) return VXRT_OK;
    CU(cudaEventRecord(h->ev_world, h->stream));
    h->world_version++;
    return VXRT_OK;
}
const VxrtHandle* world_of(const VxrtHandle* h) { return h->world_owner ? h->world_owner : h; }
// the handle that keeps the accumulation chain's event: the owner of a shared running sum, else the handle itself
VxrtHandle* accum_keeper(VxrtHandle* h) { return h->accum_borrowed ? h->world_owner : h; }
// order this handle's stream after the last write of the running sum when another handle made it
int accum_wait(VxrtHandle* h) {
    VxrtHandle* k = accum_keeper(h);
    if (k && k->ev_accum && k->accum_last && k->accum_last != h) CU(cudaStreamWaitEvent(h->stream, k->ev_accum, 0));
    return VXRT_OK;
}
int accum_written(VxrtHandle* h) {
    VxrtHandle* k = accum_keeper(h);
    if (k && k->ev_accum && (!k->sharers.empty() || k != h)) { CU(cudaEventRecord(k->ev_accum, h->stream)); k->accum_last = h; }
    return VXRT_OK;
}

int rebuild_dirty(VxrtHandle* h) {
    CU(vxrt_launch_build_minefield(h->d_materials, h->d_minefield, h->cfg.nx, h->cfg.ny, h->cfg.nz, h->d_dirty, h->stream));
    CU(cudaMemsetAsync(h->d_dirty, 0, (size_t)(h->cfg.nx >> 6) * (h->cfg.ny >> 6) * (h->cfg.nz >> 6) * 4, h->stream));
    return VXRT_OK;
}

}  // namespace

void vxrt_set_global_error(const char* msg) { g_create_error = msg ? msg : ""; }

extern "C" {

int vxrt_abi_version(void) { return VXRT_ABI_VERSION; }

// ---- image-space sharding plan (pure host arithmetic; no device needed) -------------------------------
int vxrt_plan_bands(int height, int nranks, int32_t* rows) {
    if (height <= 0 || nranks <= 0 || !rows) return VXRT_ERR_INVALID;
    // contiguous bands whose edges are multiples of 8 rows (the 8x8 shader group, constants.rs:33)
    for (int r = 0; r <= nranks; r++) {
        long long e = (long long)height * r / nranks;
        e = (e + 7) / 8 * 8;
        if (e > height || r == nranks) e = height;
        if (r < nranks) rows[2 * r] = (int32_t)e;
        if (r > 0) rows[2 * r - 1] = (int32_t)e;
    }
    return VXRT_OK;
}

int vxrt_plan_passes(int height, int row_begin, int row_end, uint32_t flags, int32_t* rows) {
    if (!rows || height <= 0 || row_begin < 0 || row_end > height || row_begin >= row_end) return VXRT_ERR_INVALID;
    int halo = 0;
    if (flags & VXRT_FLAG_DENOISE) for (int k = 0; k < 6; k++) halo += 3 * kDenoiseSizes[k];
    rows[0] = row_begin - halo < 0 ? 0 : row_begin - halo;
    rows[1] = row_end + halo > height ? height : row_end + halo;
    int remaining = halo;
    for (int k = 0; k < 6; k++) {
        remaining -= 3 * kDenoiseSizes[k];
        if (remaining < 0) remaining = 0;
        int r0 = row_begin - remaining, r1 = row_end + remaining;
        rows[2 + 2 * k] = r0 < 0 ? 0 : r0;
        rows[3 + 2 * k] = r1 > height ? height : r1;
    }
    return VXRT_OK;
}

// Load-balanced bands.  Cost model in units of one march step (from the ncu instruction counts of r1: 47
// instructions per step, ~330 per ray, ~560 per pixel of the trace pass; ~375 per filtered pixel and pass of the
// denoise at 2/3 of the trace pass's issue rate; ~220 per pixel of finalize + copies):
//   trace(row)          = steps + 7 rays + 12 W
//   denoise(row, pass)  = 15 filtered + 0.45 W
//   finalize(row)       = 4.7 W
// A band also filters its halo rows: a row at distance d outside the band is filtered by every pass whose
// remaining halo (114, 108, 96, 72, 48, 0 rows after passes 1..6) still reaches it.
int vxrt_plan_bands_weighted(int height, int nranks, uint32_t flags, const float* trace_cost, const float* denoise_cost, int32_t* rows) {
    if (height <= 0 || nranks <= 0 || !rows || !trace_cost || !denoise_cost) return VXRT_ERR_INVALID;
    if (nranks == 1) { rows[0] = 0; rows[1] = height; return VXRT_OK; }
    const bool dn = (flags & VXRT_FLAG_DENOISE) != 0;
    std::vector<double> pt(height + 1, 0.0), pd(height + 1, 0.0);
    for (int y = 0; y < height; y++) { pt[y + 1] = pt[y] + trace_cost[y]; pd[y + 1] = pd[y] + (dn ? denoise_cost[y] : 0.0); }
    auto sum = [&](const std::vector<double>& p, int a, int b) { a = a < 0 ? 0 : a; b = b > height ? height : b; return a < b ? p[b] - p[a] : 0.0; };
    int reach[6]; { int rem = 0; for (int k = 0; k < 6; k++) rem += 3 * kDenoiseSizes[k]; for (int k = 0; k < 6; k++) { rem -= 3 * kDenoiseSizes[k]; reach[k] = rem; } }
    auto cost = [&](int a, int b) {
        double c = sum(pt, a, b) + 6.0 * sum(pd, a, b);
        if (dn) for (int k = 0; k < 6; k++) c += sum(pd, a - reach[k], a) + sum(pd, b, b + reach[k]);
        return c;
    };
    // smallest T such that greedy bands of cost <= T (edges on multiples of 8 rows) cover the frame
    auto cover = [&](double T, int32_t* out) {
        int a = 0;
        for (int r = 0; r < nranks; r++) {
            int b = a + 8 > height ? height : a + 8;                  // a band has at least one 8-row group
            if (r == nranks - 1) b = height;
            else while (b < height && cost(a, b + 8 > height ? height : b + 8) <= T) b = b + 8 > height ? height : b + 8;
            // leave at least 8 rows for every later rank
            const int max_b = height - 8 * (nranks - 1 - r);
            if (b > max_b) b = max_b;
            if (b <= a) return false;
            if (out) { out[2 * r] = a; out[2 * r + 1] = b; }
            a = b;
        }
        return true;
    };
    if (height < 8 * nranks) return VXRT_ERR_INVALID;
    double lo = 0.0, hi = cost(0, height);
    std::vector<int32_t> tmp(2 * nranks);
    for (int it = 0; it < 60; it++) {
        const double T = 0.5 * (lo + hi);
        cover(T, tmp.data());
        const double last = cost(tmp[2 * (nranks - 1)], height);     // the last band takes whatever is left
        if (last <= T) hi = T; else lo = T;
    }
    cover(hi, rows);
    return VXRT_OK;
}

const char* vxrt_last_error(const VxrtHandle* h) { return h ? h->err.c_str() : g_create_error.c_str(); }

static int create_impl(const VxrtConfig* cfg, VxrtHandle* owner, VxrtHandle** out) {
    if (!out) { g_create_error = "vxrt_create: out is NULL"; return VXRT_ERR_INVALID; }
    *out = nullptr;
    if (!cfg) { g_create_error = "vxrt_create: cfg is NULL"; return VXRT_ERR_INVALID; }
    if (cfg->abi_version != VXRT_ABI_VERSION) { g_create_error = "vxrt_create: abi_version mismatch"; return VXRT_ERR_INVALID; }
    if (cfg->width <= 0 || cfg->height <= 0 || cfg->width > 32768 || cfg->height > 32768) {
        g_create_error = "vxrt_create: bad frame size"; return VXRT_ERR_INVALID;
    }
    if (!pow2(cfg->nx) || !pow2(cfg->ny) || !pow2(cfg->nz) || cfg->nx < 64 || cfg->ny < 64 || cfg->nz < 64 ||
        cfg->nx > 4096 || cfg->ny > 4096 || cfg->nz > 4096) {
        g_create_error = "vxrt_create: window dims must be powers of two in [64, 4096]"; return VXRT_ERR_INVALID;
    }
    if (cfg->mode < VXRT_MODE_PRIMARY || cfg->mode > VXRT_MODE_FULL) { g_create_error = "vxrt_create: bad mode"; return VXRT_ERR_INVALID; }
    if (cfg->max_steps < 0) { g_create_error = "vxrt_create: max_steps < 0"; return VXRT_ERR_INVALID; }
    int row0 = cfg->row_begin, row1 = cfg->row_end;
    if (row0 == 0 && row1 == 0) row1 = cfg->height;
    if (row0 < 0 || row1 > cfg->height || row0 >= row1) { g_create_error = "vxrt_create: bad row band"; return VXRT_ERR_INVALID; }
    if (cfg->flags & VXRT_FLAG_REPROJECT) {
        if (cfg->flags & VXRT_FLAG_ACCUMULATE) { g_create_error = "vxrt_create: VXRT_FLAG_REPROJECT excludes VXRT_FLAG_ACCUMULATE"; return VXRT_ERR_INVALID; }
        if (row0 != 0 || row1 != cfg->height) { g_create_error = "vxrt_create: VXRT_FLAG_REPROJECT needs the whole frame on one handle"; return VXRT_ERR_INVALID; }
    }
    if (!albedo_bytes_ok()) { g_create_error = "vxrt_create: UNORM8(n / 127) != 2n + (n >> 6) on this host (trace_fast.cu albedo image)"; return VXRT_ERR_INVALID; }
    int ndev = 0;
    cudaError_t e = cudaGetDeviceCount(&ndev);
    if (e != cudaSuccess || ndev == 0) {
        g_create_error = std::string("vxrt_create: no CUDA device (there is no CPU fallback): ") + cudaGetErrorString(e);
        return VXRT_ERR_CUDA;
    }
    if (cfg->device < 0 || cfg->device >= ndev) { g_create_error = "vxrt_create: device ordinal out of range"; return VXRT_ERR_INVALID; }
    if (owner) {
        if (owner->world_owner) { g_create_error = "vxrt_create_shared: the owner itself borrows its world"; return VXRT_ERR_INVALID; }
        if (owner->cfg.device != cfg->device || owner->cfg.nx != cfg->nx || owner->cfg.ny != cfg->ny || owner->cfg.nz != cfg->nz) {
            g_create_error = "vxrt_create_shared: device and window dims must equal the owner's"; return VXRT_ERR_INVALID;
        }
    }
    e = cudaSetDevice(cfg->device);
    if (e != cudaSuccess) { g_create_error = std::string("cudaSetDevice: ") + cudaGetErrorString(e); return VXRT_ERR_CUDA; }

    VxrtHandle* h = new (std::nothrow) VxrtHandle();
    if (!h) { g_create_error = "vxrt_create: out of host memory"; return VXRT_ERR_INVALID; }
    h->cfg = *cfg;
    if (h->cfg.max_steps == 0) h->cfg.max_steps = 2048;      // raytrace.comp:109
    h->row0 = row0; h->row1 = row1;
    // denoise needs 3*size rows of context per pass (bilateral_denoise.comp taps reach +-3*size)
    vxrt_plan_passes(cfg->height, row0, row1, cfg->flags, h->plan);
    h->trace_row0 = h->plan[0];
    h->trace_row1 = h->plan[1];
    h->stream = nullptr; h->copy_stream = nullptr; h->out_idx = 0; h->tickets = 0;
    for (int i = 0; i < 2; i++) { h->ev_frame_done[i] = h->ev_copy_done[i] = nullptr; h->copy_pending[i] = false; h->d_out_buf[i] = nullptr; h->ticket_of_buf[i] = 0; }
    for (int i = 0; i < 5; i++) h->ev[i] = nullptr;
    h->d_materials = nullptr; h->d_minefield = nullptr; h->d_noise = nullptr; h->d_lut = nullptr; h->d_dirty = nullptr;
    h->d_edits = nullptr; h->edits_cap = 0;
    h->d_lighting = h->d_pong = h->d_lighting_raw = nullptr;
    h->d_depth = nullptr; h->d_normal = nullptr; h->d_albedo = h->d_emission = h->d_fog = h->d_output = nullptr;
    h->d_hit_voxel = nullptr; h->d_hit_face = nullptr; h->d_hit_word = nullptr; h->d_hit_t = nullptr; h->d_steps = nullptr;
    h->d_rays = nullptr; h->d_accum = nullptr; h->d_counters = nullptr;
    h->have_frame = h->have_world = h->have_noise = h->rendered = false;
    h->frames_accumulated = 0;
    h->comm = nullptr; h->rank = 0; h->nranks = 1;
    memset(&h->frame, 0, sizeof h->frame);
    h->h_noise.assign((size_t)VXRT_NOISE_SIZE * VXRT_NOISE_SIZE * 4, 0);

    auto init = [&]() -> int {
        // load every kernel now: a lazy load at first launch cannot complete while a peer-mode kernel spins on a flag
        CU(vxrt_preload_trace()); CU(vxrt_preload_trace_fast()); CU(vxrt_preload_trace_pool()); CU(vxrt_preload_trace_wave()); CU(vxrt_preload_post()); CU(vxrt_preload_world()); CU(vxrt_preload_peer()); CU(vxrt_preload_temporal());
        CU(cudaStreamCreateWithFlags(&h->stream, cudaStreamNonBlocking));
        for (int i = 0; i < 5; i++) CU(cudaEventCreate(&h->ev[i]));
        CU(cudaStreamCreateWithFlags(&h->copy_stream, cudaStreamNonBlocking));
        for (int i = 0; i < 2; i++) {
            CU(cudaEventCreateWithFlags(&h->ev_frame_done[i], cudaEventDisableTiming));
            CU(cudaEventCreateWithFlags(&h->ev_copy_done[i], cudaEventDisableTiming));
        }
        const size_t V = voxels(h), P = pixels(h);
        int rc;
        CU(cudaEventCreateWithFlags(&h->ev_world, cudaEventDisableTiming));
        if (owner) {                             // the world is the owner's: no second copy in HBM
            h->d_materials = owner->d_materials; h->d_minefield = owner->d_minefield; h->d_noise = owner->d_noise;
        } else {
            if ((rc = alloc_zero(h, (void**)&h->d_materials, V * 4))) return rc;
            // the occupancy field inside ONE 4 GiB line of the address space, so that the march's 32-bit address form (E15) - and with
            // it the kernels with baked window dims, the texture forms and the wavefront - applies whatever address cudaMalloc hands
            // out (a 1 GiB field crossed a line in two of five processes on the B200: those ran the 64-bit form, 2.8 instead of 2.26 ms
            // per C4 trace pass).  A block that crosses is replaced by one of twice the size, which holds a window that does not.
            if ((rc = alloc_zero(h, (void**)&h->d_minefield_base, V))) return rc;
            h->d_minefield = h->d_minefield_base;
            {
                const unsigned long long b = (unsigned long long)h->d_minefield_base, line = (b + V - 1) & ~0xFFFFFFFFull;   // the 4 GiB line inside [b, b + V), if any
                if ((line > b || getenv("VXRT_FIELD_REALLOC") /* test hook: take this path whatever the address */) && V <= 0x80000000ull) {
                    cudaFree(h->d_minefield_base); h->d_minefield_base = nullptr;
                    if ((rc = alloc_zero(h, (void**)&h->d_minefield_base, 2 * V))) return rc;
                    const unsigned long long b2 = (unsigned long long)h->d_minefield_base, l2 = (b2 + V - 1) & ~0xFFFFFFFFull;
                    h->d_minefield = reinterpret_cast<uint8_t*>(l2 > b2 ? l2 : b2);      // the line itself is 4 GiB-aligned, and l2 + V <= b2 + 2V
                }
            }
            if ((rc = alloc_zero(h, (void**)&h->d_noise, h->h_noise.size()))) return rc;
            if ((rc = alloc_zero(h, (void**)&h->d_dirty, (size_t)(cfg->nx >> 6) * (cfg->ny >> 6) * (cfg->nz >> 6) * 4))) return rc;
            // Occupancy field through the texture unit where it does not fit L2: measured on the 1 GiB field of C4
            // 2.79 -> 2.57 ms per trace pass (L1 hits 46 -> 61 %, DRAM reads 1.44 -> 1.31 GB); on the 128 MiB field of C2/C3/C5 it is
            // a wash (0.1354 vs 0.1378 ms) and costs a second copy, so those read the linear array.  VXRT_FIELD=lin|tex|texn overrides.
            const char* fm = getenv("VXRT_FIELD");
            h->field_mode = !fm ? ((cfg->nx == 2048 && cfg->ny == 2048 && cfg->nz == 256) ? 1 : 0)      // (the window the texture forms are built for)
                                : !strcmp(fm, "tex") ? 1 : !strcmp(fm, "texn") ? 2 : !strcmp(fm, "brick") ? 3 : !strcmp(fm, "nib") ? 4 : 0;
            if (h->field_mode == 3) {
                if ((rc = alloc_zero(h, (void**)&h->d_field_brick, V))) return rc;
            } else if (h->field_mode == 4) {
                if ((rc = alloc_zero(h, (void**)&h->d_field_nib, V / 2))) return rc;
            } else if (h->field_mode) {
                cudaChannelFormatDesc cd = cudaCreateChannelDesc<unsigned char>();
                CU(cudaMalloc3DArray(&h->field_arr, &cd, make_cudaExtent((size_t)cfg->nx, (size_t)cfg->ny, (size_t)cfg->nz)));
                cudaResourceDesc rd; memset(&rd, 0, sizeof rd);
                rd.resType = cudaResourceTypeArray; rd.res.array.array = h->field_arr;
                for (int i = 0; i < 2; i++) {
                    cudaTextureDesc td; memset(&td, 0, sizeof td);
                    td.filterMode = cudaFilterModePoint; td.readMode = cudaReadModeElementType; td.normalizedCoords = i;
                    for (int a = 0; a < 3; a++) td.addressMode[a] = i ? cudaAddressModeWrap : cudaAddressModeClamp;
                    CU(cudaCreateTextureObject(&h->field_tex[i], &rd, &td, nullptr));
                }
            }
        }
        if ((rc = alloc_zero(h, (void**)&h->d_counters, 8 * sizeof(unsigned long long)))) return rc;
        std::vector<float> lut((size_t)VXRT_LUT_SCREEN + cfg->width + cfg->height);
        CU(cudaMalloc((void**)&h->d_lut, lut.size() * sizeof(float)));
        diffuse_luts(lut.data());
        value_tables(lut.data(), cfg->width, cfg->height);
        CU(cudaMemcpyAsync(h->d_lut, lut.data(), lut.size() * sizeof(float), cudaMemcpyHostToDevice, h->stream));
        CU(cudaStreamSynchronize(h->stream));
        if ((rc = alloc_zero(h, (void**)&h->d_lighting, P * 8))) return rc;
        if ((rc = alloc_zero(h, (void**)&h->d_depth, P * 2))) return rc;
        if ((rc = alloc_zero(h, (void**)&h->d_normal, P))) return rc;
        if ((rc = alloc_zero(h, (void**)&h->d_albedo, P * 4))) return rc;
        if ((rc = alloc_zero(h, (void**)&h->d_emission, P * 4))) return rc;
        if ((rc = alloc_zero(h, (void**)&h->d_fog, P * 4))) return rc;
        if ((rc = alloc_zero(h, (void**)&h->d_out_buf[0], P * 4))) return rc;
        if ((rc = alloc_zero(h, (void**)&h->d_out_buf[1], P * 4))) return rc;
        h->d_output = h->d_out_buf[0];
        if (cfg->flags & VXRT_FLAG_DENOISE) if ((rc = alloc_zero(h, (void**)&h->d_pong, P * 8))) return rc;
        if (cfg->flags & VXRT_FLAG_HIT_BUFFERS) {
            if ((rc = alloc_zero(h, (void**)&h->d_hit_voxel, P * 12))) return rc;
            if ((rc = alloc_zero(h, (void**)&h->d_hit_face, P))) return rc;
            if ((rc = alloc_zero(h, (void**)&h->d_hit_word, P * 4))) return rc;
            if ((rc = alloc_zero(h, (void**)&h->d_hit_t, P * 4))) return rc;
            if ((rc = alloc_zero(h, (void**)&h->d_steps, P * 4))) return rc;
            if ((rc = alloc_zero(h, (void**)&h->d_rays, P))) return rc;
            if ((rc = alloc_zero(h, (void**)&h->d_lighting_raw, P * 8))) return rc;
        }
        if (cfg->mode == VXRT_MODE_FULL && !(cfg->flags & (VXRT_FLAG_HIT_BUFFERS | VXRT_FLAG_REPROJECT)) && cfg->width <= 4096 && cfg->height <= 4096)
            { if ((rc = alloc_zero(h, &h->d_pool_ctx, P * 16))) return rc;       // k_trace_pool's pixel records
              int sms = 148; cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, cfg->device);
              if ((rc = alloc_zero(h, &h->d_pool_rings, vxrt_trace_pool_ring_bytes(sms)))) return rc;
              // the wavefront form's entry / ray / result arrays (trace_wave.cu): one entry per pixel of the band at most
              const size_t band = (size_t)(h->trace_row1 - h->trace_row0) * cfg->width;
              if ((rc = alloc_zero(h, &h->d_wave, vxrt_trace_wave_bytes(band)))) return rc;
              vxrt_trace_wave_carve(&h->wave, h->d_wave, band); }
        if (cfg->flags & VXRT_FLAG_ACCUMULATE) {
            if (owner) {                                                 // one running sum for the whole group of handles
                if (!owner->d_accum || owner->cfg.width != cfg->width || owner->cfg.height != cfg->height || owner->row0 != h->row0 || owner->row1 != h->row1) {
                    h->err = "vxrt_create_shared: VXRT_FLAG_ACCUMULATE needs an owner created with it and the same image and rows";
                    return VXRT_ERR_INVALID;
                }
                h->d_accum = owner->d_accum; h->accum_borrowed = true;
            } else {
                if ((rc = alloc_zero(h, (void**)&h->d_accum, P * 16))) return rc;
                CU(cudaEventCreateWithFlags(&h->ev_accum, cudaEventDisableTiming));
            }
        }
        if (cfg->flags & VXRT_FLAG_REPROJECT) {
            for (int i = 0; i < 2; i++) {
                if ((rc = alloc_zero(h, (void**)&h->d_hit_pos[i], P * 16))) return rc;
                if ((rc = alloc_zero(h, (void**)&h->d_hist[i], P * 16))) return rc;
            }
            if ((rc = alloc_zero(h, (void**)&h->d_completed, P * 8))) return rc;
        }
        CU(cudaStreamSynchronize(h->stream));
        return VXRT_OK;
    };
    const int rc = init();
    if (rc != VXRT_OK) {
        if (owner) h->d_materials = nullptr, h->d_minefield = nullptr, h->d_noise = nullptr;
        if (h->accum_borrowed) h->d_accum = nullptr;
        g_create_error = h->err; vxrt_destroy(h); return rc;
    }
    if (owner) { h->world_owner = owner; owner->sharers.push_back(h); h->world_seen = owner->world_version; }
    { const int lrc = set_l2_window(h, h->stream); if (lrc) { g_create_error = h->err; vxrt_destroy(h); return lrc; } }
    *out = h;
    return VXRT_OK;
}

int vxrt_create(const VxrtConfig* cfg, VxrtHandle** out) { return create_impl(cfg, nullptr, out); }

int vxrt_create_shared(const VxrtConfig* cfg, VxrtHandle* owner, VxrtHandle** out) {
    if (!owner) { g_create_error = "vxrt_create_shared: owner is NULL"; if (out) *out = nullptr; return VXRT_ERR_INVALID; }
    cudaSetDevice(owner->cfg.device);
    cudaStreamSynchronize(owner->stream);        // whatever the owner has uploaded so far is visible to the new handle
    return create_impl(cfg, owner, out);
}

void vxrt_destroy(VxrtHandle* h) {
    if (!h) return;
    cudaSetDevice(h->cfg.device);
    if (h->stream) cudaStreamSynchronize(h->stream);
    if (h->copy_stream) cudaStreamSynchronize(h->copy_stream);
    vxrt_comm_release(h);
    vxrt_peer_release(h);
    if (h->world_owner) {                        // borrowed world: unregister, free nothing of it
        auto& v = h->world_owner->sharers;
        for (size_t i = 0; i < v.size(); i++) if (v[i] == h) { v.erase(v.begin() + i); break; }
        h->d_materials = nullptr; h->d_minefield = nullptr; h->d_noise = nullptr;
        if (h->accum_borrowed) { h->d_accum = nullptr; if (h->world_owner->accum_last == h) h->world_owner->accum_last = nullptr; }
    }
    for (VxrtHandle* s : h->sharers) {           // an owner destroyed first leaves its borrowers without a world
        cudaStreamSynchronize(s->stream);
        s->world_owner = nullptr; s->d_materials = nullptr; s->d_minefield = nullptr; s->d_noise = nullptr;
        s->have_world = s->have_noise = false;
        if (s->accum_borrowed) s->d_accum = nullptr;     // and without a running sum: their frames fail with VXRT_ERR_STATE
    }
    for (int i = 0; i < 2; i++) if (h->field_tex[i] && !h->world_owner) cudaDestroyTextureObject(h->field_tex[i]);
    if (h->field_arr && !h->world_owner) cudaFreeArray(h->field_arr);
    if (h->d_field_brick && !h->world_owner) cudaFree(h->d_field_brick);
    if (h->d_field_nib && !h->world_owner) cudaFree(h->d_field_nib);
    for (int i = 0; i < 2; i++) { if (h->graphs[i].exec) cudaGraphExecDestroy(h->graphs[i].exec); if (h->graphs[i].graph) cudaGraphDestroy(h->graphs[i].graph); }
    if (h->ev_accum) cudaEventDestroy(h->ev_accum);
    if (h->ev_world) cudaEventDestroy(h->ev_world);
    void* ptrs[] = {h->d_materials, h->d_minefield_base, h->d_noise, h->d_lut, h->d_dirty, h->d_edits, h->d_edit_table, h->d_lighting, h->d_pong,
                    h->d_lighting_raw, h->d_depth, h->d_normal, h->d_albedo, h->d_emission, h->d_fog, h->d_out_buf[0], h->d_out_buf[1],
                    h->d_hit_voxel, h->d_hit_face, h->d_hit_word, h->d_hit_t, h->d_steps, h->d_rays, h->d_accum, h->d_counters, h->d_pool_ctx, h->d_pool_rings,
                    h->d_hit_pos[0], h->d_hit_pos[1], h->d_hist[0], h->d_hist[1], h->d_completed, h->d_wave};
    for (void* p : ptrs) if (p) cudaFree(p);
    for (int i = 0; i < 2; i++) { if (h->h_edits[i]) cudaFreeHost(h->h_edits[i]); if (h->ev_edits[i]) cudaEventDestroy(h->ev_edits[i]); }
    for (int i = 0; i < 5; i++) if (h->ev[i]) cudaEventDestroy(h->ev[i]);
    for (int i = 0; i < 2; i++) {
        if (h->ev_frame_done[i]) cudaEventDestroy(h->ev_frame_done[i]);
        if (h->ev_copy_done[i]) cudaEventDestroy(h->ev_copy_done[i]);
    }
    if (h->copy_stream) cudaStreamDestroy(h->copy_stream);
    if (h->stream) cudaStreamDestroy(h->stream);
    delete h;
}

// ---- world ------------------------------------------------------------------------------------

int vxrt_upload_chunk(VxrtHandle* h, int cx, int cy, int cz, const uint32_t* materials, const uint8_t* minefield) {
    if (!h) return VXRT_ERR_INVALID;
    if (!materials) return fail(h, VXRT_ERR_INVALID, "vxrt_upload_chunk: materials is NULL");
    if (cx < 0 || cy < 0 || cz < 0 || cx >= (h->cfg.nx >> 6) || cy >= (h->cfg.ny >> 6) || cz >= (h->cfg.nz >> 6))
        return fail(h, VXRT_ERR_INVALID, "vxrt_upload_chunk: chunk coordinate outside the window");
    CU(cudaSetDevice(h->cfg.device));
    { const int wrc = world_write_begin(h, "vxrt_upload_chunk"); if (wrc) return wrc; }
    int rc;
    if ((rc = upload_box(h, h->d_materials, materials, cx * 64, cy * 64, cz * 64, 64, 64, 64))) return rc;
    if (minefield) {
        if ((rc = upload_box(h, h->d_minefield, minefield, cx * 64, cy * 64, cz * 64, 64, 64, 64))) return rc;
    } else {
        if ((rc = mark_dirty_box(h, cx * 64, cy * 64, cz * 64, 64, 64, 64))) return rc;
        if ((rc = rebuild_dirty(h))) return rc;
    }
    CU(cudaStreamSynchronize(h->stream));     // caller may free its buffers on return
    h->have_world = true;
    const int box[6] = {cx * 64, cy * 64, cz * 64, 64, 64, 64};
    return world_write_end(h, box);
}

int vxrt_upload_world(VxrtHandle* h, const uint32_t* materials, const uint8_t* minefield) {
    if (!h) return VXRT_ERR_INVALID;
    if (!materials) return fail(h, VXRT_ERR_INVALID, "vxrt_upload_world: materials is NULL");
    CU(cudaSetDevice(h->cfg.device));
    { const int wrc = world_write_begin(h, "vxrt_upload_world"); if (wrc) return wrc; }
    const size_t V = voxels(h);
    CU(cudaMemcpyAsync(h->d_materials, materials, V * 4, cudaMemcpyHostToDevice, h->stream));
    if (minefield) CU(cudaMemcpyAsync(h->d_minefield, minefield, V, cudaMemcpyHostToDevice, h->stream));
    else CU(vxrt_launch_build_minefield(h->d_materials, h->d_minefield, h->cfg.nx, h->cfg.ny, h->cfg.nz, nullptr, h->stream));
    CU(cudaStreamSynchronize(h->stream));
    h->have_world = true;
    return world_write_end(h);
}

int vxrt_upload_slice(VxrtHandle* h, int axis, int offset, const uint32_t* materials, const uint8_t* minefield) {
    if (!h) return VXRT_ERR_INVALID;
    if (!materials) return fail(h, VXRT_ERR_INVALID, "vxrt_upload_slice: materials is NULL");
    if (axis < 0 || axis > 2) return fail(h, VXRT_ERR_INVALID, "vxrt_upload_slice: axis must be 0, 1 or 2");
    const int dims[3] = {h->cfg.nx, h->cfg.ny, h->cfg.nz};
    if (offset < 0 || offset % 16 != 0 || offset + 16 > dims[axis])
        return fail(h, VXRT_ERR_INVALID, "vxrt_upload_slice: offset must be a multiple of 16 inside the window");
    int e[3] = {dims[0], dims[1], dims[2]}, o[3] = {0, 0, 0};
    e[axis] = 16; o[axis] = offset;                           // constants.rs:30 SLICE_SIZE
    CU(cudaSetDevice(h->cfg.device));
    { const int wrc = world_write_begin(h, "vxrt_upload_slice"); if (wrc) return wrc; }
    int rc;
    if ((rc = upload_box(h, h->d_materials, materials, o[0], o[1], o[2], e[0], e[1], e[2]))) return rc;
    if (minefield) {
        if ((rc = upload_box(h, h->d_minefield, minefield, o[0], o[1], o[2], e[0], e[1], e[2]))) return rc;
    } else {
        if ((rc = mark_dirty_box(h, o[0], o[1], o[2], e[0], e[1], e[2]))) return rc;
        if ((rc = rebuild_dirty(h))) return rc;
    }
    CU(cudaStreamSynchronize(h->stream));
    h->have_world = true;
    // a rebuilt slab changes whole 64^3 chunks of the occupancy field: the box is widened to chunk edges on the slab axis
    int box[6] = {o[0], o[1], o[2], e[0], e[1], e[2]};
    box[axis] = (offset / 64) * 64; box[3 + axis] = 64;
    return world_write_end(h, box);
}

int vxrt_generate_world(VxrtHandle* h, int kind, uint32_t seed) {
    if (!h) return VXRT_ERR_INVALID;
    if (kind < 0 || kind > 1) return fail(h, VXRT_ERR_INVALID, "vxrt_generate_world: kind must be 0 (heightmap) or 1 (biome)");
    CU(cudaSetDevice(h->cfg.device));
    { const int wrc = world_write_begin(h, "vxrt_generate_world"); if (wrc) return wrc; }
    CU(vxrt_launch_generate(h->d_materials, h->cfg.nx, h->cfg.ny, h->cfg.nz, kind, seed, h->stream));
    CU(vxrt_launch_build_minefield(h->d_materials, h->d_minefield, h->cfg.nx, h->cfg.ny, h->cfg.nz, nullptr, h->stream));
    CU(cudaStreamSynchronize(h->stream));
    h->have_world = true;
    return world_write_end(h);
}

int vxrt_rebuild_minefield(VxrtHandle* h) {
    if (!h) return VXRT_ERR_INVALID;
    CU(cudaSetDevice(h->cfg.device));
    { const int wrc = world_write_begin(h, "vxrt_rebuild_minefield"); if (wrc) return wrc; }
    CU(vxrt_launch_build_minefield(h->d_materials, h->d_minefield, h->cfg.nx, h->cfg.ny, h->cfg.nz, nullptr, h->stream));
    CU(cudaStreamSynchronize(h->stream));
    return world_write_end(h);
}

int vxrt_read_world(VxrtHandle* h, uint32_t* materials, uint8_t* minefield) {
    if (!h) return VXRT_ERR_INVALID;
    if (h->world_owner) return fail(h, VXRT_ERR_STATE, "vxrt_read_world: this handle borrows its world (vxrt_create_shared); read through the owner");
    CU(cudaSetDevice(h->cfg.device));
    const size_t V = voxels(h);
    if (materials) CU(cudaMemcpyAsync(materials, h->d_materials, V * 4, cudaMemcpyDeviceToHost, h->stream));
    if (minefield) CU(cudaMemcpyAsync(minefield, h->d_minefield, V, cudaMemcpyDeviceToHost, h->stream));
    CU(cudaStreamSynchronize(h->stream));
    return VXRT_OK;
}

int vxrt_validate_chunk(VxrtHandle* h, int cx, int cy, int cz, const uint8_t* stored_minefield, uint64_t* mismatches, int64_t* first_offset) {
    if (!h) return VXRT_ERR_INVALID;
    if (!stored_minefield || !mismatches) return fail(h, VXRT_ERR_INVALID, "vxrt_validate_chunk: NULL argument");
    if (cx < 0 || cy < 0 || cz < 0 || cx >= (h->cfg.nx >> 6) || cy >= (h->cfg.ny >> 6) || cz >= (h->cfg.nz >> 6))
        return fail(h, VXRT_ERR_INVALID, "vxrt_validate_chunk: chunk coordinate outside the window");
    const VxrtHandle* wo = world_of(h);
    if (!wo->have_world) return fail(h, VXRT_ERR_STATE, "vxrt_validate_chunk: no world uploaded");
    CU(cudaSetDevice(h->cfg.device));
    uint8_t* d_stored = nullptr; unsigned long long* d_out = nullptr;
    CU(cudaMalloc((void**)&d_stored, VXRT_CHUNK * VXRT_CHUNK * VXRT_CHUNK));
    cudaError_t e = cudaMalloc((void**)&d_out, 16);
    unsigned long long init[2] = {0ull, ~0ull}, res[2] = {0, 0};
    if (e == cudaSuccess) e = cudaMemcpyAsync(d_stored, stored_minefield, VXRT_CHUNK * VXRT_CHUNK * VXRT_CHUNK, cudaMemcpyHostToDevice, h->stream);
    if (e == cudaSuccess) e = cudaMemcpyAsync(d_out, init, 16, cudaMemcpyHostToDevice, h->stream);
    if (e == cudaSuccess) e = vxrt_launch_compare_chunk(h->d_minefield, d_stored, h->cfg.nx, h->cfg.ny, cx, cy, cz, d_out, h->stream);
    if (e == cudaSuccess) e = cudaMemcpyAsync(res, d_out, 16, cudaMemcpyDeviceToHost, h->stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(h->stream);
    cudaFree(d_stored); if (d_out) cudaFree(d_out);
    CU(e);
    *mismatches = res[0];
    if (first_offset) *first_offset = res[0] ? (int64_t)res[1] : -1;
    return VXRT_OK;
}

int vxrt_upload_blue_noise(VxrtHandle* h, const uint8_t* rgba, size_t nbytes) {
    if (!h) return VXRT_ERR_INVALID;
    if (!rgba) return fail(h, VXRT_ERR_INVALID, "vxrt_upload_blue_noise: rgba is NULL");
    if (nbytes != h->h_noise.size()) return fail(h, VXRT_ERR_SIZE, "vxrt_upload_blue_noise: expected 512*512*4 bytes");
    CU(cudaSetDevice(h->cfg.device));
    { const int wrc = world_write_begin(h, "vxrt_upload_blue_noise"); if (wrc) return wrc; }
    memcpy(h->h_noise.data(), rgba, nbytes);
    CU(cudaMemcpyAsync(h->d_noise, h->h_noise.data(), nbytes, cudaMemcpyHostToDevice, h->stream));
    CU(cudaStreamSynchronize(h->stream));
    h->have_noise = true;
    return world_write_end(h);
}

// ---- edits ------------------------------------------------------------------------------------

int vxrt_apply_edits_device(VxrtHandle* h, const void* d_edits, size_t n) {
    if (!h) return VXRT_ERR_INVALID;
    if (n && !d_edits) return fail(h, VXRT_ERR_INVALID, "vxrt_apply_edits_device: d_edits is NULL");
    if (h->world_owner) return fail(h, VXRT_ERR_STATE, "vxrt_apply_edits_device: this handle borrows its world (vxrt_create_shared); write through the owner");
    if (!h->have_world) return fail(h, VXRT_ERR_STATE, "vxrt_apply_edits: no world uploaded");
    if (n == 0) return VXRT_OK;
    CU(cudaSetDevice(h->cfg.device));
    { const int wrc = world_write_begin(h, "vxrt_apply_edits_device"); if (wrc) return wrc; }
    if (h->edit_table_slots < 4 * n) {
        size_t slots = 4096; while (slots < 4 * n) slots *= 2;
        CU(cudaStreamSynchronize(h->stream));
        if (h->d_edit_table) CU(cudaFree(h->d_edit_table));
        h->d_edit_table = nullptr; h->edit_table_slots = 0;
        CU(cudaMalloc(&h->d_edit_table, slots * 12));
        h->edit_table_slots = slots;
    }
    CU(vxrt_launch_apply_edits(h->d_materials, h->d_dirty, (const VxrtEdit*)d_edits, n, h->cfg.nx, h->cfg.ny, h->cfg.nz,
                               h->d_edit_table, h->edit_table_slots, h->stream));
    { const int rc = rebuild_dirty(h); if (rc) return rc; }
    return world_write_end(h);
}

int vxrt_apply_edits(VxrtHandle* h, const VxrtEdit* edits, size_t n) {
    if (!h) return VXRT_ERR_INVALID;
    if (n && !edits) return fail(h, VXRT_ERR_INVALID, "vxrt_apply_edits: edits is NULL");
    if (h->world_owner) return fail(h, VXRT_ERR_STATE, "vxrt_apply_edits: this handle borrows its world (vxrt_create_shared); write through the owner");
    if (!h->have_world) return fail(h, VXRT_ERR_STATE, "vxrt_apply_edits: no world uploaded");
    if (n == 0) return VXRT_OK;
    CU(cudaSetDevice(h->cfg.device));
    if (n > h->edits_cap) {
        if (h->d_edits) CU(cudaFree(h->d_edits));
        h->d_edits = nullptr; h->edits_cap = 0;
        size_t cap = 1024; while (cap < n) cap *= 2;
        CU(cudaMalloc(&h->d_edits, cap * sizeof(VxrtEdit)));
        h->edits_cap = cap;
    }
    // the host list may be reused on return: it is copied into one of two page-locked staging buffers here, and the
    // device copy runs from there in stream order - no wait for the frames already queued on the stream
    if (n > h->h_edits_cap) {
        CU(cudaStreamSynchronize(h->stream));
        size_t cap = 1024; while (cap < n) cap *= 2;
        for (int i = 0; i < 2; i++) {
            if (h->h_edits[i]) CU(cudaFreeHost(h->h_edits[i]));
            h->h_edits[i] = nullptr;
            CU(cudaMallocHost(&h->h_edits[i], cap * sizeof(VxrtEdit)));
            if (!h->ev_edits[i]) CU(cudaEventCreateWithFlags(&h->ev_edits[i], cudaEventDisableTiming));
        }
        h->h_edits_cap = cap;
    }
    const int slot = h->edit_slot ^= 1;
    CU(cudaEventSynchronize(h->ev_edits[slot]));          // the copy out of this buffer two calls ago (done long since)
    memcpy(h->h_edits[slot], edits, n * sizeof(VxrtEdit));
    CU(cudaMemcpyAsync(h->d_edits, h->h_edits[slot], n * sizeof(VxrtEdit), cudaMemcpyHostToDevice, h->stream));
    CU(cudaEventRecord(h->ev_edits[slot], h->stream));
    return vxrt_apply_edits_device(h, h->d_edits, n);
}

// ---- per frame --------------------------------------------------------------------------------

int vxrt_set_frame(VxrtHandle* h, const VxrtFrameParams* p) {
    if (!h) return VXRT_ERR_INVALID;
    if (!p) return fail(h, VXRT_ERR_INVALID, "vxrt_set_frame: params is NULL");
    h->frame = *p;
    h->have_frame = true;
    return VXRT_OK;
}

int vxrt_reset_accumulation(VxrtHandle* h) {
    if (!h) return VXRT_ERR_INVALID;
    if (!h->d_accum) return fail(h, VXRT_ERR_STATE, "vxrt_reset_accumulation: handle created without VXRT_FLAG_ACCUMULATE");
    CU(cudaSetDevice(h->cfg.device));
    int rc;
    if ((rc = accum_wait(h))) return rc;
    CU(cudaMemsetAsync(h->d_accum, 0, pixels(h) * 16, h->stream));
    if ((rc = accum_written(h))) return rc;
    accum_keeper(h)->frames_accumulated = 0;
    return VXRT_OK;
}

int vxrt_reset_history(VxrtHandle* h) {
    if (!h) return VXRT_ERR_INVALID;
    if (!h->d_hist[0]) return fail(h, VXRT_ERR_STATE, "vxrt_reset_history: handle created without VXRT_FLAG_REPROJECT");
    CU(cudaSetDevice(h->cfg.device));
    // zeroed records read as "face 0 at the origin, 0 samples": a match contributes nothing
    for (int i = 0; i < 2; i++) {
        CU(cudaMemsetAsync(h->d_hit_pos[i], 0, pixels(h) * 16, h->stream));
        CU(cudaMemsetAsync(h->d_hist[i], 0, pixels(h) * 16, h->stream));
    }
    return VXRT_OK;
}

// The device work of one frame on the handle's stream: counter reset, trace pass, [reprojection], six denoise passes,
// finalize, with the four events vxrt_frame_stats reads.  Issued directly, or once into a stream capture (capturing: the host-side
// read-back wait is the caller's, events are recorded as graph nodes).
static int issue_frame(VxrtHandle* h, TraceParams& P, const VxrtFrameParams& f, bool capturing) {
    int prc;
    const unsigned evf = capturing ? cudaEventRecordExternal : cudaEventRecordDefault;
    CU(cudaMemsetAsync(h->d_counters, 0, 8 * sizeof(unsigned long long), h->stream));
    CU(cudaEventRecordWithFlags(h->ev[0], h->stream, evf));
    if (!h->peer_on) {
        CU(vxrt_launch_trace(P, h->stream));
    } else {
        // peer.cu: the rows a neighbour's denoise needs are traced first and stored into its images while this
        // band's interior is still being traced
        int32_t iv[2 * VXRT_PEER_MAX_LINKS];
        // measured at N=2 (profiles/r1_m): splitting the trace pass into boundary + interior launches costs more in partial
        // waves than the earlier push wins, so the whole band is traced first; the split stays behind a switch
        static const bool boundary_first = getenv("VXRT_PEER_BOUNDARY_FIRST") != nullptr;
        const int n = boundary_first ? vxrt_peer_boundary_rows(h, iv) : 0;
        for (int i = 0; i < n; i++) { P.row0 = iv[2 * i]; P.row1 = iv[2 * i + 1]; CU(vxrt_launch_trace(P, h->stream)); }
        if (n && (prc = vxrt_peer_push_halo(h))) return prc;
        int at = h->row0;
        for (int i = 0; i <= n; i++) {
            const int gap_end = i < n ? iv[2 * i] : h->row1;
            if (gap_end > at) { P.row0 = at; P.row1 = gap_end; CU(vxrt_launch_trace(P, h->stream)); }
            if (i < n) at = iv[2 * i + 1];
        }
        if (!n && (prc = vxrt_peer_push_halo(h))) return prc;
    }
    CU(cudaEventRecordWithFlags(h->ev[1], h->stream, evf));
    if (h->d_accum && (prc = accum_written(h))) return prc;
    if (h->peer_on && (prc = vxrt_peer_wait_halo(h))) return prc;
    if (h->d_lighting_raw)
        CU(cudaMemcpyAsync(h->d_lighting_raw, h->d_lighting, pixels(h) * 8, cudaMemcpyDeviceToDevice, h->stream));
    if (h->d_hist[0]) {                                                  // temporal.cu; timed with the denoise passes
        ReprojectParams R;
        memset(&R, 0, sizeof R);
        R.oox = f.old_origin[0]; R.ooy = f.old_origin[1]; R.ooz = f.old_origin[2];
        R.c0x = f.old_transform_c0[0]; R.c0y = f.old_transform_c0[1]; R.c0z = f.old_transform_c0[2];
        R.c1x = f.old_transform_c1[0]; R.c1y = f.old_transform_c1[1]; R.c1z = f.old_transform_c1[2];
        R.c2x = f.old_transform_c2[0]; R.c2y = f.old_transform_c2[1]; R.c2z = f.old_transform_c2[2];
        R.width = h->cfg.width; R.height = h->cfg.height;
        R.lighting = h->d_lighting; R.completed = h->d_completed;
        R.hit_pos_cur = h->d_hit_pos[h->hist_idx]; R.hit_pos_prev = h->d_hit_pos[h->hist_idx ^ 1];
        R.hist_prev = h->d_hist[h->hist_idx ^ 1]; R.hist_cur = h->d_hist[h->hist_idx];
        CU(vxrt_launch_reproject(R, h->stream));
    }
    bool fold = false;
    if (h->cfg.flags & VXRT_FLAG_DENOISE) {
        // pipeline.rs:98-115: even passes lighting -> pong, odd passes pong -> lighting.  A band shrinks
        // by 3*size rows per pass towards [row0,row1) so that only valid rows are ever read.
        // passes 2..6 follow a denoise launch directly: programmatic dependent launch (post.cu) lets a pass become
        // resident while its predecessor drains
        static const bool pdl = getenv("VXRT_NO_PDL") == nullptr;
        static const bool want_fold = getenv("VXRT_FOLD") != nullptr;
        // VXRT_FOLD=1: finalize.comp rides on the last pass (one launch, one grid tail and a read of the lighting image less).
        // Measured on the B200 (r2_h, C2): six passes + finalize 0.1915 ms apart, 0.1938 ms folded - the last pass's tiles
        // (128 x 128 pixels at dilation 16) finish later with the extra loads and the flipped-row stores than the 0.016 ms
        // finalize grid costs on its own - so the separate launch stays the default.
        fold = (h->cfg.flags & VXRT_FLAG_FINALIZE) && !h->peer_on && want_fold;
        for (int k = 0; k < 6; k++) {
            const int r0 = h->plan[2 + 2 * k], r1 = h->plan[3 + 2 * k];
            const bool odd = (k & 1) != 0;
            const int swap = (odd && (h->cfg.flags & VXRT_FLAG_PONG_SWAP)) ? 1 : 0;
            VxrtFinalizeFold ff = {h->d_albedo, h->d_emission, h->d_fog, h->d_noise, h->d_output, h->row0, h->row1};
            if (!capturing && k == 5 && fold && h->copy_pending[h->out_idx]) {            // the output image may still be on its way to the host
                CU(cudaStreamWaitEvent(h->stream, h->ev_copy_done[h->out_idx], 0));
                h->copy_pending[h->out_idx] = false;
            }
            CU(vxrt_launch_denoise(odd ? h->d_pong : h->d_lighting, odd ? h->d_lighting : h->d_pong, h->d_depth, h->d_normal,
                                   h->cfg.width, h->cfg.height, kDenoiseSizes[k], swap, r0, r1, h->stream, k > 0 && pdl,
                                   k == 5 && fold ? &ff : nullptr));
        }
    }
    if (h->peer_on && (prc = vxrt_peer_after_denoise(h))) return prc;
    CU(cudaEventRecordWithFlags(h->ev[2], h->stream, evf));
    if (!capturing && h->copy_pending[h->out_idx]) {
        CU(cudaStreamWaitEvent(h->stream, h->ev_copy_done[h->out_idx], 0));
        h->copy_pending[h->out_idx] = false;
    }
    if ((h->cfg.flags & VXRT_FLAG_FINALIZE) && !fold) {
        VxrtPeerTail tail;
        if (h->peer_on) tail = vxrt_peer_finalize_tail(h);
        CU(vxrt_launch_finalize(h->d_albedo, h->d_emission, h->d_fog, h->d_lighting, h->d_depth, h->d_noise,
                                h->peer_on ? vxrt_peer_output(h, h->out_idx) : h->d_output,       // peer mode: straight into the root's image
                                h->cfg.width, h->cfg.height, h->row0, h->row1, h->stream, h->peer_on ? &tail : nullptr));
    }
    if (h->peer_on && (prc = vxrt_peer_end_frame(h))) return prc;
    CU(cudaEventRecordWithFlags(h->ev[3], h->stream, evf));
    return VXRT_OK;
}

int vxrt_render_frame(VxrtHandle* h) {
    if (!h) return VXRT_ERR_INVALID;
    const VxrtHandle* wo = world_of(h);
    if (!wo->have_world) return fail(h, VXRT_ERR_STATE, "vxrt_render_frame: no world uploaded");
    if (!wo->have_noise) return fail(h, VXRT_ERR_STATE, "vxrt_render_frame: no blue noise uploaded");
    if (!h->have_frame) return fail(h, VXRT_ERR_STATE, "vxrt_render_frame: vxrt_set_frame not called");
    CU(cudaSetDevice(h->cfg.device));
    const VxrtFrameParams& f = h->frame;
    TraceParams P;
    memset(&P, 0, sizeof P);
    P.ox = f.origin[0]; P.oy = f.origin[1]; P.oz = f.origin[2];
    P.fx = f.forward[0]; P.fy = f.forward[1]; P.fz = f.forward[2];
    P.ux = f.up[0]; P.uy = f.up[1]; P.uz = f.up[2];
    P.rx = f.right[0]; P.ry = f.right[1]; P.rz = f.right[2];
    P.lrx = (float)f.rotation[0]; P.lry = (float)f.rotation[1]; P.lrz = (float)f.rotation[2];   // raytrace.comp:104
    H3 sd, sl;
    sun_constants(f.sun_angle, &sd, &sl);
    P.sun_x = sd.x; P.sun_y = sd.y; P.sun_z = sd.z; P.sl_r = sl.x; P.sl_g = sl.y; P.sl_b = sl.z;
    {   // sample_sky's per-frame terms (raytrace.comp:272-274, :279); colour only
        float s = ((sl.x + sl.y) + sl.z) * 0.2f - 0.02f;
        s = s < 0.0f ? 0.0f : (s > 1.0f ? 1.0f : s);
        P.sky_horizon_exp_half = 0.5f * (40.0f * (1.0f - s) + 10.0f * s);
        P.sky_halo_exp = 5.0f * (1.0f - s) + 1.0f * s;
        P.sky_bright_scale = s > 0.1f ? s : 0.1f;
    }
    {   // raytrace.comp:298-303: texel (seed % 512, seed / 512), clamp-to-edge (render_data.rs:110-128)
        uint32_t ix = f.seed % VXRT_NOISE_SIZE, iy = f.seed / VXRT_NOISE_SIZE;
        if (iy > VXRT_NOISE_SIZE - 1) iy = VXRT_NOISE_SIZE - 1;
        const uint8_t* t = wo->h_noise.data() + ((size_t)iy * VXRT_NOISE_SIZE + ix) * 4;
        P.noise_base_x = ((float)t[0] / 255.0f) * 255.0f;
        P.noise_base_y = ((float)t[1] / 255.0f) * 255.0f;
    }
    P.width = h->cfg.width; P.height = h->cfg.height; P.row0 = h->trace_row0; P.row1 = h->trace_row1;
    P.own_row0 = h->row0; P.own_row1 = h->row1;
    P.nx = h->cfg.nx; P.ny = h->cfg.ny; P.nz = h->cfg.nz;
    P.lx = ilog2(P.nx); P.ly = ilog2(P.ny); P.lz = ilog2(P.nz);
    P.hx = (float)(P.nx / 2); P.hy = (float)(P.ny / 2); P.hz = (float)(P.nz / 2);
    P.max_steps = h->cfg.max_steps; P.mode = h->cfg.mode; P.flags = h->cfg.flags;
    {   // preconditions of the instruction-lean march (trace.cu (E5)): every coordinate the march can form stays
        // below 2^21 in magnitude and texel offsets fit 32 bits; otherwise the general kernel runs
        const float lim = 2097152.0f;
        bool ok = P.lx + P.ly + P.lz <= 32 && !(h->cfg.flags & VXRT_FLAG_GENERAL_MARCH) &&
                  P.max_steps <= 200000;                  // k_trace_fast packs 5 * max_steps into 20 bits
        const float o[3] = {P.ox, P.oy, P.oz}, l[3] = {P.lrx, P.lry, P.lrz}, d[3] = {(float)P.nx, (float)P.ny, (float)P.nz};
        for (int a = 0; a < 3; a++) ok = ok && fabsf(o[a]) + d[a] + 128.0f < lim && fabsf(l[a]) + d[a] + 128.0f < lim;
        // raytrace.comp:312-315 moves the start by (space / dir.y + 1e-4) * dir when -origin.y > H: unbounded for pixels
        // whose |dir.y| is tiny, so the bound on |origin| says nothing about the relocated start.  The camera is outside
        // the window there anyway: the general kernel renders those frames.
        if (-P.oy > P.hy) ok = false;
        P.fast = ok ? 1 : 0;
        for (int a = 0; a < 3; a++) sky_thresholds(l[a], d[a] * 0.5f, &P.sky_lo[a], &P.sky_hi[a]);
    }
    P.minefield = h->d_minefield; P.materials = h->d_materials; P.noise = h->d_noise; P.lut = h->d_lut;
    P.lighting = h->d_lighting; P.depth = h->d_depth; P.normal = h->d_normal; P.albedo = h->d_albedo;
    P.emission = h->d_emission; P.fog = h->d_fog;
    P.hit_voxel = h->d_hit_voxel; P.hit_face = h->d_hit_face; P.hit_word = h->d_hit_word; P.hit_t = h->d_hit_t;
    P.steps = h->d_steps; P.rays = h->d_rays;
    P.accum = h->d_accum;
    P.pool_ctx = (h->cfg.flags & VXRT_FLAG_RAY_HANDOFF) ? nullptr : h->d_pool_ctx; P.pool_rings = h->d_pool_rings;
    P.wave = h->wave; P.trace_form = h->trace_form;
    {   // AUTO: where the caller's own vxrt_frame_stats calls have shown long rays (an edited world: C5 marches 40 steps per ray, C3 14,
        // C4 18), the second bounce ray of a pixel is handed to a finished lane of its warp ((E16), trace_fast.cu) - same rays, same
        // images and counters; measured 0.966 -> 0.899 ms per C5 trace pass, and a loss below ~25 steps per ray
        static const bool no_auto = getenv("VXRT_NO_AUTO_HANDOFF") != nullptr;
        if (!no_auto && h->trace_form == VXRT_TRACE_AUTO && P.mode == VXRT_MODE_FULL && h->steps_per_ray > 28.0f && wo->field_mode != 1)
            P.flags |= VXRT_FLAG_RAY_HANDOFF;
    }
    if (h->d_hist[0]) h->hist_idx ^= 1;                                  // this frame's records; [hist_idx ^ 1] = the frame before
    P.hit_pos = h->d_hit_pos[h->hist_idx];
    P.counters = h->d_counters;
    P.one2 = 0x3F8000003F800000ull;
    P.field = wo->field_mode;
    P.field_tex = wo->field_mode == 3 ? (unsigned long long)wo->d_field_brick : wo->field_mode == 4 ? (unsigned long long)wo->d_field_nib : wo->field_mode ? (unsigned long long)wo->field_tex[wo->field_mode - 1] : 0ull;
    P.inx = 1.0f / (float)P.nx; P.iny = 1.0f / (float)P.ny; P.inz = 1.0f / (float)P.nz;

    // the image goes to the buffer an asynchronous read-back is not using (vxrt_read_output_async)
    h->out_idx ^= 1;
    h->d_output = h->d_out_buf[h->out_idx];
    int prc;
    if (h->peer_on && (prc = vxrt_peer_begin_frame(h, h->out_idx))) return prc;
    if (h->world_owner && h->world_seen != h->world_owner->world_version) {    // the owner wrote its world since this handle's last frame
        CU(cudaStreamWaitEvent(h->stream, h->world_owner->ev_world, 0));
        h->world_seen = h->world_owner->world_version;
    }
    if (h->d_accum && (prc = accum_wait(h))) return prc;                // shared running sum: after the previous frame's trace pass
    // (8) one CUDA graph per output image replays the frame - counter memset, trace, six denoise passes, finalize, the four
    // event records - with the trace node's parameters set per frame: one submission instead of ten.  Whatever orders the frame
    // against the host or other streams stays outside: the world / running-sum waits above, the read-back wait below.
    static const bool no_graph = getenv("VXRT_NO_GRAPH") != nullptr;
    const bool shared_sum = h->d_accum && (h->accum_borrowed || !h->sharers.empty());
    h->wave_frame = vxrt_trace_wave_applies(P);      // five launches with the frame's parameters: issued launch by launch
    if (!no_graph && !h->graph_failed && !h->peer_on && !h->d_hist[0] && !shared_sum && !h->wave_frame) {
        if (h->copy_pending[h->out_idx]) {
            CU(cudaStreamWaitEvent(h->stream, h->ev_copy_done[h->out_idx], 0));
            h->copy_pending[h->out_idx] = false;
        }
        VxrtHandle::FrameGraph& g = h->graphs[h->out_idx];
        dim3 grid, block;
        size_t smem = 0;
        const void* func = vxrt_trace_kernel(P, &grid, &block, &smem);
        const bool same = g.exec && g.func == func && g.grid.x == grid.x && g.grid.y == grid.y && g.block.x == block.x && g.smem == smem;
        if (!same) {
            if (g.exec) { cudaGraphExecDestroy(g.exec); g.exec = nullptr; }
            if (g.graph) { cudaGraphDestroy(g.graph); g.graph = nullptr; }   // (kept while its exec lives: the node handle names a node of it)
            cudaGraph_t graph = nullptr;
            cudaError_t e = cudaStreamBeginCapture(h->stream, cudaStreamCaptureModeThreadLocal);
            int brc = VXRT_OK;
            if (e == cudaSuccess) {
                brc = issue_frame(h, P, f, true);
                e = cudaStreamEndCapture(h->stream, &graph);
            }
            if (e == cudaSuccess && brc == VXRT_OK) {
                size_t n = 0;
                e = cudaGraphGetNodes(graph, nullptr, &n);
                std::vector<cudaGraphNode_t> nodes(n);
                if (e == cudaSuccess) e = cudaGraphGetNodes(graph, nodes.data(), &n);
                g.trace_node = nullptr;
                for (size_t i = 0; e == cudaSuccess && i < n; i++) {
                    cudaGraphNodeType t;
                    if (cudaGraphNodeGetType(nodes[i], &t) != cudaSuccess || t != cudaGraphNodeTypeKernel) continue;
                    cudaKernelNodeParams kp;
                    if (cudaGraphKernelNodeGetParams(nodes[i], &kp) == cudaSuccess && kp.func == func) g.trace_node = nodes[i];
                }
                if (e == cudaSuccess && !g.trace_node) e = cudaErrorUnknown;
                if (e == cudaSuccess) e = cudaGraphInstantiate(&g.exec, graph, 0);
            }
            if (e != cudaSuccess || brc != VXRT_OK) {                   // no graph on this handle: frames are issued launch by launch
                cudaGetLastError();
                if (g.exec) { cudaGraphExecDestroy(g.exec); g.exec = nullptr; }
                if (graph) cudaGraphDestroy(graph);
                h->graph_failed = true;
            } else { g.graph = graph; g.func = func; g.grid = grid; g.block = block; g.smem = smem; }
        } else {
            cudaKernelNodeParams kp;
            memset(&kp, 0, sizeof kp);
            void* args[1] = {(void*)&P};
            kp.func = (void*)func; kp.gridDim = grid; kp.blockDim = block; kp.sharedMemBytes = (unsigned)smem; kp.kernelParams = args;
            CU(cudaGraphExecKernelNodeSetParams(g.exec, g.trace_node, &kp));
        }
        if (!h->graph_failed) {
            CU(cudaGraphLaunch(g.exec, h->stream));
            if (h->d_accum) accum_keeper(h)->frames_accumulated++;
            h->rendered = true;
            return VXRT_OK;
        }
    }
    if ((prc = issue_frame(h, P, f, false))) return prc;
    if (h->d_accum) accum_keeper(h)->frames_accumulated++;
    h->rendered = true;
    return VXRT_OK;
}

// peer-memory frames (peer.cu): a wait on another rank's flag that ran into its bound leaves an error word; whoever synchronises
// with the frame next learns of it (the image may hold rows that never arrived) and the word is cleared
static int peer_check(VxrtHandle* h) {
    if (!h->peer_on) return VXRT_OK;
    uint32_t perr = 0;
    const int rc = vxrt_peer_error_word(h, &perr);
    if (rc) return rc;
    if (perr) {
        vxrt_peer_clear_error(h);
        return fail(h, VXRT_ERR_COMM, "peer-memory frame: a wait on another rank's flag timed out (ranks out of step?); the frame is incomplete");
    }
    return VXRT_OK;
}

int vxrt_set_trace_form(VxrtHandle* h, int form) {
    if (!h) return VXRT_ERR_INVALID;
    if (form < VXRT_TRACE_AUTO || form > VXRT_TRACE_WAVEFRONT) return fail(h, VXRT_ERR_INVALID, "vxrt_set_trace_form: unknown form");
    h->trace_form = form;
    return VXRT_OK;
}

int vxrt_launches_per_frame(const VxrtHandle* h) {
    if (!h) return VXRT_ERR_INVALID;
    const bool dn = (h->cfg.flags & VXRT_FLAG_DENOISE) != 0, fin = (h->cfg.flags & VXRT_FLAG_FINALIZE) != 0;
    const bool fold = dn && fin && !h->peer_on && getenv("VXRT_FOLD") != nullptr;
    return (h->wave_frame ? 5 : 1) + (dn ? 6 : 0) + (fin && !fold ? 1 : 0) + (h->d_hist[0] ? 1 : 0);
}

int vxrt_sync(VxrtHandle* h) {
    if (!h) return VXRT_ERR_INVALID;
    CU(cudaSetDevice(h->cfg.device));
    CU(cudaStreamSynchronize(h->stream));
    return peer_check(h);
}

// ---- output -----------------------------------------------------------------------------------

size_t vxrt_buffer_bytes(const VxrtHandle* h, int buffer_id) { return h ? buffer_of(h, buffer_id).bytes : 0; }

void* vxrt_device_buffer(VxrtHandle* h, int buffer_id) { return h ? buffer_of(h, buffer_id).ptr : nullptr; }

void* vxrt_cuda_stream(VxrtHandle* h) { return h ? (void*)h->stream : nullptr; }

int vxrt_read_buffer(VxrtHandle* h, int buffer_id, void* dst, size_t nbytes) {
    if (!h) return VXRT_ERR_INVALID;
    if (!dst) return fail(h, VXRT_ERR_INVALID, "vxrt_read_buffer: dst is NULL");
    const BufDesc b = buffer_of(h, buffer_id);
    if (!b.ptr) return fail(h, VXRT_ERR_STATE, "vxrt_read_buffer: buffer not allocated for this handle (check the create flags)");
    if (nbytes < b.bytes) return fail(h, VXRT_ERR_SIZE, "vxrt_read_buffer: host buffer too small");
    CU(cudaSetDevice(h->cfg.device));
    int rc;
    if (buffer_id == VXRT_BUF_ACCUM && (rc = accum_wait(h))) return rc;
    CU(cudaMemcpyAsync(dst, b.ptr, b.bytes, cudaMemcpyDeviceToHost, h->stream));
    CU(cudaStreamSynchronize(h->stream));
    return peer_check(h);
}

int vxrt_read_output(VxrtHandle* h, uint8_t* rgba8, size_t nbytes) { return vxrt_read_buffer(h, VXRT_BUF_OUTPUT, rgba8, nbytes); }

int vxrt_read_output_async(VxrtHandle* h, uint8_t* rgba8, size_t nbytes, uint64_t* ticket) {
    if (!h) return VXRT_ERR_INVALID;
    if (!rgba8 || !ticket) return fail(h, VXRT_ERR_INVALID, "vxrt_read_output_async: NULL argument");
    if (!h->rendered) return fail(h, VXRT_ERR_STATE, "vxrt_read_output_async: no frame rendered yet");
    const size_t bytes = pixels(h) * 4;
    if (nbytes < bytes) return fail(h, VXRT_ERR_SIZE, "vxrt_read_output_async: host buffer too small");
    CU(cudaSetDevice(h->cfg.device));
    const int i = h->out_idx;
    CU(cudaEventRecord(h->ev_frame_done[i], h->stream));               // everything queued so far (incl. an NCCL gather)
    CU(cudaStreamWaitEvent(h->copy_stream, h->ev_frame_done[i], 0));
    CU(cudaMemcpyAsync(rgba8, h->d_out_buf[i], bytes, cudaMemcpyDeviceToHost, h->copy_stream));
    CU(cudaEventRecord(h->ev_copy_done[i], h->copy_stream));
    h->copy_pending[i] = true;
    *ticket = ++h->tickets;
    h->ticket_of_buf[i] = *ticket;
    return VXRT_OK;
}

int vxrt_wait_output(VxrtHandle* h, uint64_t ticket) {
    if (!h) return VXRT_ERR_INVALID;
    if (ticket == 0 || ticket > h->tickets) return fail(h, VXRT_ERR_INVALID, "vxrt_wait_output: unknown ticket");
    CU(cudaSetDevice(h->cfg.device));
    // copies complete in ticket order on one stream: wait for the oldest recorded copy that is not older than `ticket`
    int pick = -1;
    for (int i = 0; i < 2; i++)
        if (h->ticket_of_buf[i] >= ticket && (pick < 0 || h->ticket_of_buf[i] < h->ticket_of_buf[pick])) pick = i;
    if (pick >= 0) CU(cudaEventSynchronize(h->ev_copy_done[pick]));
    return peer_check(h);
}

int vxrt_host_register(VxrtHandle* h, void* ptr, size_t nbytes) {
    if (!h) return VXRT_ERR_INVALID;
    if (!ptr || !nbytes) return fail(h, VXRT_ERR_INVALID, "vxrt_host_register: NULL or empty buffer");
    CU(cudaSetDevice(h->cfg.device));
    CU(cudaHostRegister(ptr, nbytes, cudaHostRegisterPortable));
    return VXRT_OK;
}

int vxrt_host_unregister(VxrtHandle* h, void* ptr) {
    if (!h) return VXRT_ERR_INVALID;
    if (!ptr) return fail(h, VXRT_ERR_INVALID, "vxrt_host_unregister: NULL");
    CU(cudaSetDevice(h->cfg.device));
    CU(cudaHostUnregister(ptr));
    return VXRT_OK;
}

int vxrt_row_costs(VxrtHandle* h, float* trace_cost, float* denoise_cost, size_t n) {
    if (!h) return VXRT_ERR_INVALID;
    if (!trace_cost || !denoise_cost || n < (size_t)h->cfg.height) return fail(h, VXRT_ERR_SIZE, "vxrt_row_costs: need height floats per array");
    if (!h->rendered) return fail(h, VXRT_ERR_STATE, "vxrt_row_costs: no frame rendered yet");
    if (!h->d_steps) return fail(h, VXRT_ERR_STATE, "vxrt_row_costs: the handle needs VXRT_FLAG_HIT_BUFFERS");
    CU(cudaSetDevice(h->cfg.device));
    const int H = h->cfg.height, W = h->cfg.width;
    float* d = nullptr;
    CU(cudaMalloc((void**)&d, (size_t)H * 3 * sizeof(float)));
    std::vector<float> v((size_t)H * 3);
    cudaError_t e = vxrt_launch_row_cost(h->d_steps, h->d_rays, h->d_normal, W, H, d, h->stream);
    if (e == cudaSuccess) e = cudaMemcpyAsync(v.data(), d, v.size() * sizeof(float), cudaMemcpyDeviceToHost, h->stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(h->stream);
    cudaFree(d);
    CU(e);
    for (int y = 0; y < H; y++) {
        const bool own = y >= h->row0 && y < h->row1;                 // rows this handle did not render cost nothing here
        trace_cost[y] = own ? v[3 * y] + 7.0f * v[3 * y + 1] + 12.0f * W + 4.7f * W : 0.0f;
        denoise_cost[y] = own ? 15.0f * v[3 * y + 2] + 0.45f * W : 0.0f;
    }
    return VXRT_OK;
}

int vxrt_frame_stats(VxrtHandle* h, VxrtFrameStats* out) {
    if (!h) return VXRT_ERR_INVALID;
    if (!out) return fail(h, VXRT_ERR_INVALID, "vxrt_frame_stats: out is NULL");
    if (!h->rendered) return fail(h, VXRT_ERR_STATE, "vxrt_frame_stats: no frame rendered yet");
    CU(cudaSetDevice(h->cfg.device));
    unsigned long long c[8];
    CU(cudaMemcpyAsync(c, h->d_counters, sizeof c, cudaMemcpyDeviceToHost, h->stream));
    CU(cudaStreamSynchronize(h->stream));
    { const int rc = peer_check(h); if (rc) return rc; }
    if (c[6]) return fail(h, VXRT_ERR_STATE, "k_trace_pool: a ray ring overflowed (trace_pool.cu); the frame is void");
    out->rays = c[0]; out->steps = c[1]; out->hits = c[2]; out->noise_fetches = c[3]; out->limit_exhausted = c[4];
    if (c[0]) h->steps_per_ray = (float)((double)c[1] / (double)c[0]);                  // what AUTO picks the trace form by (vxrt_render_frame)
    CU(cudaEventElapsedTime(&out->trace_ms, h->ev[0], h->ev[1]));
    CU(cudaEventElapsedTime(&out->denoise_ms, h->ev[1], h->ev[2]));
    CU(cudaEventElapsedTime(&out->finalize_ms, h->ev[2], h->ev[3]));
    CU(cudaEventElapsedTime(&out->total_ms, h->ev[0], h->ev[3]));
    return VXRT_OK;
}

}  // extern "C"
