// trace_pool.cu — k_trace_pool: the FULL-lighting trace pass (shaders/glsl/raytrace.comp:290-385) with rays decoupled from
// lanes.
//
// k_trace_fast keeps a warp's 32 pixels together through the five rays of the lighting tree (:320-350); a round lasts as long
// as its longest ray, and bounce rays are short on the median and long in the tail (p50 = 1-2 steps, p90 = 25-31), so the
// march loop of a full-lighting frame runs at 13.9 of 32 lanes.  Here a warp is a persistent worker with three
// small rings in shared memory and no tie between a lane and a pixel:
//
//   RQ  rays ready to march (origin, direction, 1/|direction|, first occupancy texel, pixel + kind)
//   HL  hits that spawn rays: a primary hit (-> G-buffer, sun1 + dif1) or a first-bounce hit (-> sun2 + dif2)
//   FL  pixels whose light can be summed: every ray of the tree resolved, or the primary ray missed
//
// Lanes march whatever ray they hold.  When enough of them have finished (SERVICE_IDLE), the finished lanes file their result
// (one atomic on the pixel's 16-byte record for shadow / second-bounce rays, an HL entry for a hit that spawns) and take the next
// rays from RQ; HL and FL are worked off 32 entries at a time, so set-up, shading and the sky colours run on full warps.  Tiles
// of 8 x 4 pixels come from a frame-wide counter, so a warp's lanes only run dry when the frame does.
// Every ray is the ray of the sequential form - same origin, direction, steps, same fp32 operations - and a pixel's light is
// summed in the order of :320-350; images, running sum and frame counters equal k_trace_fast's bit for bit
// (tests/test_gpu_parity.py::test_ray_pool_is_invisible).
#include "trace_fast_dev.cuh"

namespace {

constexpr int PW = 4;                         // warps per CTA (each works alone)
constexpr int RQ_CAP = 128, HL_CAP = 128, FL_CAP = 128;
#ifndef VXRT_POOL_IDLE
#define VXRT_POOL_IDLE 12                     // finished lanes that trigger a service
#endif
#ifndef VXRT_POOL_CHUNK
#define VXRT_POOL_CHUNK 1
#endif
#ifndef VXRT_POOL_MINB
#define VXRT_POOL_MINB 8                       // CTAs per SM the register budget is set for (64 registers)
#endif
#ifndef VXRT_POOL_KSTEPS
#define VXRT_POOL_KSTEPS 4                    // march steps between two looks at the lane count
#endif
constexpr uint32_t TILE_CHUNK = VXRT_POOL_CHUNK;          // 8 x 4-pixel tiles a warp draws at a time (four 16 x 8 blocks)
enum { K_PRIMARY = 0, K_SUN1 = 1, K_DIF1 = 2, K_SUN2 = 3, K_DIF2 = 4 };
// ray / entry tag: px (12 bits) | py (12) | kind (3) | face code (3) | FL: bit 30 = "the primary ray missed"
__device__ __forceinline__ uint32_t tag_of(int px, int py, uint32_t kind, uint32_t face) { return (uint32_t)px | ((uint32_t)py << 12) | (kind << 24) | (face << 27); }
#define TAG_PX(t) ((int)((t) & 0xFFFu))
#define TAG_PY(t) ((int)(((t) >> 12) & 0xFFFu))
#define TAG_KIND(t) (((t) >> 24) & 7u)
#define TAG_FACE(t) (((t) >> 27) & 7u)
#define TAG_MISS 0x40000000u
// pixel record (uint4 in P.pool_ctx).  x: rays still out (bits 0-2) | sun1 saw the sky (3) | dif1 (4) | sun2 (5) | dif2 (6) |
// face of the primary hit (8-10) | face of the first-bounce hit (11-13) | noise bytes of the first level (16-31); y: voxel word of
// the first-bounce hit; z: noise bytes of the second level.  For a primary miss: x = sun-disc flag, y z w = sky colour (fp32 bits).
#define CTX_AIRBIT(kind) (1u << (2u + (kind)))

struct WarpPool {
    float rq[10][RQ_CAP];                     // pos xyz, dir xyz, 1/|dir| xyz, tag | first texel << 30 is NOT packed: [9] = tag, cur in [9]'s twin below
    uint32_t rq_cur[RQ_CAP];
    uint32_t hl[5][HL_CAP];                   // tag, pos xyz (offset hit position), voxel word
    uint32_t fl[FL_CAP];
};

template <int LR0, int DIMS>
__global__ void __launch_bounds__(32 * PW, VXRT_POOL_MINB) k_trace_pool(const __grid_constant__ TraceParams P) {
    constexpr bool SD = DIMS != 0;
    constexpr int LX = DIMS & 0xFF, LY = (DIMS >> 8) & 0xFF, LZ = (DIMS >> 16) & 0xFF;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    // the rings live in global memory, read and written past L1 (ld/st.cg): as shared memory they took 35 KB per CTA, which left the
    // occupancy field 7 % L1 hits and the SM 24 warps (r2_k); they are touched once per service, not per step
    WarpPool& W = reinterpret_cast<WarpPool*>(P.pool_rings)[blockIdx.x * PW + warp];
    const uint32_t FULL = 0xffffffffu, below = (1u << lane) - 1u;

    const float hx = SD ? (float)(1 << (LX - 1)) : P.hx, hy = SD ? (float)(1 << (LY - 1)) : P.hy, hz = SD ? (float)(1 << (LZ - 1)) : P.hz;
    const uint32_t mxm = SD ? (1u << LX) - 1u : (uint32_t)P.nx - 1u, mym = SD ? (1u << LY) - 1u : (uint32_t)P.ny - 1u,
                   mzm = SD ? (1u << LZ) - 1u : (uint32_t)P.nz - 1u, nxu = (uint32_t)P.nx, nyu = (uint32_t)P.ny;
    const uint8_t* const minefield = P.minefield;
    const uint32_t blo = (uint32_t)(unsigned long long)P.minefield, bhi = (uint32_t)((unsigned long long)P.minefield >> 32);   // (E15)
    const float M1 = 12582912.0f;             // 1.5 * 2^23
    const F3 sunlight = f3(P.sl_r, P.sl_g, P.sl_b);
    uint4* const ctx = reinterpret_cast<uint4*>(P.pool_ctx);

    auto sky_test = [&](float x, float y, float z) -> bool {                             // :138-145, (E8)
        if (LR0 == 2) return fmaxf(fmaxf(fabsf(x), fabsf(y)), fabsf(z)) >= hx;
        if (LR0) return (fabsf(x) >= hx) | (fabsf(y) >= hy) | (fabsf(z) >= hz);
        return (x >= P.sky_hi[0]) | (x <= P.sky_lo[0]) | (y >= P.sky_hi[1]) | (y <= P.sky_lo[1]) | (z >= P.sky_hi[2]) | (z <= P.sky_lo[2]);
    };
    auto texel_offset = [&](float qx, float qy, float qz) -> uint32_t {                  // (E2)+(E5): floor(p + H) & (N - 1), linear
        const uint32_t ix = (uint32_t)__float_as_int(__fadd_rd(qx, M1)) & mxm, iy = (uint32_t)__float_as_int(__fadd_rd(qy, M1)) & mym,
                       iz = (uint32_t)__float_as_int(__fadd_rd(qz, M1)) & mzm;
        return (iz * nyu + iy) * nxu + ix;
    };

    // ---- warp-uniform ring state -----------------------------------------------------------------------------------
    uint32_t rq_head = 0, rq_count = 0, hl_head = 0, hl_count = 0, fl_head = 0, fl_count = 0;
    bool tiles_left = true;
    uint32_t tile_next = 0, tile_end = 0;
    const int blocks_x = (P.width + 15) / 16, blocks_y = (P.row1 - P.row0 + 7) / 8;
    const uint32_t ntiles = (uint32_t)(blocks_x * blocks_y * 4);

    // ---- lane state: the ray this lane holds ------------------------------------------------------------------------
    // the ray this lane holds: tag == NO_RAY: none.  left > 0: marching, `left` iterations of :113's limit to go; left < 0: finished
    // and not yet filed, ~left iterations were to go when it stopped
    const uint32_t NO_RAY = 0xFFFFFFFFu;
    uint32_t tag = NO_RAY, cur = 0;
    int left = -1;
    f2 pxy = pk(0.f, 0.f), dxy = pk(0.f, 0.f), lpxy = pk(0.f, 0.f), qxy = pk(0.f, 0.f);
    float pz = 0.f, dz = 0.f, lpz = 0.f, qz = 0.f, lx = 0.f, ly = 0.f, lz = 0.f;
    uint32_t c_steps = 0, c_rays_hits = 0, c_noise_limit = 0;                           // frame counters of this lane
    const f2 one2 = P.one2;
    const f2 hxy = pk(hx, hy);

    // ---- ring pushes (whole warp calls; `on` lanes contribute) --------------------------------------------------------
    auto push_ray = [&](bool on, F3 o, F3 dir, F3 lpa, uint32_t c, uint32_t t) {
        const uint32_t m = __ballot_sync(FULL, on);
        if (on) {
            const uint32_t s = (rq_head + rq_count + (uint32_t)__popc(m & below)) & (RQ_CAP - 1);
            __stcg(&W.rq[0][s], o.x); __stcg(&W.rq[1][s], o.y); __stcg(&W.rq[2][s], o.z); __stcg(&W.rq[3][s], dir.x); __stcg(&W.rq[4][s], dir.y); __stcg(&W.rq[5][s], dir.z);
            __stcg(&W.rq[6][s], lpa.x); __stcg(&W.rq[7][s], lpa.y); __stcg(&W.rq[8][s], lpa.z); __stcg(&W.rq[9][s], __uint_as_float(t)); __stcg(&W.rq_cur[s], c);
        }
        rq_count += (uint32_t)__popc(m);
    };
    auto push_hit = [&](bool on, uint32_t t, F3 p, uint32_t word) {
        const uint32_t m = __ballot_sync(FULL, on);
        if (on) {
            const uint32_t s = (hl_head + hl_count + (uint32_t)__popc(m & below)) & (HL_CAP - 1);
            __stcg(&W.hl[0][s], t); __stcg(&W.hl[1][s], __float_as_uint(p.x)); __stcg(&W.hl[2][s], __float_as_uint(p.y)); __stcg(&W.hl[3][s], __float_as_uint(p.z)); __stcg(&W.hl[4][s], word);
        }
        hl_count += (uint32_t)__popc(m);
    };
    auto push_done = [&](bool on, uint32_t t) {
        const uint32_t m = __ballot_sync(FULL, on);
        if (on) __stcg(&W.fl[(fl_head + fl_count + (uint32_t)__popc(m & below)) & (FL_CAP - 1)], t);
        fl_count += (uint32_t)__popc(m);
    };
    // a ray that leaves the tree without spawning: one atomic on the pixel's record carries its sky bit and takes it off the
    // count of rays still out; the lane that takes the last one off files the pixel for its light
    auto resolve = [&](bool on, uint32_t t, bool air) {
        bool last = false;
        if (on) {
            const size_t pi = (size_t)TAG_PY(t) * P.width + TAG_PX(t);
            const uint32_t old = atomicAdd(&ctx[pi].x, (air ? CTX_AIRBIT(TAG_KIND(t)) : 0u) - 1u);
            last = (old & 7u) == 1u;
        }
        push_done(last, t & 0x00FFFFFFu);
    };

    // trace_ray's prologue (:83-107) for a ray (o, d): what the march needs, or - start inside a solid / NaN, (E4) - the hit of
    // word 0 after one iteration that IEEE makes of it
    struct Prep { F3 dir, lpa; uint32_t cur; bool instant; };
    auto prepare = [&](F3 o, F3 d) -> Prep {
        Prep r;
        r.dir = normalize3_shared(d);                                                   // :83
        r.lpa = rcp_abs3(r.dir);                                                        // :88
        const float qx0 = fadd(o.x, hx), qy0 = fadd(o.y, hy), qz0 = fadd(o.z, hz);
        const bool bad = !(qx0 == qx0) | !(qy0 == qy0) | !(qz0 == qz0) | !(r.dir.x == r.dir.x) | !(r.dir.y == r.dir.y) | !(r.dir.z == r.dir.z);
        r.cur = 0;
        if (!bad) r.cur = __ldg(minefield + texel_offset(qx0, qy0, qz0));               // :106
        r.instant = bad | (r.cur == 0);
        return r;
    };
    // the two rays a hit spawns (:325-330 / :337-342): towards the sun and the diffuse bounce, both from the offset hit position
    auto spawn_pair = [&](bool on, int px, int py, F3 hp, uint32_t nvx, uint32_t nvy, uint32_t face, uint32_t kind_sun, uint32_t kind_dif) {
        F3 ds = f3(0, 0, 1), dd = f3(0, 0, 1);
        Prep a, b;
        a.instant = b.instant = false; a.cur = b.cur = 1; a.dir = a.lpa = b.dir = b.lpa = f3(0, 0, 0);
        if (on) {
            ds = sun_ray_dir(P, make_uchar4((unsigned char)nvx, (unsigned char)nvy, 0, 0));
            dd = diffuse_direction_shared(P, nvx, nvy, face);
            a = prepare(hp, ds); b = prepare(hp, dd);
            c_rays_hits += 2u;
        }
        // (E4) rays: a "hit" of word 0 after one step - not the sky; a bounce ray that ends so spawns from a NaN position
        const bool ia = on && a.instant, ib = on && b.instant;
        if (ia | ib) { c_steps += (ia ? 1u : 0u) + (ib ? 1u : 0u); c_rays_hits += ((ia ? 1u : 0u) + (ib ? 1u : 0u)) << 16; }
        push_ray(on && !a.instant, hp, a.dir, a.lpa, a.cur, tag_of(px, py, kind_sun, 0));
        push_ray(on && !b.instant, hp, b.dir, b.lpa, b.cur, tag_of(px, py, kind_dif, 0));
        resolve(ia, tag_of(px, py, kind_sun, 0), false);
        if (kind_dif == K_DIF1) {
            const float qn = CUDART_NAN_F;
            if (ib) atomicAdd(&ctx[(size_t)py * P.width + px].x, 1u);                     // itself off, its two rays on
            push_hit(ib, tag_of(px, py, K_DIF1, b.dir.z > 0.0f ? 5u : 4u), f3(qn, qn, qn), 0u);
        } else {
            resolve(ib, tag_of(px, py, kind_dif, 0), false);
        }
    };

    // ---- a tile of 8 x 4 pixels: fog image, primary rays into RQ (:296-315, :380-385) --------------------------------------
    auto generate_tile = [&]() {
        // tiles are drawn TILE_CHUNK at a time from the frame-wide counter.  Measured (r2_l, C3 trace): 1 tile 0.80 ms, 2 tiles 0.86,
        // 4 tiles 0.94, 16 tiles 1.39 - the warps that draw the last, ground-heavy chunks finish alone, so the finest grain wins
        if (tile_next == tile_end) {
            uint32_t t0 = 0;
            if (lane == 0) t0 = (uint32_t)atomicAdd(P.counters + 5, (unsigned long long)TILE_CHUNK);
            t0 = __shfl_sync(FULL, t0, 0);
            if (t0 >= ntiles) { tiles_left = false; return; }
            tile_next = t0; tile_end = t0 + TILE_CHUNK < ntiles ? t0 + TILE_CHUNK : ntiles;
        }
        const uint32_t t = tile_next++;
        const uint32_t b = t >> 2, sub = t & 3u;
        const int bx = (int)(b % (uint32_t)blocks_x), by = (int)(b / (uint32_t)blocks_x);
        const int px = bx * 16 + (int)(sub & 1u) * 8 + (lane & 7), py = P.row0 + by * 8 + (int)(sub >> 1) * 4 + (lane >> 3);
        const bool on = px < P.width && py < P.row1;
        F3 o = f3(0, 0, 0), d = f3(0, 0, 1);
        Prep r;
        r.instant = false; r.cur = 1; r.dir = r.lpa = f3(0, 0, 0);
        if (on) {
            const size_t pi = (size_t)py * P.width + px;
            c_noise_limit += 2u;
            c_rays_hits += 1u;
            primary_ray(P, px, py, o, d);
            const Sky2 sk = sky_colours(P, d);
            P.fog[pi] = make_uchar4((unsigned char)unorm_store(fmul(sk.plain.x, 0.5f), 255.0f), (unsigned char)unorm_store(fmul(sk.plain.y, 0.5f), 255.0f),
                                    (unsigned char)unorm_store(fmul(sk.plain.z, 0.5f), 255.0f), 255);
            // the colour a miss will need (:322) waits in the pixel's record; a hit overwrites the record
            ctx[pi] = make_uint4(sk.sun ? 1u : 0u, __float_as_uint(sk.plain.x), __float_as_uint(sk.plain.y), __float_as_uint(sk.plain.z));
            r = prepare(o, d);
        }
        const bool inst = on && r.instant;
        if (inst) { c_steps += 1u; c_rays_hits += 1u << 16; }
        push_ray(on && !r.instant, o, r.dir, r.lpa, r.cur, tag_of(px, py, K_PRIMARY, 0));
        const float qn = CUDART_NAN_F;
        push_hit(inst, tag_of(px, py, K_PRIMARY, r.dir.z > 0.0f ? 5u : 4u), f3(qn, qn, qn), 0u);
    };

    // ---- 32 hits: G-buffer of a primary hit (:359-378), then the level's two rays ---------------------------------------------
    auto work_hits = [&]() {
        uint32_t n = hl_count < 32u ? hl_count : 32u;
        const uint32_t room = (RQ_CAP - rq_count) >> 1;
        if (n > room) n = room;
        if (n == 0) return;
        const bool on = (uint32_t)lane < n;
        const uint32_t s = (hl_head + (uint32_t)lane) & (HL_CAP - 1);
        uint32_t t = 0, word = 0; F3 hp = f3(0, 0, 0);
        if (on) { t = __ldcg(&W.hl[0][s]); hp = f3(__uint_as_float(__ldcg(&W.hl[1][s])), __uint_as_float(__ldcg(&W.hl[2][s])), __uint_as_float(__ldcg(&W.hl[3][s]))); word = __ldcg(&W.hl[4][s]); }
        __syncwarp();
        hl_head = (hl_head + n) & (HL_CAP - 1); hl_count -= n;
        const int px = TAG_PX(t), py = TAG_PY(t);
        const uint32_t face = TAG_FACE(t);
        const bool first = TAG_KIND(t) == K_PRIMARY;
        uint32_t nvx = 0, nvy = 0;
        if (on) {
            const size_t pi = (size_t)py * P.width + px;
            float nox, noy;
            noise_offset(P, px, py, nox, noy);
            c_noise_limit += 1u;
            if (first) {
                const uint32_t dq = __float2uint_rz(fmul(length3(sub3(f3(P.ox, P.oy, P.oz), hp)), 32.0f));   // :359-363 (NaN -> 0, as the sequential form)
                P.depth[pi] = (uint16_t)dq;
                P.normal[pi] = (uint8_t)face;
                const uint32_t f = ((word >> 14) & 0x7Fu) | ((word << 1) & 0x7F00u) | ((word << 16) & 0x7F0000u);
                reinterpret_cast<uint32_t*>(P.albedo)[pi] = ((f << 1) + ((f >> 6) & 0x010101u)) | 0xFF000000u;
                P.emission[pi] = make_uchar4(0, 0, 0, 255);
                const uchar4 nv = noise_texel(P.noise, gmod512(nox), gmod512(noy));                           // :324
                nvx = nv.x; nvy = nv.y;
                ctx[pi] = make_uint4(2u | (face << 8) | (nvx << 16) | (nvy << 24), 0u, 0u, 0u);
            } else {
                const float two_over = fdiv(2.0f, 512.0f);
                const uchar4 nv = noise_texel(P.noise, gmod512(fadd(nox, two_over)), gmod512(fadd(noy, two_over)));   // :336
                nvx = nv.x; nvy = nv.y;
                ctx[pi].y = word; ctx[pi].z = nvx | (nvy << 8);
                atomicOr(&ctx[pi].x, face << 11);
            }
        }
        __syncwarp();
        // both levels in one call each would double the code: the level only picks the kinds
        const uint32_t m_first = __ballot_sync(FULL, on && first), m_second = __ballot_sync(FULL, on && !first);
        if (m_first) spawn_pair(on && first, px, py, hp, nvx, nvy, face, K_SUN1, K_DIF1);
        if (m_second) spawn_pair(on && !first, px, py, hp, nvx, nvy, face, K_SUN2, K_DIF2);
    };

    // ---- 32 pixels whose light can be summed (:320-350), or whose primary ray missed (:322, :359-378) ---------------------------
    auto work_done = [&]() {
        const uint32_t n = fl_count < 32u ? fl_count : 32u;
        if (n == 0) return;
        const bool on = (uint32_t)lane < n;
        uint32_t t = 0;
        if (on) t = __ldcg(&W.fl[(fl_head + (uint32_t)lane) & (FL_CAP - 1)]);
        __syncwarp();
        fl_head = (fl_head + n) & (FL_CAP - 1); fl_count -= n;
        if (!on) return;
        const int px = TAG_PX(t), py = TAG_PY(t);
        const size_t pi = (size_t)py * P.width + px;
        const uint4 c = __ldcg(&ctx[pi]);                    // written by atomics at L2 and by other lanes: not through L1
        F3 light;
        if (t & TAG_MISS) {
            light = f3(__uint_as_float(c.y), __uint_as_float(c.z), __uint_as_float(c.w));
            if (c.x) light = f3(light.x + P.sl_r, light.y + P.sl_g, light.z + P.sl_b);
            P.depth[pi] = 0xFFFF; P.normal[pi] = 16;
            reinterpret_cast<uint32_t*>(P.albedo)[pi] = 0xFFFFFFFFu;
            P.emission[pi] = make_uchar4(0, 0, 0, 0);
        } else {
            const bool sun1 = (c.x & CTX_AIRBIT(K_SUN1)) != 0, d1 = (c.x & CTX_AIRBIT(K_DIF1)) != 0, sun2 = (c.x & CTX_AIRBIT(K_SUN2)) != 0,
                       d2 = (c.x & CTX_AIRBIT(K_DIF2)) != 0;
            const F3 l1 = sun1 ? sunlight : f3(0, 0, 0);                                                       // 0 + sunlight, :326-328
            if (d1) {
                light = add3(l1, sky_with_sun(P, diffuse_direction_shared(P, (c.x >> 16) & 0xFFu, c.x >> 24, (c.x >> 8) & 7u)));   // :331-332
            } else {
                F3 light2 = sun2 ? sunlight : f3(0, 0, 0);                                                     // :338-340
                if (d2) light2 = add3(light2, sky_with_sun(P, diffuse_direction_shared(P, c.z & 0xFFu, (c.z >> 8) & 0xFFu, (c.x >> 11) & 7u)));   // :343-345
                light2 = mul3(light2, decode_albedo(P, c.y));                                                  // :346
                light = add3(l1, light2);                                                                      // :348
            }
        }
        if (P.accum) {   // temporal accumulation (extension): running sum, the mean is stored
            float4 a = P.accum[pi];
            a.x = fadd(a.x, light.x); a.y = fadd(a.y, light.y); a.z = fadd(a.z, light.z); a.w = fadd(a.w, 1.0f);
            P.accum[pi] = a;
            light = f3(fdiv(a.x, a.w), fdiv(a.y, a.w), fdiv(a.z, a.w));
        }
        P.lighting[pi] = make_ushort4((unsigned short)unorm_store(fmul(light.x, 0.0625f), 65535.0f), (unsigned short)unorm_store(fmul(light.y, 0.0625f), 65535.0f),
                                      (unsigned short)unorm_store(fmul(light.z, 0.0625f), 65535.0f), (unsigned short)unorm_store(0.0625f, 65535.0f));
    };

    // ---- the worker loop ------------------------------------------------------------------------------------------------------------
    for (int guard = 0; guard < (1 << 28); guard++) {
        // (1) finished lanes file their result
        const bool fin = tag != NO_RAY && left < 0;
        if (__any_sync(FULL, fin)) {
            bool r_plain = false, r_air = false, r_spawn = false, r_miss = false;
            uint32_t word = 0, face = 0, tag_done = 0;
            F3 hp = f3(0, 0, 0);
            if (fin) {
                const float x = lo2(pxy), y = hi2(pxy);
                const uint32_t steps = (uint32_t)(P.max_steps - ~left);
                c_steps += steps;
                const bool sky = sky_test(x, y, pz);
                const bool hit = !sky && cur == 0;                                       // sky first, :138-146
                if (!sky && cur != 0) c_noise_limit += 1u << 16;                        // loop limit exhausted: reads as a miss
                if (hit) c_rays_hits += 1u << 16;
                const uint32_t kind = TAG_KIND(tag);
                r_air = !hit;
                if (hit && (kind == K_PRIMARY || kind == K_DIF1)) {
                    r_spawn = true;
                    word = __ldg(P.materials + texel_offset(lo2(qxy), hi2(qxy), qz));   // :146-159, (E3)
                    const bool dpx = lo2(dxy) > 0.0f, dpy = hi2(dxy) > 0.0f, dpz = dz > 0.0f;   // :89-93
                    if (lx < ly) face = (lx < lz) ? (dpx ? 1u : 0u) : (dpz ? 5u : 4u);   // (E7) :120-136 of the last step
                    else face = (ly < lz) ? (dpy ? 3u : 2u) : (dpz ? 5u : 4u);
                    const float off = (face & 1u) ? -0.001f : 0.001f;                    // :166-180
                    hp = f3(x, y, pz);
                    if ((face >> 1) == 0) hp.x = fadd(hp.x, off);
                    else if ((face >> 1) == 1) hp.y = fadd(hp.y, off);
                    else hp.z = fadd(hp.z, off);
                    if (kind == K_DIF1) atomicAdd(&ctx[(size_t)TAG_PY(tag) * P.width + TAG_PX(tag)].x, 1u);   // itself off, its two rays on
                } else if (kind == K_PRIMARY) r_miss = true;
                else r_plain = true;
                tag_done = tag; tag = NO_RAY;
            }
            push_hit(r_spawn, (tag_done & 0x07FFFFFFu) | (face << 27), hp, word);
            push_done(r_miss, (tag_done & 0x00FFFFFFu) | TAG_MISS);
            resolve(r_plain, tag_done, r_air);
            __syncwarp();
        }
        // (2) keep RQ stocked: deeper levels first (they free pixels), new tiles when nothing else is ready
        for (int it = 0; it < 8 && (rq_count < 32u || hl_count >= 64u); it++) {
            if (hl_count >= 32u || (hl_count > 0u && !tiles_left)) work_hits();
            else if (tiles_left && rq_count < 32u) generate_tile();
            else break;
            __syncwarp();
        }
        while (fl_count >= 32u) { work_done(); __syncwarp(); }
        if (fl_count > 0u && rq_count == 0u && hl_count == 0u && !tiles_left) { work_done(); __syncwarp(); }
        if ((rq_count > (uint32_t)RQ_CAP) | (hl_count > (uint32_t)HL_CAP) | (fl_count > (uint32_t)FL_CAP)) {   // a ring overflowed: results are void
            if (lane == 0) atomicAdd(P.counters + 6, 1ull);
            break;
        }
        // (3) idle lanes take the next rays
        {
            const uint32_t idle = __ballot_sync(FULL, tag == NO_RAY);
            const uint32_t rank = (uint32_t)__popc(idle & below);
            if (tag == NO_RAY && rank < rq_count) {
                const uint32_t s = (rq_head + rank) & (RQ_CAP - 1);
                pxy = pk(__ldcg(&W.rq[0][s]), __ldcg(&W.rq[1][s])); pz = __ldcg(&W.rq[2][s]);
                dxy = pk(__ldcg(&W.rq[3][s]), __ldcg(&W.rq[4][s])); dz = __ldcg(&W.rq[5][s]);
                lpxy = pk(__ldcg(&W.rq[6][s]), __ldcg(&W.rq[7][s])); lpz = __ldcg(&W.rq[8][s]);
                tag = __float_as_uint(__ldcg(&W.rq[9][s])); cur = __ldcg(&W.rq_cur[s]);
                qxy = add2(pxy, hxy); qz = fadd(pz, hz);
                left = P.max_steps;
            }
            const uint32_t take = (uint32_t)__popc(idle) < rq_count ? (uint32_t)__popc(idle) : rq_count;
            __syncwarp();
            rq_head = (rq_head + take) & (RQ_CAP - 1); rq_count -= take;
        }
        // (4) nothing marching and nothing queued: the frame's tiles are gone and every list is empty
        if (!__any_sync(FULL, tag != NO_RAY)) {
            if (rq_count == 0u && hl_count == 0u && fl_count == 0u && !tiles_left) break;
            continue;
        }
        // (5) march until SERVICE_IDLE lanes have finished (or everything has)
        float Ms = __int_as_float((int)(cur << 23) + 0x4AC00000);                        // 1.5 * 2^(22+cur)
        for (;;) {
#pragma unroll
            for (int k = 0; k < VXRT_POOL_KSTEPS; k++) {
                if (left > 0) {
                    // one iteration of trace_ray's loop (:113-162), the arithmetic of k_trace_fast's packed form
                    const float ax = flip_if_positive(lo2(qxy), lo2(dxy)), ay = flip_if_positive(hi2(qxy), hi2(dxy)), az = flip_if_positive(qz, dz);
                    const f2 axy = pk(ax, ay), Ms2 = pk(Ms, Ms);
                    const f2 mxy = sub2(axy, sub2(add2_rd(axy, Ms2), Ms2));               // (E5)
                    const float mz = fsub(az, fsub(__fadd_rd(az, Ms), Ms));
                    const f2 lxy = mul2(add2(mxy, pk(0.0001f, 0.0001f)), lpxy);           // :119
                    lz = fmul(fadd(0.0001f, mz), lpz);
                    lx = lo2(lxy); ly = hi2(lxy);
                    const float t = fminf(fminf(lx, ly), lz);                             // (E7)
                    const f2 sxy = mul2(dxy, pk(t, t));
                    pxy = fma2(sxy, one2, pxy);                                           // (E14)
                    pz = fadd(pz, fmul(dz, t));
                    qxy = add2(pxy, hxy); qz = fadd(pz, hz);
                    {
                        const f2 fxy = add2_rd(qxy, pk(M1, M1));                              // (E5)
                        const uint32_t ix = (uint32_t)__float_as_int(lo2(fxy)) & mxm, iy = (uint32_t)__float_as_int(hi2(fxy)) & mym,
                                       iz = (uint32_t)__float_as_int(__fadd_rd(qz, M1)) & mzm;
                        cur = ldg_u8_halves((iz * nyu + iy) * nxu + (ix | blo), bhi);         // :137, (E15)
                    }
                    left--;
                    if (sky_test(lo2(pxy), hi2(pxy), pz) | (cur == 0) | (left == 0)) left = ~left;   // :138-146, :113
                    else Ms = __int_as_float((int)(cur << 23) + 0x4AC00000);              // :161
                }
            }
            const uint32_t marching = __ballot_sync(FULL, left > 0);
            if (marching == 0u) break;
            // leave for a service when enough rays have finished: VXRT_POOL_IDLE of a full warp, half of what a thin one holds (their
            // results are what new rays come from; empty lanes alone cannot be helped)
            const uint32_t holding = (uint32_t)__popc(__ballot_sync(FULL, tag != NO_RAY)), nfin = holding - (uint32_t)__popc(marching);
            const uint32_t need = holding / 2u < (uint32_t)VXRT_POOL_IDLE ? (holding / 2u > 0u ? holding / 2u : 1u) : (uint32_t)VXRT_POOL_IDLE;
            if (nfin >= need) break;
        }
    }
    // ---- frame counters: rays, steps, hits, noise fetches, limit exhausted ---------------------------------------------------------------
    {
        const uint32_t v[5] = {c_rays_hits & 0xFFFFu, c_steps, c_rays_hits >> 16, c_noise_limit & 0xFFFFu, c_noise_limit >> 16};
#pragma unroll
        for (int k = 0; k < 5; k++) {
            const uint32_t s = __reduce_add_sync(FULL, v[k]);
            if (lane == 0 && s) atomicAdd(P.counters + k, (unsigned long long)s);
        }
    }
}

}  // namespace

#define VXRT_PRELOAD(k) do { cudaFuncAttributes a_; cudaError_t e_ = cudaFuncGetAttributes(&a_, k); if (e_ != cudaSuccess) return e_; } while (0)
#define VXRT_DIMS(lx, ly, lz) ((lx) | ((ly) << 8) | ((lz) << 16))
static const void* pool_kernel(int form, int dims) {
    if (dims == VXRT_DIMS(9, 9, 9)) return form == 2 ? (const void*)k_trace_pool<2, VXRT_DIMS(9, 9, 9)> : form == 1 ? (const void*)k_trace_pool<1, VXRT_DIMS(9, 9, 9)> : (const void*)k_trace_pool<0, VXRT_DIMS(9, 9, 9)>;
    if (dims == VXRT_DIMS(11, 11, 8)) return form == 1 ? (const void*)k_trace_pool<1, VXRT_DIMS(11, 11, 8)> : form == 0 ? (const void*)k_trace_pool<0, VXRT_DIMS(11, 11, 8)> : nullptr;
    return form == 2 ? (const void*)k_trace_pool<2, 0> : form == 1 ? (const void*)k_trace_pool<1, 0> : (const void*)k_trace_pool<0, 0>;
}
static const size_t kPoolSmem = 0;
size_t vxrt_trace_pool_ring_bytes(int sms) { return sizeof(WarpPool) * PW * (size_t)sms * VXRT_POOL_MINB; }

cudaError_t vxrt_preload_trace_pool() {
    for (int form = 0; form < 3; form++)
        for (int dims : {0, VXRT_DIMS(9, 9, 9), VXRT_DIMS(11, 11, 8)}) {
            const void* k = pool_kernel(form, dims);
            if (!k) continue;
            cudaFuncAttributes a;
            cudaError_t e = cudaFuncGetAttributes(&a, k);
            if (e != cudaSuccess) return e;
        }
    return cudaSuccess;
}

// the pool form of the trace pass for these parameters, or nullptr where k_trace_fast has to run it (see the conditions)
const void* vxrt_trace_pool_kernel(const TraceParams& p, dim3* grid, dim3* block, size_t* smem) {
    // opt-in: slower than the sequential form as built.  C3 trace 0.79 ms against 0.436 (r2_k, r2_l): lanes per instruction
    // 16.9 -> 23.2 and 369 M warp instructions against 318 M, but 30 % issue utilisation - a warp's 32 rays all wait while a service
    // (ring traffic, the pixel-record atomics, spawn set-up with its first occupancy fetches) runs its chain of L2 round trips,
    // every ~12 finished rays.  Rings in shared memory (7 % L1 hits, 24 warps) or in global memory (24-32 % hits, 30 warps),
    // 40 / 51 / 64 registers, service thresholds 12-24 and 4-8 steps between lane counts all land within 5 % of each other.
    static const bool on_env = getenv("VXRT_POOL") != nullptr;
    if (!(p.trace_form == VXRT_TRACE_POOL || (p.trace_form == VXRT_TRACE_AUTO && on_env)) || !p.pool_ctx || !p.pool_rings || p.mode != VXRT_MODE_FULL || p.hit_voxel || p.hit_pos || p.field) return nullptr;
    if (p.own_row0 > p.row0 || p.own_row1 < p.row1 || p.width > 4096 || p.height > 4096 || p.max_steps > 200000) return nullptr;
    const unsigned long long bytes = (unsigned long long)p.nx * p.ny * p.nz;
    const unsigned long long base = (unsigned long long)p.minefield;
    if ((base & (unsigned long long)(p.nx - 1)) != 0 || (base & 0xFFFFFFFFull) + bytes > 0x100000000ull) return nullptr;   // (E15)
    const bool lr0 = p.lrx == 0.0f && p.lry == 0.0f && p.lrz == 0.0f;
    const int form = !lr0 ? 0 : (p.nx == p.ny && p.ny == p.nz) ? 2 : 1;
    int dev = 0, sms = 148;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    const int tiles = ((p.width + 15) / 16) * ((p.row1 - p.row0 + 7) / 8) * 4;
    int ctas = sms * VXRT_POOL_MINB;
    if (ctas * PW > tiles) ctas = (tiles + PW - 1) / PW;
    *grid = dim3((unsigned)ctas); *block = dim3(32 * PW); *smem = kPoolSmem;
    return pool_kernel(form, VXRT_DIMS(p.lx, p.ly, p.lz));
}
