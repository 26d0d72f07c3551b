// world.cu — device-resident voxel window: occupancy-LOD ("minefield") build, voxel edits and the
// seeded synthetic world generators.
//   k_build_minefield  replaces UnpackedChunkData::pack_into (src/world/chunk.rs:125-184)
//   k_apply_edits      no reference equivalent beyond set_block (chunk.rs:115-117)
//   k_generate_*       replaces src/world/generate.rs for SYNTHETIC, SEEDED inputs (the reference
//                      generator draws from thread_rng and is not reproducible)
#include "vxrt_internal.h"
#include <mutex>
#include <cooperative_groups.h>

namespace {

// ---------------------------------------------------------------------------------------------
// pack_into: per 64^3 chunk, minefield = 0 for solid voxels (bit 15), else the smallest k in 1..6
// whose 2^k-aligned cube contains a solid voxel; a wholly empty chunk becomes all 6 / words 0.
// One 512-thread block per chunk.  The words are read ONCE, 16 bytes per thread and load; the solid bits of a voxel
// row (64 voxels along x) are one 64-bit mask in shared memory, and level k is kept as masks at (y,z) resolution 2^k
// with x left expanded (bit x = "the 2^k-cube containing x holds a solid"), so a level is three ORs of the level below
// plus one in-word spread.  The levels are nested (solid in L1 in L2 ...), hence
// minefield(x) = 6 - #{S, L1..L5 with bit x set}: four voxels are a byte-wise sum of six spread nibbles and one
// 4-byte store.  HBM traffic: 4 B read + 1 B written per voxel (+4 B written when an empty chunk has to be wiped).
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ unsigned long long spread_or(unsigned long long m, int shift, unsigned long long hi_mask) {
    return m | ((m & hi_mask) >> shift) | ((m & ~hi_mask) << shift);      // OR of the two halves of every aligned 2*shift-bit group
}
__device__ __forceinline__ uint32_t nibble_to_bytes(uint32_t n) { return (n * 0x00204081u) & 0x01010101u; }   // bit i -> byte i

__global__ void __launch_bounds__(512) k_build_minefield(uint32_t* __restrict__ materials, uint8_t* __restrict__ minefield,
                                                         int nx, int ny, int nz, const uint32_t* __restrict__ dirty) {
    if (dirty && !dirty[blockIdx.x]) return;
    __shared__ unsigned long long S[64 * 64];                // solid bits of voxel row (z, y)
    __shared__ unsigned long long L1[32 * 32], L2[16 * 16], L3[8 * 8], L4[4 * 4], L5[2 * 2];
    __shared__ uint32_t s_any_solid, s_any_word;
    const int cxn = nx >> 6, cyn = ny >> 6;
    const int cx = blockIdx.x % cxn, cy = (blockIdx.x / cxn) % cyn, cz = blockIdx.x / (cxn * cyn);
    const size_t base = ((size_t)(cz * 64) * ny + cy * 64) * nx + cx * 64;
    const int tid = threadIdx.x;
    uint32_t* S32 = reinterpret_cast<uint32_t*>(S);
    for (int i = tid; i < 64 * 64 * 2; i += 512) S32[i] = 0;
    if (tid == 0) { s_any_solid = 0; s_any_word = 0; }
    __syncthreads();
    // pass 1: four voxels per load; q = z*1024 + y*16 + x4
    uint32_t any_word = 0, any_solid = 0;
#pragma unroll 8
    for (int it = 0; it < 128; it++) {
        const int q = tid + it * 512, x4 = q & 15, y = (q >> 4) & 63, z = q >> 10;
        const uint4 w = __ldcs(reinterpret_cast<const uint4*>(materials + base + ((size_t)z * ny + y) * nx + x4 * 4));
        const uint32_t nib = ((w.x >> 15) & 1u) | ((w.y >> 14) & 2u) | ((w.z >> 13) & 4u) | ((w.w >> 12) & 8u);
        any_word |= w.x | w.y | w.z | w.w;
        any_solid |= nib;
        if (nib) atomicOr(&S32[(z * 64 + y) * 2 + (x4 >> 3)], nib << ((x4 & 7) * 4));
    }
    if (__any_sync(0xffffffffu, any_solid != 0) && (tid & 31) == 0) s_any_solid = 1;
    if (__any_sync(0xffffffffu, any_word != 0) && (tid & 31) == 0) s_any_word = 1;
    __syncthreads();
    const bool empty = s_any_solid == 0;                                        // chunk.rs:155-161
    if (empty) {
        // wholly empty chunk: every voxel 6, every word 0 (written only if some word is not already 0)
        const bool wipe = s_any_word != 0;
        for (int it = 0; it < 128; it++) {
            const int q = tid + it * 512, x4 = q & 15, y = (q >> 4) & 63, z = q >> 10;
            const size_t gi = base + ((size_t)z * ny + y) * nx + x4 * 4;
            *reinterpret_cast<uint32_t*>(minefield + gi) = 0x06060606u;
            if (wipe) *reinterpret_cast<uint4*>(materials + gi) = make_uint4(0, 0, 0, 0);
        }
        return;
    }
    // levels: OR of the 2x2 rows below, then the in-word spread of that level's group size
    for (int c = tid; c < 32 * 32; c += 512) {
        const int y = c & 31, z = c >> 5;
        const unsigned long long m = S[(2 * z) * 64 + 2 * y] | S[(2 * z) * 64 + 2 * y + 1] | S[(2 * z + 1) * 64 + 2 * y] | S[(2 * z + 1) * 64 + 2 * y + 1];
        L1[c] = spread_or(m, 1, 0xAAAAAAAAAAAAAAAAull);
    }
    __syncthreads();
    if (tid < 16 * 16) {
        const int y = tid & 15, z = tid >> 4;
        const unsigned long long m = L1[(2 * z) * 32 + 2 * y] | L1[(2 * z) * 32 + 2 * y + 1] | L1[(2 * z + 1) * 32 + 2 * y] | L1[(2 * z + 1) * 32 + 2 * y + 1];
        L2[tid] = spread_or(m, 2, 0xCCCCCCCCCCCCCCCCull);
    }
    __syncthreads();
    if (tid < 8 * 8) {
        const int y = tid & 7, z = tid >> 3;
        const unsigned long long m = L2[(2 * z) * 16 + 2 * y] | L2[(2 * z) * 16 + 2 * y + 1] | L2[(2 * z + 1) * 16 + 2 * y] | L2[(2 * z + 1) * 16 + 2 * y + 1];
        L3[tid] = spread_or(m, 4, 0xF0F0F0F0F0F0F0F0ull);
    }
    __syncthreads();
    if (tid < 4 * 4) {
        const int y = tid & 3, z = tid >> 2;
        const unsigned long long m = L3[(2 * z) * 8 + 2 * y] | L3[(2 * z) * 8 + 2 * y + 1] | L3[(2 * z + 1) * 8 + 2 * y] | L3[(2 * z + 1) * 8 + 2 * y + 1];
        L4[tid] = spread_or(m, 8, 0xFF00FF00FF00FF00ull);
    }
    __syncthreads();
    if (tid < 2 * 2) {
        const int y = tid & 1, z = tid >> 1;
        const unsigned long long m = L4[(2 * z) * 4 + 2 * y] | L4[(2 * z) * 4 + 2 * y + 1] | L4[(2 * z + 1) * 4 + 2 * y] | L4[(2 * z + 1) * 4 + 2 * y + 1];
        L5[tid] = spread_or(m, 16, 0xFFFF0000FFFF0000ull);
    }
    __syncthreads();
    // pass 2: minefield(x) = 6 - #{S, L1..L5 with bit x set}
#pragma unroll 4
    for (int it = 0; it < 128; it++) {
        const int q = tid + it * 512, x4 = q & 15, y = (q >> 4) & 63, z = q >> 10;
        const int sh = x4 * 4;
        const uint32_t cnt = nibble_to_bytes((uint32_t)(S[z * 64 + y] >> sh) & 15u) + nibble_to_bytes((uint32_t)(L1[(z >> 1) * 32 + (y >> 1)] >> sh) & 15u) +
                             nibble_to_bytes((uint32_t)(L2[(z >> 2) * 16 + (y >> 2)] >> sh) & 15u) + nibble_to_bytes((uint32_t)(L3[(z >> 3) * 8 + (y >> 3)] >> sh) & 15u) +
                             nibble_to_bytes((uint32_t)(L4[(z >> 4) * 4 + (y >> 4)] >> sh) & 15u) + nibble_to_bytes((uint32_t)(L5[(z >> 5) * 2 + (y >> 5)] >> sh) & 15u);
        __stcs(reinterpret_cast<uint32_t*>(minefield + base + ((size_t)z * ny + y) * nx + x4 * 4), 0x06060606u - cnt);
    }
}

// ---------------------------------------------------------------------------------------------
// k_build_minefield_half: the same result from half the instructions and twice the CTAs.  ncu on the kernel above (C5, 512 chunks,
// profiles/r2_z): 104 M warp instructions at 53 % issue, 0.86 waves of 512-thread CTAs with every CTA in the same phase (read,
// levels, write), DRAM at 34 % of peak.  Three changes:
//   * a chunk is two 32-slab halves in z, one 256-thread CTA each: the levels up to L5 (32^3 cubes) never look across the halves,
//     so 1024 CTAs of mixed phases keep the memory system busy.  The one chunk-wide fact - pack_into wipes the words of a wholly
//     empty chunk (chunk.rs:155-161) - is exchanged between the two CTAs of a thread-block cluster through distributed shared
//     memory; the occupancy value is 6 whether its 32-cube or its whole chunk is empty, so it needs no exchange;
//   * the levels are nested (S in L1 in ... in L5), so the value - the index of the first level whose bit is set, 6 if none - has
//     the bit planes  b0 = L1&~S | L3&~L2 | L5&~L4,  b1 = L3&~L1 | ~L5,  b2 = ~L3  as 64-bit row masks: a dozen logic operations per
//     ROW instead of six nibble spreads and five adds per four voxels;
//   * a thread stores 16 voxels (one 16-byte store) from one 16-bit field of each plane.
// ---------------------------------------------------------------------------------------------
#ifndef VXRT_BUILD_MINB
#define VXRT_BUILD_MINB 8                          // 32 registers, 2048 threads per SM: the 1024 CTAs of a 512-chunk window are one wave (measured 6 / 7 / 8: 136 / 113 / 110 us)
#endif
__device__ __forceinline__ uint32_t planes_to_bytes(uint32_t f0, uint32_t f1, uint32_t f2, int g) {   // voxels 4g .. 4g+3 of a 16-bit field triple
    return nibble_to_bytes((f0 >> (4 * g)) & 15u) | (nibble_to_bytes((f1 >> (4 * g)) & 15u) << 1) | (nibble_to_bytes((f2 >> (4 * g)) & 15u) << 2);
}
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(256, VXRT_BUILD_MINB) k_build_minefield_half(uint32_t* __restrict__ materials, uint8_t* __restrict__ minefield,
                                                                                       int nx, int ny, int nz, const uint32_t* __restrict__ dirty) {
    const int chunk = blockIdx.x >> 1, half = blockIdx.x & 1;                  // the cluster's two CTAs: z slabs [0,32) and [32,64) of one chunk
    if (dirty && !dirty[chunk]) return;                                        // (both CTAs of the cluster leave)
    __shared__ unsigned long long S[32 * 64];                                  // solid bits of voxel row (z, y), z in the half
    __shared__ unsigned long long L1[16 * 32], L2[8 * 16], L3[4 * 8], L4[2 * 4], L5[1 * 2];
    __shared__ uint32_t s_flags[2];                                            // this half: [0] any solid voxel, [1] any nonzero word
    cooperative_groups::cluster_group cluster = cooperative_groups::this_cluster();
    const int cxn = nx >> 6, cyn = ny >> 6;
    const int cx = chunk % cxn, cy = (chunk / cxn) % cyn, cz = chunk / (cxn * cyn);
    const size_t base = ((size_t)(cz * 64 + half * 32) * ny + cy * 64) * nx + cx * 64;
    const int tid = threadIdx.x;
    uint32_t* S32 = reinterpret_cast<uint32_t*>(S);
    for (int i = tid; i < 32 * 64 * 2; i += 256) S32[i] = 0;
    if (tid < 2) s_flags[tid] = 0;
    __syncthreads();
    // pass 1: four voxels per load; q = z*1024 + y*16 + x4
    uint32_t any_word = 0, any_solid = 0;
#pragma unroll 8
    for (int it = 0; it < 128; it++) {
        const int q = tid + it * 256, x4 = q & 15, y = (q >> 4) & 63, z = q >> 10;
        const uint4 w = __ldcs(reinterpret_cast<const uint4*>(materials + base + ((size_t)z * ny + y) * nx + x4 * 4));
        const uint32_t nib = ((w.x >> 15) & 1u) | ((w.y >> 14) & 2u) | ((w.z >> 13) & 4u) | ((w.w >> 12) & 8u);
        any_word |= w.x | w.y | w.z | w.w;
        any_solid |= nib;
        if (nib) atomicOr(&S32[(z * 64 + y) * 2 + (x4 >> 3)], nib << ((x4 & 7) * 4));
    }
    if (__any_sync(0xffffffffu, any_solid != 0) && (tid & 31) == 0) s_flags[0] = 1;
    if (__any_sync(0xffffffffu, any_word != 0) && (tid & 31) == 0) s_flags[1] = 1;
    cluster.sync();                                                            // (also this CTA's barrier after pass 1)
    const uint32_t* peer = cluster.map_shared_rank(s_flags, cluster.block_rank() ^ 1u);
    const bool half_empty = s_flags[0] == 0, chunk_empty = half_empty && peer[0] == 0;          // chunk.rs:155-161
    const bool wipe = chunk_empty && s_flags[1] != 0;
    cluster.sync();                                                            // the peer's flags are not read after this point: either CTA may exit
    if (half_empty) {
        // no solid in this half: every voxel 6 (its 32-cubes are empty; so is the chunk, or the value is 6 anyway); words wiped
        // only for a wholly empty chunk that still holds nonzero words
        for (int it = 0; it < 32; it++) {
            const int q = tid + it * 256, x16 = q & 3, y = (q >> 2) & 63, z = q >> 8;
            const size_t gi = base + ((size_t)z * ny + y) * nx + x16 * 16;
            __stcs(reinterpret_cast<uint4*>(minefield + gi), make_uint4(0x06060606u, 0x06060606u, 0x06060606u, 0x06060606u));
            if (wipe) for (int k = 0; k < 4; k++) *reinterpret_cast<uint4*>(materials + gi + 4 * k) = make_uint4(0, 0, 0, 0);
        }
        return;
    }
    // levels: OR of the 2x2 rows below, then the in-word spread of that level's group size
    for (int c = tid; c < 16 * 32; c += 256) {
        const int y = c & 31, z = c >> 5;
        const unsigned long long m = S[(2 * z) * 64 + 2 * y] | S[(2 * z) * 64 + 2 * y + 1] | S[(2 * z + 1) * 64 + 2 * y] | S[(2 * z + 1) * 64 + 2 * y + 1];
        L1[c] = spread_or(m, 1, 0xAAAAAAAAAAAAAAAAull);
    }
    __syncthreads();
    if (tid < 8 * 16) {
        const int y = tid & 15, z = tid >> 4;
        const unsigned long long m = L1[(2 * z) * 32 + 2 * y] | L1[(2 * z) * 32 + 2 * y + 1] | L1[(2 * z + 1) * 32 + 2 * y] | L1[(2 * z + 1) * 32 + 2 * y + 1];
        L2[tid] = spread_or(m, 2, 0xCCCCCCCCCCCCCCCCull);
    }
    __syncthreads();
    if (tid < 4 * 8) {
        const int y = tid & 7, z = tid >> 3;
        const unsigned long long m = L2[(2 * z) * 16 + 2 * y] | L2[(2 * z) * 16 + 2 * y + 1] | L2[(2 * z + 1) * 16 + 2 * y] | L2[(2 * z + 1) * 16 + 2 * y + 1];
        L3[tid] = spread_or(m, 4, 0xF0F0F0F0F0F0F0F0ull);
    }
    __syncthreads();
    if (tid < 2 * 4) {
        const int y = tid & 3, z = tid >> 2;
        const unsigned long long m = L3[(2 * z) * 8 + 2 * y] | L3[(2 * z) * 8 + 2 * y + 1] | L3[(2 * z + 1) * 8 + 2 * y] | L3[(2 * z + 1) * 8 + 2 * y + 1];
        L4[tid] = spread_or(m, 8, 0xFF00FF00FF00FF00ull);
    }
    __syncthreads();
    if (tid < 2) {
        const int y = tid;                                                      // z = 0: the half is one 32-cube deep
        const unsigned long long m = L4[2 * y] | L4[2 * y + 1] | L4[4 + 2 * y] | L4[4 + 2 * y + 1];
        L5[tid] = spread_or(m, 16, 0xFFFF0000FFFF0000ull);
    }
    __syncthreads();
    // pass 2: 16 voxels per thread; value = index of the first set level, as three bit planes of the row
#pragma unroll 2
    for (int it = 0; it < 32; it++) {
        const int q = tid + it * 256, x16 = q & 3, y = (q >> 2) & 63, z = q >> 8;
        const unsigned long long s0 = S[z * 64 + y], l1 = L1[(z >> 1) * 32 + (y >> 1)], l2 = L2[(z >> 2) * 16 + (y >> 2)], l3 = L3[(z >> 3) * 8 + (y >> 3)],
                                 l4 = L4[(z >> 4) * 4 + (y >> 4)], l5 = L5[y >> 5];
        const unsigned long long b0 = (l1 & ~s0) | (l3 & ~l2) | (l5 & ~l4), b1 = (l3 & ~l1) | ~l5, b2 = ~l3;
        const int sh = x16 * 16;
        const uint32_t f0 = (uint32_t)(b0 >> sh) & 0xFFFFu, f1 = (uint32_t)(b1 >> sh) & 0xFFFFu, f2 = (uint32_t)(b2 >> sh) & 0xFFFFu;
        const uint4 v = make_uint4(planes_to_bytes(f0, f1, f2, 0), planes_to_bytes(f0, f1, f2, 1), planes_to_bytes(f0, f1, f2, 2), planes_to_bytes(f0, f1, f2, 3));
        __stcs(reinterpret_cast<uint4*>(minefield + base + ((size_t)z * ny + y) * nx + x16 * 16), v);
    }
}

// Stored-chunk validation (chunk_storage.rs:56-68 reads materials || minefield from a file): the 64^3 occupancy bytes a file
// carries against the ones k_build_minefield derives from the window's words.  One thread per 4 voxels; mismatching bytes are
// counted, the first mismatching texel offset (in the chunk) is kept.
__global__ void __launch_bounds__(256) k_compare_chunk(const uint8_t* __restrict__ minefield, const uint8_t* __restrict__ stored,
                                                       int nx, int ny, int cx, int cy, int cz, unsigned long long* __restrict__ out) {
    const int q = blockIdx.x * 256 + threadIdx.x;              // 65536 threads: q = z*1024 + y*16 + x4
    const int x4 = q & 15, y = (q >> 4) & 63, z = q >> 10;
    const size_t gi = ((size_t)(cz * 64 + z) * ny + cy * 64 + y) * nx + cx * 64 + x4 * 4;
    const uint32_t a = *reinterpret_cast<const uint32_t*>(minefield + gi), b = __ldg(reinterpret_cast<const uint32_t*>(stored) + q);
    uint32_t bad = 0, first = 4;
#pragma unroll
    for (int k = 3; k >= 0; k--) if (((a >> (8 * k)) & 0xFFu) != ((b >> (8 * k)) & 0xFFu)) { bad++; first = k; }
    if (bad) { atomicAdd(out, (unsigned long long)bad); atomicMin(out + 1, (unsigned long long)(q * 4 + first)); }
}

// Voxel edits.  world coordinate -> window texel: (w + N/2) mod N (raytrace.comp:105-106).  A list may name a voxel more than
// once; the reference's set_block (chunk.rs:115-117) applied in list order makes the LAST entry win, and every replica of a
// multi-GPU run must resolve it the same way.  Two passes over an open-addressing table keyed by the texel offset: pass 1
// leaves the highest list index of each voxel in the table (atomicMax), pass 2 lets exactly that entry write.
__device__ __forceinline__ unsigned long long edit_key(const VxrtEdit& e, int nx, int ny, int nz) {
    const uint32_t x = (uint32_t)(e.x + nx / 2) & (uint32_t)(nx - 1), y = (uint32_t)(e.y + ny / 2) & (uint32_t)(ny - 1),
                   z = (uint32_t)(e.z + nz / 2) & (uint32_t)(nz - 1);
    return ((unsigned long long)z * ny + y) * nx + x;
}
__device__ __forceinline__ uint32_t edit_slot(unsigned long long key, uint32_t mask) {
    key ^= key >> 29; key *= 0xBF58476D1CE4E5B9ull; key ^= key >> 32;
    return (uint32_t)key & mask;
}
__global__ void k_claim_edits(const VxrtEdit* __restrict__ edits, size_t n, int nx, int ny, int nz,
                              unsigned long long* __restrict__ keys, uint32_t* __restrict__ winner, uint32_t mask) {
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const unsigned long long key = edit_key(edits[i], nx, ny, nz) + 1ull;       // 0 = empty slot
    for (uint32_t h = edit_slot(key, mask);; h = (h + 1u) & mask) {
        const unsigned long long prev = atomicCAS(keys + h, 0ull, key);
        if (prev == 0ull || prev == key) { atomicMax(winner + h, (uint32_t)i + 1u); return; }
    }
}
__global__ void k_apply_edits(uint32_t* __restrict__ materials, uint32_t* __restrict__ dirty, const VxrtEdit* __restrict__ edits,
                              size_t n, int nx, int ny, int nz, const unsigned long long* __restrict__ keys,
                              const uint32_t* __restrict__ winner, uint32_t mask) {
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const VxrtEdit e = edits[i];
    const unsigned long long off = edit_key(e, nx, ny, nz), key = off + 1ull;
    uint32_t h = edit_slot(key, mask);
    while (keys[h] != key) h = (h + 1u) & mask;                               // present: pass 1 put it there
    if (winner[h] != (uint32_t)i + 1u) return;                                // a later entry of the list names this voxel
    materials[off] = e.word;
    const uint32_t x = (uint32_t)(off % (unsigned long long)nx), y = (uint32_t)((off / (unsigned long long)nx) % (unsigned long long)ny),
                   z = (uint32_t)(off / ((unsigned long long)nx * ny));
    dirty[((z >> 6) * (ny >> 6) + (y >> 6)) * (nx >> 6) + (x >> 6)] = 1u;
}

// ---------------------------------------------------------------------------------------------
// Seeded synthetic worlds (integer-only so that raytrace_b200/worldgen.py reproduces them bit for
// bit on the CPU).  Spec: DESIGN.md §5.
// ---------------------------------------------------------------------------------------------
__host__ __device__ inline uint32_t mix32(uint32_t a) {
    a ^= a >> 16; a *= 0x7feb352du; a ^= a >> 15; a *= 0x846ca68bu; a ^= a >> 16; return a;
}
__host__ __device__ inline uint32_t hash3(uint32_t x, uint32_t y, uint32_t z, uint32_t s) {
    return mix32(x * 0x9E3779B1u ^ mix32(y * 0x85EBCA77u ^ mix32(z * 0xC2B2AE3Du ^ s)));
}
__device__ inline uint32_t vnoise(uint32_t x, uint32_t y, uint32_t o, uint32_t S, uint32_t seed) {
    const uint32_t i = x / S, fx = x % S, j = y / S, fy = y % S;
    const uint64_t S3 = (uint64_t)S * S * S;
    const uint64_t wx = (uint64_t)fx * fx * (3 * S - 2 * fx), wy = (uint64_t)fy * fy * (3 * S - 2 * fy);
    const uint64_t v00 = hash3(i, j, o, seed) >> 16, v10 = hash3(i + 1, j, o, seed) >> 16,
                   v01 = hash3(i, j + 1, o, seed) >> 16, v11 = hash3(i + 1, j + 1, o, seed) >> 16;
    const uint64_t a = (v00 * (S3 - wx) + v10 * wx) / S3, b = (v01 * (S3 - wx) + v11 * wx) / S3;
    return (uint32_t)((a * (S3 - wy) + b * wy) / S3);
}

struct GenParams { int nx, ny, nz, kind; uint32_t seed; uint32_t words[7]; int sx[64], sy[64], stype[64], samp[64]; };

__device__ inline void biome_of(const GenParams& G, int x, int y, int* type, int* amp8) {
    if (G.kind == 0) { *type = 0; *amp8 = 8; return; }
    long long best = -1; int bk = 0;
    for (int k = 0; k < 64; k++) {
        const long long dx = x - G.sx[k], dy = y - G.sy[k], d = dx * dx + dy * dy;
        if (best < 0 || d < best) { best = d; bk = k; }
    }
    *type = G.stype[bk]; *amp8 = G.samp[bk];
}
__device__ inline int column_height(const GenParams& G, int x, int y, int* type) {
    int amp8; biome_of(G, x, y, type, &amp8);
    const uint32_t S0 = (uint32_t)min(128, G.nx / 2);
    const uint32_t amp = (uint32_t)(G.nz * 3 / 8);
    uint32_t h = (uint32_t)(G.nz / 16);
    uint32_t sum = 0;
    for (uint32_t o = 0; o < 4; o++) sum += ((amp >> o) * vnoise((uint32_t)x, (uint32_t)y, o, S0 >> o, G.seed)) >> 16;
    h += (sum * (uint32_t)amp8) >> 3;
    return min((int)h, G.nz * 3 / 4);
}
__device__ inline int band_material(int z, uint32_t r, int type) {            // shape of generate.rs:31-51
    const int lo = (type == 1) ? 4 : (type == 2) ? 5 : 2;
    const int mid = (type >= 2) ? 4 : 5;
    const int hi = 6;
    if (z < 20) return lo;
    if (z < 80) return (int)(r % 60u) < z - 20 ? mid : lo;
    if (z < 160) return (int)(r % 80u) < z - 80 ? hi : mid;
    return hi;
}

__global__ void __launch_bounds__(256) k_generate(uint32_t* __restrict__ materials, const __grid_constant__ GenParams G) {
    const int x = blockIdx.x * 32 + (threadIdx.x & 31), y = blockIdx.y * 8 + (threadIdx.x >> 5);
    if (x >= G.nx || y >= G.ny) return;
    int type; const int ht = column_height(G, x, y, &type);
    // tree of this column's 16x16 cell
    const uint32_t cx = (uint32_t)x >> 4, cy = (uint32_t)y >> 4;
    const uint32_t th = hash3(cx, cy, 7u, G.seed);
    int trunk_lo = 0, trunk_hi = 0, can_lo = 0, can_hi = 0;
    if ((th & 3u) == 0u) {
        const int tx = (int)(cx * 16 + 3 + ((th >> 2) & 7u)), ty = (int)(cy * 16 + 3 + ((th >> 5) & 7u));
        const int tall = 5 + (int)((th >> 8) & 7u);
        int ttype; const int tb = column_height(G, tx, ty, &ttype);
        if (x == tx && y == ty) { trunk_lo = tb; trunk_hi = tb + tall; }
        if (abs(x - tx) <= 2 && abs(y - ty) <= 2) { can_lo = tb + tall; can_hi = tb + tall + 3; }
    }
    for (int z = 0; z < G.nz; z++) {
        uint32_t w = 0;
        if (z < ht) w = G.words[band_material(z, hash3((uint32_t)x, (uint32_t)y, (uint32_t)z, G.seed ^ 0xA5A5A5A5u), type)];
        else if (z >= trunk_lo && z < trunk_hi) w = G.words[3];
        else if (z >= can_lo && z < can_hi) w = G.words[2];
        materials[((size_t)z * G.ny + y) * G.nx + x] = w;
    }
}


// Layout A/B (VXRT_FIELD=brick): the occupancy field once more, as 8x8x8-voxel bricks of 512 contiguous bytes
// (bricks x fastest; inside a brick x fastest).  One thread moves one 8-byte x-row of a brick; a warp reads 32 consecutive rows
// of one (y, z) line - coalesced - and stores them 512 bytes apart.  Only the bricks that meet the written box are rewritten.
__global__ void __launch_bounds__(256) k_field_to_bricks(const uint8_t* __restrict__ lin, uint8_t* __restrict__ bricks, int nx, int ny,
                                                        int bx0, int y0, int z0, int nbx, int ey, int ez) {
    const size_t n = (size_t)nbx * ey * ez;
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
        const int bx = bx0 + (int)(i % nbx); const size_t r = i / nbx;
        const int y = y0 + (int)(r % ey), z = z0 + (int)(r / ey);
        const uint2 v = *reinterpret_cast<const uint2*>(lin + ((size_t)z * ny + y) * nx + (size_t)bx * 8);
        const size_t brick = ((size_t)(z >> 3) * (ny >> 3) + (y >> 3)) * (nx >> 3) + bx;
        *reinterpret_cast<uint2*>(bricks + brick * 512 + (z & 7) * 64 + (y & 7) * 8) = v;
    }
}


// Layout A/B (VXRT_FIELD=nib): the occupancy field at 4 bits per voxel (its values are 0..6), two x-neighbours per
// byte, even x in the low nibble: half the footprint - the 128 MiB field of the 512^3 window becomes 64 MiB and fits L2.  One
// thread packs 16 voxels (one 16-byte load, one 8-byte store); only the rows that meet the written box are rewritten.
__global__ void __launch_bounds__(256) k_field_to_nibbles(const uint8_t* __restrict__ lin, uint8_t* __restrict__ nib, int nx, int ny,
                                                         int gx0, int y0, int z0, int ngx, int ey, int ez) {
    const size_t n = (size_t)ngx * ey * ez;
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
        const int gx = gx0 + (int)(i % ngx); const size_t r = i / ngx;
        const int y = y0 + (int)(r % ey), z = z0 + (int)(r / ey);
        const size_t row = ((size_t)z * ny + y) * nx;
        const uint4 v = *reinterpret_cast<const uint4*>(lin + row + (size_t)gx * 16);
        auto pack = [](uint32_t w) { return (w & 0xFu) | ((w >> 4) & 0xF0u) | ((w >> 8) & 0xF00u) | ((w >> 12) & 0xF000u); };   // 4 bytes -> 16 bits
        uint2 o;
        o.x = pack(v.x) | (pack(v.y) << 16); o.y = pack(v.z) | (pack(v.w) << 16);
        *reinterpret_cast<uint2*>(nib + (row >> 1) + (size_t)gx * 8) = o;
    }
}

}  // namespace

// see peer.cu: every kernel is loaded when a handle is created (a lazy load cannot complete under a spinning kernel)
#define VXRT_PRELOAD(k) do { cudaFuncAttributes a_; cudaError_t e_ = cudaFuncGetAttributes(&a_, k); if (e_ != cudaSuccess) return e_; } while (0)
cudaError_t vxrt_preload_world() { VXRT_PRELOAD(k_build_minefield); VXRT_PRELOAD(k_build_minefield_half); VXRT_PRELOAD(k_claim_edits); VXRT_PRELOAD(k_apply_edits); VXRT_PRELOAD(k_generate); VXRT_PRELOAD(k_compare_chunk); VXRT_PRELOAD(k_field_to_bricks); VXRT_PRELOAD(k_field_to_nibbles); return cudaSuccess; }

cudaError_t vxrt_launch_compare_chunk(const uint8_t* minefield, const uint8_t* d_stored, int nx, int ny, int cx, int cy, int cz,
                                      unsigned long long* d_out, cudaStream_t s) {
    k_compare_chunk<<<256, 256, 0, s>>>(minefield, d_stored, nx, ny, cx, cy, cz, d_out);
    return cudaGetLastError();
}

cudaError_t vxrt_launch_build_minefield(uint32_t* materials, uint8_t* minefield, int nx, int ny, int nz,
                                        const uint32_t* dirty, cudaStream_t s) {
    const int chunks = (nx >> 6) * (ny >> 6) * (nz >> 6);
    // two 256-thread CTAs (one thread-block cluster) per chunk; VXRT_BUILD_WHOLE=1 runs the one-CTA-per-chunk kernel (A/B, and what
    // vxrt_apply_edits_static keeps using)
    static const bool whole = getenv("VXRT_BUILD_WHOLE") != nullptr;
    if (whole) k_build_minefield<<<chunks, 512, 0, s>>>(materials, minefield, nx, ny, nz, dirty);
    else k_build_minefield_half<<<2 * chunks, 256, 0, s>>>(materials, minefield, nx, ny, nz, dirty);
    return cudaGetLastError();
}

cudaError_t vxrt_launch_apply_edits(uint32_t* materials, uint32_t* dirty, const VxrtEdit* d_edits, size_t n,
                                    int nx, int ny, int nz, void* table, size_t table_slots, cudaStream_t s) {
    if (n == 0) return cudaSuccess;
    // table: table_slots (a power of two >= 4 n) keys of 8 bytes, then as many 4-byte winners
    unsigned long long* keys = (unsigned long long*)table;
    uint32_t* winner = (uint32_t*)(keys + table_slots);
    cudaError_t e = cudaMemsetAsync(table, 0, table_slots * 12, s);
    if (e != cudaSuccess) return e;
    const unsigned blocks = (unsigned)((n + 255) / 256);
    k_claim_edits<<<blocks, 256, 0, s>>>(d_edits, n, nx, ny, nz, keys, winner, (uint32_t)(table_slots - 1));
    k_apply_edits<<<blocks, 256, 0, s>>>(materials, dirty, d_edits, n, nx, ny, nz, keys, winner, (uint32_t)(table_slots - 1));
    return cudaGetLastError();
}

cudaError_t vxrt_launch_field_to_bricks(const uint8_t* minefield, uint8_t* bricks, int nx, int ny, int nz, const int* box, cudaStream_t s) {
    const int whole[6] = {0, 0, 0, nx, ny, nz};
    const int* b = box ? box : whole;
    const int bx0 = b[0] >> 3, bx1 = (b[0] + b[3] + 7) >> 3;              // whole 8-byte rows; y and z per row, so no rounding there
    const size_t n = (size_t)(bx1 - bx0) * b[4] * b[5];
    if (n == 0) return cudaSuccess;
    const size_t blocks = (n + 255) / 256;
    k_field_to_bricks<<<(unsigned)(blocks > 148 * 32 ? 148 * 32 : blocks), 256, 0, s>>>(minefield, bricks, nx, ny, bx0, b[1], b[2], bx1 - bx0, b[4], b[5]);
    return cudaGetLastError();
}

cudaError_t vxrt_launch_field_to_nibbles(const uint8_t* minefield, uint8_t* nibbles, int nx, int ny, int nz, const int* box, cudaStream_t s) {
    const int whole[6] = {0, 0, 0, nx, ny, nz};
    const int* b = box ? box : whole;
    const int gx0 = b[0] >> 4, gx1 = (b[0] + b[3] + 15) >> 4;             // whole groups of 16 voxels
    const size_t n = (size_t)(gx1 - gx0) * b[4] * b[5];
    if (n == 0) return cudaSuccess;
    const size_t blocks = (n + 255) / 256;
    k_field_to_nibbles<<<(unsigned)(blocks > 148 * 32 ? 148 * 32 : blocks), 256, 0, s>>>(minefield, nibbles, nx, ny, gx0, b[1], b[2], gx1 - gx0, b[4], b[5]);
    return cudaGetLastError();
}

cudaError_t vxrt_launch_generate(uint32_t* materials, int nx, int ny, int nz, int kind, uint32_t seed, cudaStream_t s) {
    GenParams G;
    G.nx = nx; G.ny = ny; G.nz = nz; G.kind = kind; G.seed = seed;
    // voxel words of misc/materials.csv through the CSV rule build.rs:50-79 + pack GEN_MATERIALS.rs:44-51
    const uint32_t words[7] = {0x0u, 0x1FC07Fu, 0x9F73Du, 0xCD319u, 0xCD9B3u, 0xF8D96u, 0x1BBA73u};
    for (int i = 0; i < 7; i++) G.words[i] = words[i];
    for (uint32_t k = 0; k < 64; k++) {
        G.sx[k] = (int)(hash3(k, 0u, 1u, seed) % (uint32_t)nx); G.sy[k] = (int)(hash3(k, 1u, 1u, seed) % (uint32_t)ny);
        G.stype[k] = (int)(hash3(k, 2u, 1u, seed) % 4u);        G.samp[k] = 4 + (int)(hash3(k, 3u, 1u, seed) % 8u);
    }
    dim3 grid((nx + 31) / 32, (ny + 7) / 8);
    k_generate<<<grid, 256, 0, s>>>(materials, G);
    return cudaGetLastError();
}

// ---------------------------------------------------------------------------------------------
// vxrt_apply_edits_static: the edit path (k_claim_edits + k_apply_edits + k_build_minefield of the touched chunks) for callers
// that may not allocate - code under CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc.  No counterpart
// in the reference.  The scratch vxrt_apply_edits keeps per handle in cudaMalloc'd
// memory (the last-wins table and the per-chunk dirty flags) lives here in ONE file-scope device array of fixed size, which every
// GPU that loads the library holds until exit (896 KiB).  Every call shares it: it is cleared on the caller's stream at the start
// of each call, and two calls that overlap on the GPU race on it.
// ---------------------------------------------------------------------------------------------
constexpr size_t kStaticEditSlots = (size_t)4 * VXRT_STATIC_MAX_EDITS;            // table slots: 4 per edit (load factor <= 1/4)
constexpr size_t kStaticScratchBytes = kStaticEditSlots * 12 + (size_t)VXRT_STATIC_MAX_CHUNKS * 4;
static __device__ __align__(256) unsigned char g_static_edit_scratch[kStaticScratchBytes];   // keys (8 B) | winners (4 B) | dirty flags (4 B)

// The array's name is a host-side stand-in in host code, not a device address: the address is asked for once per device, with
// that device current, and cached.
static int static_scratch_address(int device, unsigned char** out, std::string* err) {
    static std::mutex mu;
    static unsigned char* cached[64] = {nullptr};
    if (device < 0 || device >= 64) { *err = "vxrt_apply_edits_static: device ordinal out of range"; return VXRT_ERR_INVALID; }
    std::lock_guard<std::mutex> lock(mu);
    if (!cached[device]) {
        cudaError_t e = cudaSetDevice(device);
        void* ptr = nullptr;
        if (e == cudaSuccess) e = cudaGetSymbolAddress(&ptr, g_static_edit_scratch);
        if (e != cudaSuccess || !ptr) { *err = std::string("vxrt_apply_edits_static: cudaGetSymbolAddress: ") + cudaGetErrorString(e); return VXRT_ERR_CUDA; }
        cached[device] = (unsigned char*)ptr;
    }
    *out = cached[device];
    return VXRT_OK;
}

extern "C" int vxrt_apply_edits_static(int device, uint32_t* d_materials, uint8_t* d_minefield, int nx, int ny, int nz,
                                       const VxrtEdit* d_edits, size_t n, void* cuda_stream) {
    std::string err;
    auto fail = [&](int code, const std::string& msg) { vxrt_set_global_error(msg.c_str()); return code; };
    auto pow2 = [](int v) { return v >= 64 && v <= 4096 && (v & (v - 1)) == 0; };
    if (!d_materials || !d_minefield || (n && !d_edits)) return fail(VXRT_ERR_INVALID, "vxrt_apply_edits_static: NULL device pointer");
    if (!pow2(nx) || !pow2(ny) || !pow2(nz)) return fail(VXRT_ERR_INVALID, "vxrt_apply_edits_static: window dims must be powers of two in [64, 4096]");
    const size_t chunks = (size_t)(nx >> 6) * (ny >> 6) * (nz >> 6);
    if (n > (size_t)VXRT_STATIC_MAX_EDITS)
        return fail(VXRT_ERR_SIZE, "vxrt_apply_edits_static: more than VXRT_STATIC_MAX_EDITS (16384) edits in one call; split the list or use vxrt_apply_edits");
    if (chunks > (size_t)VXRT_STATIC_MAX_CHUNKS)
        return fail(VXRT_ERR_SIZE, "vxrt_apply_edits_static: window has more than VXRT_STATIC_MAX_CHUNKS (32768) chunks; use vxrt_apply_edits");
    if (n == 0) return VXRT_OK;
    unsigned char* scratch = nullptr;
    const int rc = static_scratch_address(device, &scratch, &err);
    if (rc != VXRT_OK) return fail(rc, err);
    cudaError_t e = cudaSetDevice(device);
    if (e != cudaSuccess) return fail(VXRT_ERR_CUDA, std::string("vxrt_apply_edits_static: cudaSetDevice: ") + cudaGetErrorString(e));
    cudaStream_t s = (cudaStream_t)cuda_stream;
    unsigned long long* keys = (unsigned long long*)scratch;
    uint32_t* winner = (uint32_t*)(scratch + kStaticEditSlots * 8);
    uint32_t* dirty = (uint32_t*)(scratch + kStaticEditSlots * 12);
    // the array keeps the last call's contents: table and flags start at zero
    e = cudaMemsetAsync(scratch, 0, kStaticEditSlots * 12 + chunks * 4, s);
    if (e != cudaSuccess) return fail(VXRT_ERR_CUDA, std::string("vxrt_apply_edits_static: cudaMemsetAsync: ") + cudaGetErrorString(e));
    const unsigned blocks = (unsigned)((n + 255) / 256);
    k_claim_edits<<<blocks, 256, 0, s>>>(d_edits, n, nx, ny, nz, keys, winner, (uint32_t)(kStaticEditSlots - 1));
    k_apply_edits<<<blocks, 256, 0, s>>>(d_materials, dirty, d_edits, n, nx, ny, nz, keys, winner, (uint32_t)(kStaticEditSlots - 1));
    k_build_minefield<<<(unsigned)chunks, 512, 0, s>>>(d_materials, d_minefield, nx, ny, nz, dirty);
    e = cudaGetLastError();
    if (e != cudaSuccess) return fail(VXRT_ERR_CUDA, std::string("vxrt_apply_edits_static: launch: ") + cudaGetErrorString(e));
    return VXRT_OK;
}
