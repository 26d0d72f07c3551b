// chunk_store.cpp — the reference's chunk file format (src/world/chunk_storage.rs:37-68), host side only:
//   file name  = "{:016X}{:016X}{:016X}" of the signed chunk coordinates (two's complement, :38-41)
//   file bytes = LZ4 frame( materials u32 LE x 64^3  ||  minefield u8 x 64^3 )           (:43-56, level 4)
// The reference links the `lz4` crate 1.23.1 (liblz4, not in this tree); this file restates the published LZ4 frame
// format (magic 0x184D2204, FLG/BD/HC descriptor, blocks with a 31-bit size + "stored" bit, end mark, optional xxHash32
// block / content checksums) and the LZ4 block format (token, literals, 16-bit offset, match length) from their
// specifications.  The decoder accepts everything liblz4 emits (linked or independent blocks, any block size, content
// size, checksums — which it verifies); the encoder writes independent 4 MiB blocks with a greedy hash-chain matcher
// and a content checksum, which any LZ4 frame reader (the reference's Decoder included) accepts.
#include <stdint.h>
#include <stdio.h>
#include <string.h>
#include <vector>
#include "../../include/vxrt.h"

namespace {

constexpr size_t kChunkVoxels = 64u * 64u * 64u, kChunkBytes = kChunkVoxels * 5u;

uint32_t rd32(const uint8_t* p) { return (uint32_t)p[0] | ((uint32_t)p[1] << 8) | ((uint32_t)p[2] << 16) | ((uint32_t)p[3] << 24); }
void wr32(std::vector<uint8_t>& v, uint32_t x) { for (int i = 0; i < 4; i++) v.push_back((uint8_t)(x >> (8 * i))); }
uint32_t rotl(uint32_t x, int r) { return (x << r) | (x >> (32 - r)); }

// xxHash32 (specification: github.com/Cyan4973/xxHash/blob/dev/doc/xxhash_spec.md)
uint32_t xxh32(const uint8_t* p, size_t n, uint32_t seed) {
    const uint32_t P1 = 2654435761u, P2 = 2246822519u, P3 = 3266489917u, P4 = 668265263u, P5 = 374761393u;
    const uint8_t* end = p + n;
    uint32_t h;
    if (n >= 16) {
        uint32_t v1 = seed + P1 + P2, v2 = seed + P2, v3 = seed, v4 = seed - P1;
        for (; p + 16 <= end; p += 16) {
            v1 = rotl(v1 + rd32(p) * P2, 13) * P1; v2 = rotl(v2 + rd32(p + 4) * P2, 13) * P1;
            v3 = rotl(v3 + rd32(p + 8) * P2, 13) * P1; v4 = rotl(v4 + rd32(p + 12) * P2, 13) * P1;
        }
        h = rotl(v1, 1) + rotl(v2, 7) + rotl(v3, 12) + rotl(v4, 18);
    } else {
        h = seed + P5;
    }
    h += (uint32_t)n;
    for (; p + 4 <= end; p += 4) h = rotl(h + rd32(p) * P3, 17) * P4;
    for (; p < end; p++) h = rotl(h + *p * P5, 11) * P1;
    h ^= h >> 15; h *= P2; h ^= h >> 13; h *= P3; h ^= h >> 16;
    return h;
}

// One LZ4 block into out[pos..]; matches may reach back into earlier blocks (linked mode): `out` is the whole frame.
bool lz4_block_decode(const uint8_t* src, size_t n, std::vector<uint8_t>& out, size_t cap) {
    size_t i = 0;
    while (i < n) {
        const uint8_t token = src[i++];
        size_t lit = token >> 4;
        if (lit == 15) { uint8_t b; do { if (i >= n) return false; b = src[i++]; lit += b; } while (b == 255); }
        if (i + lit > n || out.size() + lit > cap) return false;
        out.insert(out.end(), src + i, src + i + lit);
        i += lit;
        if (i >= n) break;                                   // the last sequence has no match part
        if (i + 2 > n) return false;
        const size_t off = (size_t)src[i] | ((size_t)src[i + 1] << 8);
        i += 2;
        size_t len = token & 15;
        if (len == 15) { uint8_t b; do { if (i >= n) return false; b = src[i++]; len += b; } while (b == 255); }
        len += 4;
        if (off == 0 || off > out.size() || out.size() + len > cap) return false;
        size_t from = out.size() - off;
        for (size_t k = 0; k < len; k++) out.push_back(out[from + k]);   // byte-wise: overlapping matches repeat
    }
    return true;
}

// Greedy LZ4 block compressor: 4-byte hash table, matches extended forwards; the block format's end conditions
// (last 5 bytes are literals, no match starts within the last 12 bytes) are respected.
void lz4_block_encode(const uint8_t* src, size_t n, std::vector<uint8_t>& out) {
    std::vector<int32_t> table(1 << 16, -1);
    size_t anchor = 0, i = 0;
    auto emit = [&](size_t lit_end, size_t match_len, size_t offset) {
        const size_t lit = lit_end - anchor;
        const size_t ml = match_len ? match_len - 4 : 0;
        out.push_back((uint8_t)(((lit < 15 ? lit : 15) << 4) | (match_len ? (ml < 15 ? ml : 15) : 0)));
        if (lit >= 15) { size_t r = lit - 15; while (r >= 255) { out.push_back(255); r -= 255; } out.push_back((uint8_t)r); }
        out.insert(out.end(), src + anchor, src + lit_end);
        if (match_len) {
            out.push_back((uint8_t)(offset & 255)); out.push_back((uint8_t)(offset >> 8));
            if (ml >= 15) { size_t r = ml - 15; while (r >= 255) { out.push_back(255); r -= 255; } out.push_back((uint8_t)r); }
        }
    };
    if (n > 12) {
        const size_t match_limit = n - 5, start_limit = n - 12;
        while (i < start_limit) {
            const uint32_t h = (rd32(src + i) * 2654435761u) >> 16;
            const int32_t cand = table[h];
            table[h] = (int32_t)i;
            if (cand >= 0 && i - (size_t)cand <= 65535 && rd32(src + cand) == rd32(src + i)) {
                size_t len = 4;
                while (i + len < match_limit && src[cand + len] == src[i + len]) len++;
                emit(i, len, i - (size_t)cand);
                i += len;
                anchor = i;
            } else {
                i++;
            }
        }
    }
    emit(n, 0, 0);                                           // trailing literals
}

}  // namespace

extern "C" {

int vxrt_chunk_file_name(int64_t cx, int64_t cy, int64_t cz, char* out, size_t nbytes) {
    if (!out || nbytes < 49) return VXRT_ERR_SIZE;
    snprintf(out, nbytes, "%016llX%016llX%016llX", (unsigned long long)cx, (unsigned long long)cy, (unsigned long long)cz);
    return VXRT_OK;
}

size_t vxrt_chunk_encode_bound(void) { return kChunkBytes + kChunkBytes / 255 + 64; }

int vxrt_chunk_encode(const uint32_t* materials, const uint8_t* minefield, void* out, size_t cap, size_t* nbytes) {
    if (!materials || !minefield || !out || !nbytes) return VXRT_ERR_INVALID;
    std::vector<uint8_t> raw(kChunkBytes);
    for (size_t i = 0; i < kChunkVoxels; i++) { const uint32_t w = materials[i]; raw[4 * i] = (uint8_t)w; raw[4 * i + 1] = (uint8_t)(w >> 8); raw[4 * i + 2] = (uint8_t)(w >> 16); raw[4 * i + 3] = (uint8_t)(w >> 24); }
    memcpy(raw.data() + 4 * kChunkVoxels, minefield, kChunkVoxels);
    std::vector<uint8_t> f;
    wr32(f, 0x184D2204u);
    const uint8_t flg = 0x40 | 0x20 | 0x04;                 // version 01, independent blocks, content checksum
    const uint8_t bd = 0x70;                                 // 4 MiB maximum block size
    f.push_back(flg); f.push_back(bd);
    f.push_back((uint8_t)(xxh32(f.data() + 4, 2, 0) >> 8));  // header checksum
    std::vector<uint8_t> blk;
    lz4_block_encode(raw.data(), raw.size(), blk);
    if (blk.size() < raw.size()) { wr32(f, (uint32_t)blk.size()); f.insert(f.end(), blk.begin(), blk.end()); }
    else { wr32(f, (uint32_t)raw.size() | 0x80000000u); f.insert(f.end(), raw.begin(), raw.end()); }   // stored
    wr32(f, 0);                                              // end mark
    wr32(f, xxh32(raw.data(), raw.size(), 0));
    *nbytes = f.size();
    if (f.size() > cap) return VXRT_ERR_SIZE;
    memcpy(out, f.data(), f.size());
    return VXRT_OK;
}

int vxrt_chunk_decode(const void* file, size_t nbytes, uint32_t* materials, uint8_t* minefield) {
    if (!file || !materials || !minefield) return VXRT_ERR_INVALID;
    const uint8_t* p = (const uint8_t*)file;
    size_t i = 0;
    if (nbytes < 7 || rd32(p) != 0x184D2204u) return VXRT_ERR_INVALID;
    const uint8_t flg = p[4], bd = p[5];
    if ((flg >> 6) != 1 || (bd & 0x8F)) return VXRT_ERR_INVALID;
    const bool block_checksum = flg & 0x10, has_size = flg & 0x08, content_checksum = flg & 0x04, has_dict = flg & 0x01;
    i = 6 + (has_size ? 8 : 0) + (has_dict ? 4 : 0);
    if (i + 1 > nbytes || (uint8_t)(xxh32(p + 4, i - 4, 0) >> 8) != p[i]) return VXRT_ERR_INVALID;
    i++;
    std::vector<uint8_t> raw;
    raw.reserve(kChunkBytes);
    for (;;) {
        if (i + 4 > nbytes) return VXRT_ERR_INVALID;
        const uint32_t bs = rd32(p + i);
        i += 4;
        if (bs == 0) break;
        const size_t n = bs & 0x7FFFFFFFu;
        if (i + n + (block_checksum ? 4 : 0) > nbytes) return VXRT_ERR_INVALID;
        if (bs & 0x80000000u) { if (raw.size() + n > kChunkBytes) return VXRT_ERR_SIZE; raw.insert(raw.end(), p + i, p + i + n); }
        else if (!lz4_block_decode(p + i, n, raw, kChunkBytes)) return VXRT_ERR_INVALID;
        if (block_checksum && xxh32(p + i, n, 0) != rd32(p + i + n)) return VXRT_ERR_INVALID;
        i += n + (block_checksum ? 4 : 0);
    }
    if (raw.size() != kChunkBytes) return VXRT_ERR_SIZE;      // read_exact of both arrays (:58-68)
    if (content_checksum) { if (i + 4 > nbytes || xxh32(raw.data(), raw.size(), 0) != rd32(p + i)) return VXRT_ERR_INVALID; }
    for (size_t k = 0; k < kChunkVoxels; k++) materials[k] = rd32(raw.data() + 4 * k);
    memcpy(minefield, raw.data() + 4 * kChunkVoxels, kChunkVoxels);
    return VXRT_OK;
}

}  // extern "C"
