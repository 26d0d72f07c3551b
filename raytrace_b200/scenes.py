"""Seeded scenes shared by tests/, bench.py and __graft_entry__.smoke(): BASELINE.json configs C1..C5
scaled by a size argument.  World + camera + frame blocks + edit streams; no reference, oracle or library code is
needed to define them (bench.py's reference arm loads this file by path, without importing the package)."""
import math
import struct

# name -> (window dims, generator kind, world seed)
WORLDS = {
    "c1": ((128, 128, 128), 0, 1),            # 128^3 heightmap (BASELINE config 1)
    "c2": ((512, 512, 512), 1, 2),            # 512^3 biome world (configs 2, 3, 5)
    "c4": ((2048, 2048, 256), 0, 4),          # 2048x2048x256 multi-chunk world (config 4)
    "small": ((64, 64, 64), 0, 3),            # single chunk, unit tests
    "mid": ((256, 256, 256), 1, 5),           # reference window size
}


def camera(window, k=0, n=1):
    """Camera k of an n-frame path: stands just inside the -Y face, looks along +Y and down;
    the heading sweeps +-0.5 rad over the path (BASELINE config 2: 64-frame camera path)."""
    nx, ny, nz = window
    sweep = 0.0 if n <= 1 else (k / (n - 1) - 0.5)
    origin = (0.3 + 0.2 * nx * sweep, -ny / 2 + 1.5, 0.3 * nz + 0.25)
    return dict(origin=origin, heading=math.pi / 2 + sweep, pitch=-0.35, sun_angle=0.7)


def _f32(x):
    return struct.unpack("<f", struct.pack("<f", x))[0]


def frame_block(cam, seed, lr=(0, 0, 0)):
    """The 192-byte RaytraceUniformData image of one frame (structs.rs:3-31, filled as pipeline.rs:195-207 does from
    util.rs:9-22's basis): origin, forward, up * 0.4, right * 0.4, seed, sun_angle, rotation = space_offset = lr.
    Trigonometry in double, rounded once to fp32, products in fp32 - a deterministic input of the boundary on every
    host (parity is defined on identical blocks)."""
    h, p = _f32(cam["heading"]), _f32(cam["pitch"])
    q = _f32(p + _f32(math.pi / 2))
    ch, sh, cp, sp, cq, sq = (_f32(v) for v in (math.cos(h), math.sin(h), math.cos(p), math.sin(p), math.cos(q), math.sin(q)))
    fwd = [_f32(ch * cp), _f32(sh * cp), sp]
    up = [_f32(ch * cq), _f32(sh * cq), sq]
    right = [_f32(_f32(fwd[1] * up[2]) - _f32(fwd[2] * up[1])), _f32(_f32(fwd[2] * up[0]) - _f32(fwd[0] * up[2])),
             _f32(_f32(fwd[0] * up[1]) - _f32(fwd[1] * up[0]))]
    k = _f32(0.4)
    b = bytearray(192)
    struct.pack_into("<fI", b, 0, cam["sun_angle"], seed)
    struct.pack_into("<3f", b, 16, *cam["origin"])
    struct.pack_into("<3f", b, 32, *fwd)
    struct.pack_into("<3f", b, 48, *[_f32(v * k) for v in up])
    struct.pack_into("<3f", b, 64, *[_f32(v * k) for v in right])
    struct.pack_into("<3i", b, 160, *lr)
    struct.pack_into("<3i", b, 176, *lr)
    return bytes(b)


def path_blocks(window, n_path, count, first=0, stride=1):
    """Blocks of frames first, first + stride, ... of an n_path-frame camera path; seed = frame index + 1
    (pipeline.rs:201 pre-increments)."""
    return [frame_block(camera(window, i % n_path, n_path), i + 1) for i in range(first, first + count * stride, stride)]


# ---- edit streams (BASELINE config 5) ---------------------------------------------------------------------------
_M = 0xFFFFFFFF
SOLID_EDIT_WORD = 0xCD9B3                      # material 4 of misc/materials.csv through GEN_MATERIALS.rs:44-51


def _mix32(a):
    a ^= a >> 16; a = (a * 0x7feb352d) & _M; a ^= a >> 15; a = (a * 0x846ca68b) & _M; a ^= a >> 16
    return a


def _hash3(x, y, z, s):
    return _mix32(((x * 0x9E3779B1) & _M) ^ _mix32(((y * 0x85EBCA77) & _M) ^ _mix32(((z * 0xC2B2AE3D) & _M) ^ s)))


def edit_stream(n, window, seed, frame):
    """n edits of frame `frame` as int32[n][4] = (x, y, z, word) in WORLD coordinates: n/2 distinct voxels of one
    64^3 hot cube, the rest scattered over the window (never inside the hot cube), alternately set (material 4) and
    cleared; every voxel appears once.  Integer-only (affine permutations of the index), so the oracle's
    orc_generate_edits yields the same list bit for bit (tests/test_oracle.py)."""
    import numpy as np
    nx, ny, nz = window
    hbx, hby = _hash3(frame, 0, 11, seed) % (nx - 63), _hash3(frame, 1, 11, seed) % (ny - 63)
    hbz = nz // 4 if nz // 4 <= nz - 64 else 0
    V = nx * ny * nz
    mul_hot, add_hot = _hash3(frame, 2, 11, seed) | 1, _hash3(frame, 3, 11, seed)
    mul_all, add_all = (_hash3(frame, 4, 11, seed) << 1) | 1, _hash3(frame, 5, 11, seed)
    n_hot = min(n // 2, 1 << 18)
    out = np.empty((n, 4), np.int32)
    i = np.arange(n_hot, dtype=np.uint64)
    v = (np.uint64(mul_hot) * i + np.uint64(add_hot)) & np.uint64(0x3FFFF)
    out[:n_hot, 0] = (hbx + (v & np.uint64(63))).astype(np.int64) - nx // 2
    out[:n_hot, 1] = (hby + ((v >> np.uint64(6)) & np.uint64(63))).astype(np.int64) - ny // 2
    out[:n_hot, 2] = (hbz + (v >> np.uint64(12))).astype(np.int64) - nz // 2
    need, span = n - n_hot, 2 * (n - n_hot) + 64
    if need > 0 and V <= 64 ** 3:
        raise ValueError("edit_stream: the window is one hot cube; no voxel outside it to scatter edits over")
    while need > 0:                                           # the whole-window permutation in index order, hot-cube voxels skipped
        j = np.arange(span, dtype=np.uint64)
        with np.errstate(over="ignore"):
            v = (np.uint64(mul_all) * j + np.uint64(add_all)) & np.uint64(V - 1)
        x, y, z = v % np.uint64(nx), (v // np.uint64(nx)) % np.uint64(ny), v // np.uint64(nx * ny)
        keep = ~((x >= hbx) & (x < hbx + 64) & (y >= hby) & (y < hby + 64) & (z >= hbz) & (z < hbz + 64))
        if int(keep.sum()) < need:
            span *= 2
            continue
        xyz = np.stack([x[keep], y[keep], z[keep]], 1)[:need].astype(np.int64)
        out[n_hot:, 0] = xyz[:, 0] - nx // 2; out[n_hot:, 1] = xyz[:, 1] - ny // 2; out[n_hot:, 2] = xyz[:, 2] - nz // 2
        need = 0
    out[:, 3] = np.where(np.arange(n) & 1, 0, SOLID_EDIT_WORD).astype(np.int32)
    return out
