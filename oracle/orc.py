"""ctypes binding of the CPU ORACLE (oracle/_build/liborc.so).

TEST INFRASTRUCTURE ONLY: imported by tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
--impl reference legs.  Nothing under raytrace_b200/ may import this module.
"""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "_build", "liborc.so")

MODE_PRIMARY, MODE_SHADOW, MODE_FULL = 0, 1, 2


def build(force=False):
    srcs = [os.path.join(_HERE, f) for f in os.listdir(_HERE) if f.endswith((".cpp", ".h"))]
    if not force and os.path.exists(LIB_PATH) and all(os.path.getmtime(LIB_PATH) >= os.path.getmtime(s) for s in srcs):
        return
    subprocess.run(["make", "-C", _HERE, "_build/liborc.so"], check=True, stdout=subprocess.DEVNULL)


class OrcScene(C.Structure):
    _fields_ = [("nx", C.c_int32), ("ny", C.c_int32), ("nz", C.c_int32), ("materials", C.c_void_p),
                ("minefield", C.c_void_p), ("blue_noise", C.c_void_p), ("max_steps", C.c_int32)]


class OrcBuffers(C.Structure):
    _fields_ = [("width", C.c_int32), ("height", C.c_int32), ("lighting", C.c_void_p), ("depth", C.c_void_p),
                ("normal", C.c_void_p), ("albedo", C.c_void_p), ("emission", C.c_void_p), ("fog", C.c_void_p),
                ("hit_voxel", C.c_void_p), ("hit_face", C.c_void_p), ("hit_word", C.c_void_p), ("hit_t", C.c_void_p),
                ("steps", C.c_void_p), ("rays", C.c_void_p), ("accum", C.c_void_p), ("hit_pos", C.c_void_p)]


class OrcStats(C.Structure):
    _fields_ = [("rays", C.c_uint64), ("steps", C.c_uint64), ("hits", C.c_uint64), ("noise_fetches", C.c_uint64),
                ("ub_limit_exhausted", C.c_uint64), ("ub_start_in_solid", C.c_uint64), ("ub_nan_dir", C.c_uint64),
                ("ub_coord_eq_n", C.c_uint64)]


_lib = None


def lib():
    global _lib
    if _lib is None:
        build()
        L = C.CDLL(LIB_PATH)
        L.orc_pack_word.restype = C.c_uint32
        L.orc_pack_word.argtypes = [C.c_uint32, C.c_uint32, C.c_uint32, C.c_int]
        L.orc_material_word_from_csv.restype = C.c_uint32
        L.orc_material_word_from_csv.argtypes = [C.c_int] * 4
        L.orc_pack_chunk.restype = None
        L.orc_pack_chunk.argtypes = [C.c_void_p] * 3
        L.orc_build_minefield.restype = None
        L.orc_build_minefield.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int]
        L.orc_fill_frame.restype = None
        L.orc_fill_frame.argtypes = [C.c_void_p, C.c_void_p, C.c_float, C.c_float, C.c_float, C.c_uint32, C.c_void_p]
        L.orc_fill_old_transform.restype = None
        L.orc_fill_old_transform.argtypes = [C.c_void_p, C.c_void_p]
        L.orc_reproject.restype = None
        L.orc_reproject.argtypes = [C.c_void_p, C.c_int, C.c_int] + [C.c_void_p] * 6 + [C.c_int]
        L.orc_diffuse_luts.restype = None
        L.orc_diffuse_luts.argtypes = [C.c_void_p] * 4
        L.orc_trace.restype = C.c_int
        L.orc_trace.argtypes = [C.POINTER(OrcScene), C.c_void_p, C.c_int, C.POINTER(OrcBuffers), C.c_int, C.c_int,
                                C.POINTER(OrcStats), C.c_int]
        L.orc_trace_ray.restype = None
        L.orc_trace_ray.argtypes = [C.POINTER(OrcScene), C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
        L.orc_denoise_pass.restype = None
        L.orc_denoise_pass.argtypes = [C.c_void_p] * 4 + [C.c_int] * 7
        L.orc_finalize.restype = None
        L.orc_finalize.argtypes = [C.POINTER(OrcBuffers), C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int]
        L.orc_wg_to_pixel.restype = None
        L.orc_wg_to_pixel.argtypes = [C.c_int] * 4 + [C.POINTER(C.c_int)] * 2
        L.orc_pixel_to_wg.restype = None
        L.orc_pixel_to_wg.argtypes = [C.c_int] * 2 + [C.POINTER(C.c_int)] * 2
        L.orc_generate_world.restype = None
        L.orc_generate_world.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_uint32]
        L.orc_generate_edits.restype = None
        L.orc_generate_edits.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_uint32, C.c_uint32]
        L.orc_apply_edits.restype = None
        L.orc_apply_edits.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int]
        _lib = L
    return _lib


def _p(a):
    return a.ctypes.data_as(C.c_void_p) if a is not None else None


DENOISE_SIZES = (1, 2, 4, 8, 8, 16)      # pipeline.rs:103


def blue_noise(seed=0x5EED):
    """Seeded 512x512 RGBA8 stand-in for blue_noise_512.png (parity is oracle<->kernel on equal bytes)."""
    rng = np.random.Generator(np.random.PCG64(seed))
    return rng.integers(0, 256, size=(512, 512, 4), dtype=np.uint8)


def generate_world(nx, ny, nz, kind, seed):
    m = np.empty((nz, ny, nx), np.uint32)
    lib().orc_generate_world(_p(m), nx, ny, nz, kind, seed)
    return m


def build_minefield(materials):
    """pack_into over every 64^3 chunk; returns (materials', minefield)."""
    nz, ny, nx = materials.shape
    m = np.ascontiguousarray(materials.copy())
    f = np.empty((nz, ny, nx), np.uint8)
    lib().orc_build_minefield(_p(m), _p(f), nx, ny, nz)
    return m, f


def pack_chunk(words):
    w = np.ascontiguousarray(words, dtype=np.uint32).reshape(64, 64, 64)
    m = np.empty_like(w)
    f = np.empty(w.shape, np.uint8)
    lib().orc_pack_chunk(_p(w), _p(m), _p(f))
    return m, f


def generate_edits(n, nx, ny, nz, seed, frame):
    e = np.empty((n, 4), np.int32)
    lib().orc_generate_edits(_p(e), n, nx, ny, nz, seed, frame)
    return e


def apply_edits(materials, edits):
    nz, ny, nx = materials.shape
    e = np.ascontiguousarray(edits, dtype=np.int32)
    lib().orc_apply_edits(_p(materials), _p(e), len(e), nx, ny, nz)


def fill_frame(origin, heading, pitch, sun_angle, seed, lr=(0, 0, 0)):
    block = np.zeros(192, np.uint8)
    o = np.asarray(origin, np.float32)
    l = np.asarray(lr, np.int32)
    lib().orc_fill_frame(_p(block), _p(o), heading, pitch, sun_angle, seed, _p(l))
    return block


def fill_old_transform(next_block, prev_block):
    """pipeline.rs:213-227: old_origin / old_transform_c0..2 of `next_block` from the frame submitted before it (in place)."""
    assert next_block.dtype == np.uint8 and next_block.size == 192
    lib().orc_fill_old_transform(_p(next_block), _p(np.ascontiguousarray(prev_block, np.uint8)))
    return next_block


class History:
    """State of the temporal-reprojection extension between frames (double-buffered on the device; two arrays here)."""

    def __init__(self, width, height):
        self.hit_pos = np.zeros((height, width, 4), np.float32)     # previous frame's records
        self.hist = np.zeros((height, width, 4), np.float32)        # previous frame's {mean rgb, count}
        self.completed = np.zeros((height, width, 4), np.uint16)


def reproject(frame_block, fr, history, threads=0):
    """orc_reproject on Frame `fr` (needs fr.hit_pos) against `history`; updates both in place."""
    h, w = fr.height, fr.width
    cur = np.zeros_like(history.hist)
    blk = np.ascontiguousarray(frame_block, np.uint8)
    lib().orc_reproject(_p(blk), w, h, _p(fr.lighting), _p(fr.hit_pos), _p(history.hit_pos), _p(history.hist), _p(cur),
                        _p(history.completed), threads)
    history.hist = cur
    history.hit_pos = fr.hit_pos.copy()
    return history


def diffuse_luts():
    t = [np.empty(256, np.float32) for _ in range(4)]
    lib().orc_diffuse_luts(*[_p(a) for a in t])
    return t


def trace_ray(materials, minefield, noise, lr, origin, direction, max_steps=2048):
    nz, ny, nx = minefield.shape
    sc = OrcScene(nx, ny, nz, _p(materials), _p(minefield), _p(noise), max_steps)
    l = np.asarray(lr, np.int32); o = np.asarray(origin, np.float32); d = np.asarray(direction, np.float32)
    io = np.zeros(7, np.int32); fo = np.zeros(4, np.float32)
    lib().orc_trace_ray(C.byref(sc), _p(l), _p(o), _p(d), _p(io), _p(fo))
    return dict(air=bool(io[0]), face=int(io[1]), steps=int(io[2]), word=int(io[3]) & 0xFFFFFFFF,
                voxel=tuple(int(v) for v in io[4:7]), t=float(fo[0]), position=tuple(float(v) for v in fo[1:4]))


class Frame:
    """All per-pixel outputs of one oracle frame (numpy arrays)."""

    def __init__(self, width, height, accumulate=False, hit_pos=False):
        P = (height, width)
        self.hit_pos = np.zeros(P + (4,), np.float32) if hit_pos else None
        self.width, self.height = width, height
        self.lighting = np.zeros(P + (4,), np.uint16)
        self.depth = np.zeros(P, np.uint16)
        self.normal = np.zeros(P, np.uint8)
        self.albedo = np.zeros(P + (4,), np.uint8)
        self.emission = np.zeros(P + (4,), np.uint8)
        self.fog = np.zeros(P + (4,), np.uint8)
        self.hit_voxel = np.zeros(P + (3,), np.int32)
        self.hit_face = np.zeros(P, np.uint8)
        self.hit_word = np.zeros(P, np.uint32)
        self.hit_t = np.zeros(P, np.float32)
        self.steps = np.zeros(P, np.uint32)
        self.rays = np.zeros(P, np.uint8)
        self.accum = np.zeros(P + (4,), np.float32) if accumulate else None
        self.lighting_raw = None
        self.output = np.zeros(P + (4,), np.uint8)
        self.stats = OrcStats()

    def buffers(self):
        return OrcBuffers(self.width, self.height, _p(self.lighting), _p(self.depth), _p(self.normal), _p(self.albedo),
                          _p(self.emission), _p(self.fog), _p(self.hit_voxel), _p(self.hit_face), _p(self.hit_word),
                          _p(self.hit_t), _p(self.steps), _p(self.rays), _p(self.accum), _p(self.hit_pos))


def render(materials, minefield, noise, frame_block, mode, width, height, *, denoise=False, finalize=False,
           pong_swap=False, frame=None, rows=None, max_steps=2048, threads=0, history=None):
    """raytrace.comp -> [6x bilateral_denoise.comp] -> [finalize.comp] on the CPU (pipeline.rs:86-123)."""
    nz, ny, nx = minefield.shape
    sc = OrcScene(nx, ny, nz, _p(materials), _p(minefield), _p(noise), max_steps)
    fr = frame if frame is not None else Frame(width, height, hit_pos=history is not None)
    r0, r1 = rows if rows is not None else (0, height)
    b = fr.buffers()
    blk = np.ascontiguousarray(frame_block, np.uint8)
    lib().orc_trace(C.byref(sc), _p(blk), mode, C.byref(b), r0, r1, C.byref(fr.stats), threads)
    fr.lighting_raw = fr.lighting.copy()
    if history is not None:                      # temporal reprojection (extension): between trace and denoise
        reproject(blk, fr, history, threads)
    if denoise:
        pong = np.zeros_like(fr.lighting)
        for k, size in enumerate(DENOISE_SIZES):
            src, dst = (fr.lighting, pong) if k % 2 == 0 else (pong, fr.lighting)
            lib().orc_denoise_pass(_p(src), _p(dst), _p(fr.depth), _p(fr.normal), width, height, size,
                                   1 if (pong_swap and k % 2 == 1) else 0, r0, r1, threads)
    if finalize:
        lib().orc_finalize(C.byref(b), _p(fr.lighting), _p(noise), _p(fr.output), r0, r1, threads)
    return fr


# ---- oracle/_ref: the reference's own shader text compiled as C++ (oracle/make_ref.py) -------------------
REF_LIB_PATH = os.path.join(_HERE, "_ref", "libref.so")
_ref = None


def ref_available():
    return os.path.exists(REF_LIB_PATH)


def ref_lib():
    global _ref
    if _ref is None:
        L = C.CDLL(REF_LIB_PATH)
        L.ref_trace.restype = C.c_int
        L.ref_trace.argtypes = [C.c_void_p] * 4 + [C.c_int, C.c_int] + [C.c_void_p] * 6 + [C.c_int]
        L.ref_denoise.restype = C.c_int
        L.ref_denoise.argtypes = [C.c_void_p] * 4 + [C.c_int] * 4
        L.ref_finalize.restype = C.c_int
        L.ref_finalize.argtypes = [C.c_void_p] * 7 + [C.c_int, C.c_int]
        _ref = L
    return _ref


def ref_render(materials, minefield, noise, frame_block, width, height, *, denoise=True, finalize=True,
               pong_swap=False, cover_all_rows=True):
    """The reference's three shaders over a 256^3 window (its compile-time size), full lighting.
    Returns a Frame holding the G-buffer, lighting_raw, lighting after denoise and the output image."""
    assert minefield.shape == (256, 256, 256), "the reference shader hard-codes ROOT_BLOCK_WIDTH = 256"
    fr = Frame(width, height)
    blk = np.ascontiguousarray(frame_block, np.uint8)
    L = ref_lib()
    L.ref_trace(_p(materials), _p(minefield), _p(noise), _p(blk), width, height, _p(fr.lighting), _p(fr.depth),
                _p(fr.normal), _p(fr.albedo), _p(fr.emission), _p(fr.fog), 1 if cover_all_rows else 0)
    fr.lighting_raw = fr.lighting.copy()
    if denoise:
        pong = np.zeros_like(fr.lighting)
        for k, size in enumerate(DENOISE_SIZES):
            src, dst = (fr.lighting, pong) if k % 2 == 0 else (pong, fr.lighting)
            L.ref_denoise(_p(src), _p(dst), _p(fr.depth), _p(fr.normal), width, height, size,
                          1 if (pong_swap and k % 2 == 1) else 0)
    if finalize:
        L.ref_finalize(_p(fr.albedo), _p(fr.emission), _p(fr.fog), _p(fr.lighting), _p(fr.depth), _p(noise),
                       _p(fr.output), width, height)
    return fr
