"""Host-side world containers mirroring src/world: PackedChunkData (chunk.rs:52-101) and the part of
ChunkStorage the renderer calls (chunk_storage.rs:147-151).  No arithmetic of the hot path lives here:
occupancy ("minefield") data a chunk does not carry is built on the GPU (k_build_minefield)."""
import numpy as np

CHUNK_SIZE = 64


def default_blue_noise(seed=0x5EED):
    """Seeded 512x512 RGBA8 noise.  The reference embeds blue_noise_512.png (render_data.rs:130); it is
    an input of the boundary (vxrt_upload_blue_noise), and parity is defined on equal input bytes."""
    rng = np.random.Generator(np.random.PCG64(seed))
    return rng.integers(0, 256, size=(512, 512, 4), dtype=np.uint8)


class PackedChunkData:
    """chunk.rs:52-56: materials u32 x 64^3 (x fastest), minefield u8 x 64^3 (None -> built on device)."""
    __slots__ = ("materials", "minefield")

    def __init__(self, materials=None, minefield=None):
        self.materials = np.zeros((64, 64, 64), np.uint32) if materials is None else \
            np.ascontiguousarray(materials, np.uint32).reshape(64, 64, 64)
        self.minefield = None if minefield is None else np.ascontiguousarray(minefield, np.uint8).reshape(64, 64, 64)


class ChunkStorage:
    """Chunk provider keyed by signed chunk coordinates, as chunk_storage.rs:147-151 hands them to
    make_world_upload_buffers (render_data.rs:221-241).  Chunks never stored read as empty air."""

    def __init__(self):
        self.chunks = {}

    @classmethod
    def from_window(cls, materials, minefield=None):
        """Cut an [nz][ny][nx] window (texel = world + N/2) into chunks at coords -n/2 .. n/2-1."""
        st = cls()
        nz, ny, nx = materials.shape
        for cz in range(nz // 64):
            for cy in range(ny // 64):
                for cx in range(nx // 64):
                    sl = (slice(cz * 64, cz * 64 + 64), slice(cy * 64, cy * 64 + 64), slice(cx * 64, cx * 64 + 64))
                    st.chunks[(cx - nx // 128, cy - ny // 128, cz - nz // 128)] = PackedChunkData(
                        materials[sl], None if minefield is None else minefield[sl])
        return st

    def store(self, coord, data):
        self.chunks[tuple(coord)] = data

    def borrow_packed_chunk_data(self, coord):
        c = tuple(coord)
        if c not in self.chunks:
            self.chunks[c] = PackedChunkData()
        return self.chunks[c]


# ---- chunk files (src/world/chunk_storage.rs:37-68) ---------------------------------------------------------------
def chunk_file_name(coord):
    """{:016X}{:016X}{:016X} of the signed chunk coordinates (chunk_storage.rs:38-41)."""
    import ctypes as C
    from . import _native as N
    buf = C.create_string_buffer(64)
    assert N.lib.vxrt_chunk_file_name(int(coord[0]), int(coord[1]), int(coord[2]), buf, 64) == 0
    return buf.value.decode()


def encode_chunk(data):
    """PackedChunkData -> bytes of the reference's chunk file (LZ4 frame of materials || minefield)."""
    import ctypes as C
    from . import _native as N
    if data.minefield is None:
        raise ValueError("a stored chunk carries its minefield (chunk_storage.rs:43-56)")
    cap = N.lib.vxrt_chunk_encode_bound()
    out = (C.c_uint8 * cap)()
    n = C.c_size_t(0)
    m = np.ascontiguousarray(data.materials, np.uint32); f = np.ascontiguousarray(data.minefield, np.uint8)
    rc = N.lib.vxrt_chunk_encode(m.ctypes.data, f.ctypes.data, out, cap, C.byref(n))
    if rc != 0:
        raise ValueError(f"vxrt_chunk_encode failed ({rc})")
    return bytes(out[:n.value])


def decode_chunk(blob):
    """bytes of a chunk file -> PackedChunkData (chunk_storage.rs:58-68)."""
    from . import _native as N
    m = np.empty((64, 64, 64), np.uint32); f = np.empty((64, 64, 64), np.uint8)
    rc = N.lib.vxrt_chunk_decode(blob, len(blob), m.ctypes.data, f.ctypes.data)
    if rc != 0:
        raise ValueError(f"not a chunk file (vxrt_chunk_decode returned {rc})")
    return PackedChunkData(m, f)


class DiskChunkStorage(ChunkStorage):
    """ChunkStorage backed by a directory of chunk files in the reference's format: a chunk missing from memory is
    read from disk (chunk_storage.rs:118-141); `save` writes one.  Generation of missing chunks stays with the caller."""

    def __init__(self, storage_dir):
        super().__init__()
        self.storage_dir = storage_dir

    def path_for(self, coord):
        import os
        return os.path.join(self.storage_dir, chunk_file_name(coord))

    def save(self, coord, data):
        with open(self.path_for(coord), "wb") as fh:
            fh.write(encode_chunk(data))

    def borrow_packed_chunk_data(self, coord):
        import os
        c = tuple(coord)
        if c not in self.chunks and os.path.exists(self.path_for(c)):
            with open(self.path_for(c), "rb") as fh:
                self.chunks[c] = decode_chunk(fh.read())
        return super().borrow_packed_chunk_data(c)
