"""Sliding-window terrain streaming: the reference's TerrainUploadManager
(src/render/pipeline/terrain_upload.rs) over the C ABI's vxrt_upload_slice.

Same request policy and bookkeeping (Position :24-47, request_increase :292-322, request_decrease :324-349,
request_move_towards :351-367, at most one 16-voxel slab per frame :277-286, render offset :29-36); the window size
is the handle's (the reference fixes ROOT_BLOCK_SIZE = 256).  What differs is how a slab is put together: the
reference copies chunk pieces into a staging buffer (:84-211); here the slab is gathered as ONE world box and rolled
into texel order, using the identity the piece arithmetic implements:

    texel_i = (world_i - origin_i * 64) mod N_i ,   slab box starts at world  origin_i * 64 + 16 * num_slices_i

The toroidal scatter itself is vxrt_upload_slice (a strided device copy).
"""
import numpy as np

CHUNK_SIZE = 64
SLICE_SIZE = 16
AXIS_X, AXIS_Y, AXIS_Z = 0, 1, 2


class Position:
    """terrain_upload.rs:22-47: origin of the current region in chunks + slices of the next region already loaded."""

    def __init__(self, origin, num_loaded_slices=(0, 0, 0)):
        self.origin = tuple(int(v) for v in origin)
        self.num_loaded_slices = tuple(int(v) for v in num_loaded_slices)

    def clone(self):
        return Position(self.origin, self.num_loaded_slices)

    def render_offset(self, root_chunks):
        """:29-36 — the window centre `lr` the shader subtracts (raytrace.comp:104)."""
        return tuple((o + rc // 2) * CHUNK_SIZE + s * SLICE_SIZE for o, rc, s in zip(self.origin, root_chunks, self.num_loaded_slices))


class TerrainUploadRequest:
    """:11-19"""
    __slots__ = ("origin", "num_slices", "axis", "new_position")

    def __init__(self, origin, num_slices, axis, new_position):
        self.origin, self.num_slices, self.axis, self.new_position = tuple(origin), tuple(num_slices), axis, new_position


def gather_world_box(chunks, start, shape):
    """Materials and minefield of the world box [start, start + shape) (x, y, z order) out of a ChunkStorage
    (chunk_storage.rs:147: borrow_packed_chunk_data).  Returns ([z][y][x] u32, [z][y][x] u8 or None when some chunk
    carries no minefield -> the device rebuilds it)."""
    sx, sy, sz = shape
    m = np.empty((sz, sy, sx), np.uint32)
    f = np.empty((sz, sy, sx), np.uint8)
    have_f = True
    lo = [s // CHUNK_SIZE for s in start]
    hi = [(s + n - 1) // CHUNK_SIZE for s, n in zip(start, shape)]
    for cz in range(lo[2], hi[2] + 1):
        for cy in range(lo[1], hi[1] + 1):
            for cx in range(lo[0], hi[0] + 1):
                data = chunks.borrow_packed_chunk_data((cx, cy, cz))
                a = [max(s, c * CHUNK_SIZE) for s, c in zip(start, (cx, cy, cz))]
                b = [min(s + n, (c + 1) * CHUNK_SIZE) for s, n, c in zip(start, shape, (cx, cy, cz))]
                src = tuple(slice(a[i] - c * CHUNK_SIZE, b[i] - c * CHUNK_SIZE) for i, c in ((2, cz), (1, cy), (0, cx)))
                dst = tuple(slice(a[i] - start[i], b[i] - start[i]) for i in (2, 1, 0))
                m[dst] = data.materials[src]
                if data.minefield is None:
                    have_f = False
                else:
                    f[dst] = data.minefield[src]
    return m, (f if have_f else None)


class TerrainUploadManager:
    def __init__(self, window):
        self.window = tuple(int(v) for v in window)
        self.root_chunks = tuple(n // CHUNK_SIZE for n in self.window)         # ROOT_CHUNK_SIZE per axis
        self.slices_per_axis = tuple(n // SLICE_SIZE for n in self.window)
        start = tuple(-(rc // 2) for rc in self.root_chunks)                   # Position::default, :39-46
        self.request_queue = []
        self.cpu_position = Position(start)
        self.gpu_position = Position(start)

    # -- request policy ------------------------------------------------------------------------------------
    def request_increase(self, axis):
        """:292-322 — load the next slab from the NEXT region, then count it."""
        old = self.cpu_position.clone()
        slices, origin = list(self.cpu_position.num_loaded_slices), list(self.cpu_position.origin)
        slices[axis] += 1
        if slices[axis] == self.slices_per_axis[axis]:
            slices[axis] = 0
            origin[axis] += self.root_chunks[axis]
        self.cpu_position = Position(origin, slices)
        req_origin = list(old.origin)
        req_origin[axis] += self.root_chunks[axis]
        self.request_queue.append(TerrainUploadRequest(req_origin, old.num_loaded_slices, axis, self.cpu_position.clone()))

    def request_decrease(self, axis):
        """:324-349 — rewind to the previous slab and load it from the CURRENT region."""
        slices, origin = list(self.cpu_position.num_loaded_slices), list(self.cpu_position.origin)
        if slices[axis] == 0:
            slices[axis] = self.slices_per_axis[axis]
            origin[axis] -= self.root_chunks[axis]
        slices[axis] -= 1
        self.cpu_position = Position(origin, slices)
        self.request_queue.append(TerrainUploadRequest(origin, slices, axis, self.cpu_position.clone()))

    def request_move_towards(self, desired_center):
        """:351-367 — at most one request per call, X before Y before Z."""
        cur = self.cpu_position.render_offset(self.root_chunks)
        delta = [int(d) - c for d, c in zip(desired_center, cur)]
        for axis in (AXIS_X, AXIS_Y, AXIS_Z):
            if delta[axis] > SLICE_SIZE:
                return self.request_increase(axis)
            if -delta[axis] > SLICE_SIZE:
                return self.request_decrease(axis)

    def get_render_offset(self):
        """:288-290 — what goes into RaytraceUniformData.rotation / space_offset (pipeline.rs:204-207)."""
        return self.gpu_position.render_offset(self.root_chunks)

    # -- slab assembly + upload ----------------------------------------------------------------------------
    def build_slab(self, chunks, request):
        """(axis texel offset, materials, minefield) of one request, slab layout of vxrt_upload_slice."""
        a = request.axis
        start = [o * CHUNK_SIZE + s * SLICE_SIZE for o, s in zip(request.origin, request.num_slices)]
        shape = list(self.window)
        shape[a] = SLICE_SIZE
        m, f = gather_world_box(chunks, start, shape)
        for i in range(3):                       # off axes: world order -> texel order, texel = (16 s_i + j) mod N_i
            if i != a:
                shift = (request.num_slices[i] * SLICE_SIZE) % self.window[i]
                m = np.roll(m, shift, axis=2 - i)
                f = None if f is None else np.roll(f, shift, axis=2 - i)
        offset = (request.num_slices[a] % self.slices_per_axis[a]) * SLICE_SIZE   # :201-207
        return offset, np.ascontiguousarray(m), None if f is None else np.ascontiguousarray(f)

    def setup_next_request(self, chunks, renderer):
        """:277-286 — consume at most one request: one slab per frame."""
        if not self.request_queue:
            return False
        req = self.request_queue.pop(0)
        offset, m, f = self.build_slab(chunks, req)
        renderer.upload_slice(req.axis, offset, m, f)
        self.gpu_position = req.new_position
        return True


def expected_window(chunks, window, render_offset):
    """The window the shader sees for centre `lr` (raytrace.comp:104-106,137): texel (w + N/2) mod N holds world w
    for w in [lr - N/2, lr + N/2).  Independent of the slab arithmetic above; tests compare the device against it."""
    start = [lr - n // 2 for lr, n in zip(render_offset, window)]
    m, f = gather_world_box(chunks, start, window)
    for i in range(3):
        shift = (start[i] + window[i] // 2) % window[i]
        m = np.roll(m, shift, axis=2 - i)
        f = None if f is None else np.roll(f, shift, axis=2 - i)
    return m, f
