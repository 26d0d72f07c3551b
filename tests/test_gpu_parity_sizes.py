"""GPU parity at the sizes BASELINE.json names (north_star: "bit-exact hit/material buffers and in-tolerance colour on
every config") and on the kernel forms the small scenes never reach:

  C2  1920x1080 primary + shadow on the 512^3 biome world, two frames of the 64-frame camera path
  C3  1920x1080 full lighting, 4 accumulated frames
  C4  3840x2160 full lighting on the 2048x2048x256 world (non-cubic: k_trace_fast<1, ...>)
  C5  10 000 edits on the 512^3 world (touched chunks rebuilt), then a 1080p full-lighting frame
  non-cubic window at lr = 0 in every addressing form (A32, VXRT_NO_A32, VXRT_FLAG_GENERAL_MARCH)

The worlds come from vxrt_generate_world + vxrt_read_world (the generator is pinned against the oracle's in
test_gpu_parity.py::test_generator_and_minefield_match_oracle); the oracle renders the same arrays on the host cores.
Tolerances are those of test_gpu_parity.py: integer buffers bit-exact, hit t <= 1e-4, colour <= 1/255."""
import os
import subprocess
import sys

import numpy as np
import pytest

from test_gpu_parity import C8_TOL, _compare

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _block(orc, dims, k, n, seed, lr=(0, 0, 0)):
    from raytrace_b200 import scenes
    cam = scenes.camera(dims, k, n)
    return orc.fill_frame(cam["origin"], cam["heading"], cam["pitch"], cam["sun_angle"], seed, lr)


@pytest.fixture(scope="module")
def world512():
    """The 512^3 biome world of C2/C3/C5, generated on the device (k_generate + k_build_minefield) and read back."""
    import raytrace_b200 as R
    from raytrace_b200 import scenes
    dims, kind, seed = scenes.WORLDS["c2"]
    with R.Renderer(64, 64, dims, 0, 0) as r:
        r.generate_world(kind, seed)
        return r.read_world()


def test_c2_at_size(orc, noise, world512):
    """BASELINE config 2 at its own size: 1920x1080, primary + 1 shadow ray, 512^3 biome world; frames 0 and 37 of the
    64-frame path.  Also the A32 precondition on the 128 MiB occupancy field and the frame counters."""
    import raytrace_b200 as R
    from raytrace_b200 import _native as N, scenes
    dims = scenes.WORLDS["c2"][0]
    m, f = world512
    W, H = 1920, 1080
    with R.Renderer(W, H, dims, N.MODE_SHADOW, N.FLAG_HIT_BUFFERS | N.FLAG_DENOISE | N.FLAG_FINALIZE) as r:
        r.upload_blue_noise(noise); r.upload_world(m, f)
        for k in (0, 37):
            blk = _block(orc, dims, k, 64, k + 1)
            ref = orc.render(m, f, noise, blk, N.MODE_SHADOW, W, H, denoise=True, finalize=True)
            r.set_frame(blk); r.render_frame()
            _compare(ref, r, N, denoise=True, finalize=True)
            s = r.frame_stats()
            assert (s.rays, s.steps, s.hits, s.noise_fetches) == (ref.stats.rays, ref.stats.steps, ref.stats.hits, ref.stats.noise_fetches)
            assert s.limit_exhausted == ref.stats.ub_limit_exhausted == 0


def test_c2_two_frames_in_flight_at_size(orc, noise, world512):
    """vxrt_create_shared on the 640 MiB world: frames dealt to two handles equal the oracle's (how bench.py runs C2)."""
    import raytrace_b200 as R
    from raytrace_b200 import _native as N, scenes
    dims = scenes.WORLDS["c2"][0]
    m, f = world512
    W, H = 1920, 1080
    with R.FramePipeline(W, H, dims, N.MODE_SHADOW, N.FLAG_DENOISE | N.FLAG_FINALIZE, depth=2) as fp:
        fp.owner.upload_blue_noise(noise); fp.owner.upload_world(m, f)
        blks = [_block(orc, dims, k, 64, k + 1) for k in (10, 11)]
        used = [fp.submit(b) for b in blks]
        for r, b in zip(used, blks):
            ref = orc.render(m, f, noise, b, N.MODE_SHADOW, W, H, denoise=True, finalize=True)
            d = np.abs(r.read_output().astype(np.int16) - ref.output.astype(np.int16)).max()
            assert d <= C8_TOL
            assert np.array_equal(r.read_buffer(N.BUF_DEPTH), ref.depth) and np.array_equal(r.read_buffer(N.BUF_NORMAL), ref.normal)
            assert np.array_equal(r.read_buffer(N.BUF_ALBEDO), ref.albedo)


def test_c3_at_size(orc, noise, world512):
    """BASELINE config 3: 1080p full lighting, running mean over 4 frames (seeds 1..4), fixed camera."""
    import raytrace_b200 as R
    from raytrace_b200 import _native as N, scenes
    dims = scenes.WORLDS["c2"][0]
    m, f = world512
    W, H = 1920, 1080
    fr = orc.Frame(W, H, accumulate=True)
    with R.Renderer(W, H, dims, N.MODE_FULL, N.FLAG_HIT_BUFFERS | N.FLAG_DENOISE | N.FLAG_FINALIZE | N.FLAG_ACCUMULATE) as r:
        r.upload_blue_noise(noise); r.upload_world(m, f)
        for seed in range(1, 5):
            blk = _block(orc, dims, 0, 1, seed)
            ref = orc.render(m, f, noise, blk, N.MODE_FULL, W, H, denoise=True, finalize=True, frame=fr)
            r.set_frame(blk); r.render_frame()
        _compare(ref, r, N, denoise=True, finalize=True)
        acc = r.read_buffer(N.BUF_ACCUM)
        assert np.array_equal(acc[..., 3], fr.accum[..., 3])
        assert np.allclose(acc[..., :3], fr.accum[..., :3], rtol=1e-5, atol=1e-5)
        s = r.frame_stats()
        assert (s.rays, s.steps, s.hits) == (ref.stats.rays, ref.stats.steps, ref.stats.hits)
        assert s.limit_exhausted == ref.stats.ub_limit_exhausted == 0


def _compare_images(ref, r, N):
    """A handle without the parity buffers: the G-buffer bit for bit, colour within the tolerances of test_gpu_parity.py."""
    from test_gpu_parity import C16_TOL
    B = r.read_buffer
    for buf, want in ((N.BUF_DEPTH, ref.depth), (N.BUF_NORMAL, ref.normal), (N.BUF_ALBEDO, ref.albedo), (N.BUF_EMISSION, ref.emission)):
        assert np.array_equal(B(buf), want), buf
    d = lambda a, b: int(np.max(np.abs(a.astype(np.int64) - b.astype(np.int64)), initial=0))
    assert d(B(N.BUF_FOG), ref.fog) <= C8_TOL and d(B(N.BUF_LIGHTING), ref.lighting) <= C16_TOL
    assert d(B(N.BUF_OUTPUT), ref.output) <= C8_TOL


def test_c3_wavefront_at_size(orc, noise, world512):
    """BASELINE config 3 without parity buffers and with the trace pass in its wavefront form (trace_wave.cu):
    running mean over 4 frames at 1080p against the oracle, frame counters included."""
    import raytrace_b200 as R
    from raytrace_b200 import _native as N, scenes
    dims = scenes.WORLDS["c2"][0]
    m, f = world512
    W, H = 1920, 1080
    fr = orc.Frame(W, H, accumulate=True)
    with R.Renderer(W, H, dims, N.MODE_FULL, N.FLAG_DENOISE | N.FLAG_FINALIZE | N.FLAG_ACCUMULATE) as r:
        r.set_trace_form(N.TRACE_WAVEFRONT)
        r.upload_blue_noise(noise); r.upload_world(m, f)
        for seed in range(1, 5):
            blk = _block(orc, dims, 0, 1, seed)
            ref = orc.render(m, f, noise, blk, N.MODE_FULL, W, H, denoise=True, finalize=True, frame=fr)
            r.set_frame(blk); r.render_frame()
        assert r.launches_per_frame() == 5 + 6 + 1, "the wavefront form did not run"
        _compare_images(ref, r, N)
        acc = r.read_buffer(N.BUF_ACCUM)
        assert np.array_equal(acc[..., 3], fr.accum[..., 3])
        assert np.allclose(acc[..., :3], fr.accum[..., :3], rtol=1e-5, atol=1e-5)
        s = r.frame_stats()
        assert (s.rays, s.steps, s.hits, s.noise_fetches) == (ref.stats.rays, ref.stats.steps, ref.stats.hits, ref.stats.noise_fetches)
        assert s.limit_exhausted == ref.stats.ub_limit_exhausted == 0


@pytest.mark.parametrize("field,mode", [("brick", 1), ("brick", 2), ("nib", 1), ("nib", 2), ("tex", 2), ("texn", 2)])
def test_occupancy_field_layouts(orc, noise, world512, field, mode):
    """Layout A/B (VXRT_FIELD): the occupancy field read from a copy in 8x8x8-voxel bricks, at 4 bits per voxel, or through the
    texture unit.  Same texels at other addresses, so every buffer equals the oracle's as with the linear array; an edit list goes
    through the box-limited refresh of the copy."""
    import raytrace_b200 as R
    from raytrace_b200 import _native as N, scenes
    dims = scenes.WORLDS["c2"][0]
    m, f = world512
    W, H = 640, 360
    old = os.environ.get("VXRT_FIELD")
    os.environ["VXRT_FIELD"] = field
    try:
        with R.Renderer(W, H, dims, mode, N.FLAG_HIT_BUFFERS | N.FLAG_DENOISE | N.FLAG_FINALIZE) as r:
            r.upload_blue_noise(noise); r.upload_world(m, f)
            blk = _block(orc, dims, 9, 64, 3)
            ref = orc.render(m, f, noise, blk, mode, W, H, denoise=True, finalize=True)
            r.set_frame(blk); r.render_frame()
            _compare(ref, r, N, denoise=True, finalize=True)
            if mode == 2:
                e = orc.generate_edits(2000, *dims, 11, 0)
                r.apply_edits(e)
                m2, f2 = r.read_world()
                ref = orc.render(m2, f2, noise, blk, mode, W, H, denoise=True, finalize=True)
                r.set_frame(blk); r.render_frame()
                _compare(ref, r, N, denoise=True, finalize=True)
    finally:
        if old is None: del os.environ["VXRT_FIELD"]
        else: os.environ["VXRT_FIELD"] = old


def test_c5_at_size(orc, noise, world512):
    """BASELINE config 5: one frame's 10 000 edits on the 512^3 world (device: scatter + rebuild of the touched chunks;
    oracle: set_block + pack_into of the same chunks), the whole window compared, then a 1080p full-lighting frame."""
    import raytrace_b200 as R
    from raytrace_b200 import _native as N, scenes
    dims = scenes.WORLDS["c2"][0]
    m, f = world512
    m, f = m.copy(), f.copy()
    W, H = 1920, 1080
    e = orc.generate_edits(10000, *dims, 7, 0)
    orc.apply_edits(m, e)
    nx, ny, nz = dims
    tex = (e[:, :3] + np.array([nx // 2, ny // 2, nz // 2], np.int32)) % np.array([nx, ny, nz], np.int32)
    touched = {tuple(c) for c in (tex // 64).tolist()}
    for cx, cy, cz in touched:
        sl = (slice(cz * 64, cz * 64 + 64), slice(cy * 64, cy * 64 + 64), slice(cx * 64, cx * 64 + 64))
        m[sl], f[sl] = orc.pack_chunk(m[sl])
    with R.Renderer(W, H, dims, N.MODE_FULL, N.FLAG_HIT_BUFFERS | N.FLAG_DENOISE | N.FLAG_FINALIZE) as r:
        r.upload_blue_noise(noise); r.upload_world(*world512)
        r.apply_edits(e)
        gm, gf = r.read_world()
        assert np.array_equal(gm, m) and np.array_equal(gf, f), f"{len(touched)} chunks touched"
        blk = _block(orc, dims, 5, 64, 6)
        ref = orc.render(m, f, noise, blk, N.MODE_FULL, W, H, denoise=True, finalize=True)
        r.set_frame(blk); r.render_frame()
        _compare(ref, r, N, denoise=True, finalize=True)
        assert r.frame_stats().limit_exhausted == ref.stats.ub_limit_exhausted


def _noncubic_case(orc, noise, flags_extra=0, W=320, H=200, dims=(512, 256, 128), lr=(0, 0, 0)):
    import raytrace_b200 as R
    from raytrace_b200 import _native as N
    with R.Renderer(W, H, dims, N.MODE_FULL, N.FLAG_HIT_BUFFERS | N.FLAG_DENOISE | N.FLAG_FINALIZE | flags_extra) as r:
        r.upload_blue_noise(noise); r.generate_world(0, 6)
        m, f = r.read_world()
        for k in (0, 2):
            blk = _block(orc, dims, k, 3, 30 + k, lr)
            ref = orc.render(m, f, noise, blk, N.MODE_FULL, W, H, denoise=True, finalize=True)
            r.set_frame(blk); r.render_frame()
            _compare(ref, r, N, denoise=True, finalize=True)
            s = r.frame_stats()
            assert (s.rays, s.steps, s.hits) == (ref.stats.rays, ref.stats.steps, ref.stats.hits)


def test_noncubic_window(orc, noise):
    """A 512x256x128 window at lr = 0: the k_trace_fast<1, ...> instantiation (per-axis sky test), which every cubic scene
    skips and C4 runs."""
    _noncubic_case(orc, noise)


def test_noncubic_window_general_march(orc, noise):
    from raytrace_b200 import _native as N
    _noncubic_case(orc, noise, N.FLAG_GENERAL_MARCH)


def test_noncubic_window_slid(orc, noise):
    """The same window after it slid (lr != 0): k_trace_fast<0, ...> with host-bisected sky thresholds on unequal dims."""
    import raytrace_b200 as R
    from raytrace_b200 import _native as N, scenes
    dims = (512, 256, 128)
    W, H = 256, 160
    with R.Renderer(W, H, dims, N.MODE_FULL, N.FLAG_HIT_BUFFERS | N.FLAG_DENOISE | N.FLAG_FINALIZE) as r:
        r.upload_blue_noise(noise); r.generate_world(0, 6)
        m, f = r.read_world()
        lr = (48, -16, 16)
        cam = scenes.camera(dims)
        o = tuple(c + l for c, l in zip(cam["origin"], lr))
        blk = orc.fill_frame(o, cam["heading"], cam["pitch"], cam["sun_angle"], 9, lr)
        ref = orc.render(m, f, noise, blk, N.MODE_FULL, W, H, denoise=True, finalize=True)
        r.set_frame(blk); r.render_frame()
        _compare(ref, r, N, denoise=True, finalize=True)


_NO_A32_SCRIPT = r"""
import sys
sys.path.insert(0, {root!r}); sys.path.insert(0, {tests!r})
from oracle import orc
import test_gpu_parity_sizes as T
T._noncubic_case(orc, orc.blue_noise())
T._noncubic_case(orc, orc.blue_noise(), dims=(128, 128, 128), W=200, H=120)
print("no-a32 ok")
"""


def test_noncubic_window_64bit_addressing():
    """VXRT_NO_A32=1 (read once per process, hence the subprocess): the A32 = false addressing form of the march."""
    env = dict(os.environ, VXRT_NO_A32="1")
    p = subprocess.run([sys.executable, "-c", _NO_A32_SCRIPT.format(root=ROOT, tests=os.path.join(ROOT, "tests"))],
                       env=env, capture_output=True, text=True, timeout=600)
    assert p.returncode == 0 and "no-a32 ok" in p.stdout, p.stdout[-2000:] + p.stderr[-4000:]


def test_field_allocation_inside_one_4gib_line(orc, noise, monkeypatch):
    """The occupancy field is placed so that it never crosses a 4 GiB line of the address space (the 32-bit address form, the kernels
    with baked dims, the texture forms and the wavefront all need that); a block that would cross is replaced by one of twice the
    size.  VXRT_FIELD_REALLOC=1 takes that path whatever address cudaMalloc returned; the frame must not notice."""
    monkeypatch.setenv("VXRT_FIELD_REALLOC", "1")
    _noncubic_case(orc, noise)
    monkeypatch.delenv("VXRT_FIELD_REALLOC")


def test_start_relocation_with_tiny_dir_y(orc, world, noise):
    """raytrace.comp:312-315 with the camera outside the -y face and a pixel column whose |dir.y| is ~1e-5: the relocated
    start is millions of voxels away, beyond what the lean march's floor trick covers.  vxrt_render_frame
    sends such frames (-origin.y > H) to the general kernel; results equal the oracle's."""
    import math
    import raytrace_b200 as R
    from raytrace_b200 import _native as N, scenes
    dims = scenes.WORLDS["mid"][0]
    m, f = world("mid")
    W, H = 136, 72
    sx0 = np.float32(100) / np.float32(W) * np.float32(2) - np.float32(1)
    for eps in (2e-5, -3e-5, 0.0):
        heading = math.atan(0.4 * float(sx0)) + eps            # dir.y = sin h - 0.4 sx cos h ~ eps at column 100 (pitch 0)
        blk = orc.fill_frame((3.3, -400.0, 90.7), heading, 0.0, 0.7, 13)
        ref = orc.render(m, f, noise, blk, N.MODE_FULL, W, H, denoise=True, finalize=True)
        assert np.abs(ref.hit_voxel[:, 100].astype(np.int64)).max() > (1 << 21)   # the case the lean march cannot address
        with R.Renderer(W, H, dims, N.MODE_FULL, N.FLAG_HIT_BUFFERS | N.FLAG_DENOISE | N.FLAG_FINALIZE) as r:
            r.upload_blue_noise(noise); r.upload_world(m, f); r.set_frame(blk); r.render_frame()
            _compare(ref, r, N, denoise=True, finalize=True)


def test_c4_at_size(orc, noise):
    """BASELINE config 4 at its own size: 3840x2160 full lighting on the 2048x2048x256 world (1 GiB occupancy field,
    4 GiB of voxel words: ~5 GiB of host memory for the read-back the oracle renders from)."""
    import raytrace_b200 as R
    from raytrace_b200 import _native as N, scenes
    dims, kind, seed = scenes.WORLDS["c4"]
    W, H = 3840, 2160
    with R.Renderer(W, H, dims, N.MODE_FULL, N.FLAG_HIT_BUFFERS | N.FLAG_DENOISE | N.FLAG_FINALIZE) as r:
        r.upload_blue_noise(noise); r.generate_world(kind, seed)
        try:
            m, f = r.read_world()
        except MemoryError:
            pytest.skip("host memory: the 5 GiB world of C4 does not fit")
        blk = _block(orc, dims, 3, 16, 4)
        ref = orc.render(m, f, noise, blk, N.MODE_FULL, W, H, denoise=True, finalize=True)
        r.set_frame(blk); r.render_frame()
        _compare(ref, r, N, denoise=True, finalize=True)
        s = r.frame_stats()
        assert (s.rays, s.steps, s.hits, s.noise_fetches) == (ref.stats.rays, ref.stats.steps, ref.stats.hits, ref.stats.noise_fetches)
        assert s.limit_exhausted == ref.stats.ub_limit_exhausted
    # the same frame without parity buffers, trace pass in its wavefront form, occupancy field through the texture unit
    with R.Renderer(W, H, dims, N.MODE_FULL, N.FLAG_DENOISE | N.FLAG_FINALIZE) as r:
        r.set_trace_form(N.TRACE_WAVEFRONT)
        r.upload_blue_noise(noise); r.generate_world(kind, seed)
        r.set_frame(blk); r.render_frame()
        assert r.launches_per_frame() == 5 + 6 + 1, "the wavefront form did not run"
        _compare_images(ref, r, N)
        s = r.frame_stats()
        assert (s.rays, s.steps, s.hits, s.noise_fetches) == (ref.stats.rays, ref.stats.steps, ref.stats.hits, ref.stats.noise_fetches)
        assert s.limit_exhausted == ref.stats.ub_limit_exhausted
