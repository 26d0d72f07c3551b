"""Sliding-window streaming: the TerrainUploadManager policy of terrain_upload.rs over
vxrt_upload_slice.  The invariant checked everywhere: after every consumed request the window equals what the shader
addresses for the render offset the manager reports — texel (w + N/2) mod N holds world voxel w for w in
[lr - N/2, lr + N/2) (raytrace.comp:104-106,137)."""
import numpy as np
import pytest


def _storage(seed, lo=-3, hi=3, with_minefield=True, orc=None):
    import raytrace_b200 as R
    rng = np.random.default_rng(seed)
    st = R.ChunkStorage()
    for cz in range(lo, hi):
        for cy in range(lo, hi):
            for cx in range(lo, hi + 1):
                if orc is None:
                    m = rng.integers(0, 2 ** 20, size=(64, 64, 64), dtype=np.uint32)
                    f = rng.integers(0, 7, size=(64, 64, 64), dtype=np.uint8) if with_minefield else None
                else:                              # terrain-like chunk: solid below a noisy height, packed by the oracle
                    h = (20 + 6 * (cx + cy) + rng.integers(0, 12, size=(64, 64))) - cz * 64
                    z = np.arange(64)[:, None, None]
                    words = np.where(z < h[None], np.uint32(0x9F73D), np.uint32(0)).astype(np.uint32)
                    m, f = orc.pack_chunk(words)
                st.store((cx, cy, cz), R.PackedChunkData(m, f))
    return st


PATH = [(40, 0, 0)] * 4 + [(40, 0, -50)] * 5 + [(-60, 0, -50)] * 12 + [(200, 0, 30)] * 14


def test_request_policy_and_slab_arithmetic():
    """CPU: the slabs the manager builds keep a host window equal to expected_window through increases, decreases and
    region roll-overs on two axes; one request per call, X before Z, render offset = (origin + n/2) * 64 + slices * 16."""
    from raytrace_b200.terrain_upload import TerrainUploadManager, expected_window
    st = _storage(0)
    win = (128, 128, 128)
    tum = TerrainUploadManager(win)
    assert tum.get_render_offset() == (0, 0, 0)
    m, f = expected_window(st, win, (0, 0, 0))

    class Host:
        def upload_slice(self, axis, off, sm, sf):
            sl = [slice(None)] * 3; sl[2 - axis] = slice(off, off + 16)
            m[tuple(sl)] = sm; f[tuple(sl)] = sf
    seen = set()
    for tgt in PATH:
        before = len(tum.request_queue)
        tum.request_move_towards(tgt)
        assert len(tum.request_queue) - before <= 1
        tum.setup_next_request(st, Host())
        lr = tum.get_render_offset()
        seen.add(lr)
        em, ef = expected_window(st, win, lr)
        assert np.array_equal(m, em) and np.array_equal(f, ef), (tgt, lr)
    assert tum.get_render_offset() == (176, 0, -48) and len(seen) >= 15
    assert tum.gpu_position.origin == (1, -1, -3) and tum.gpu_position.num_loaded_slices == (3, 0, 5)
    tum.request_move_towards((176, 0, -48))      # within one slice of the centre: nothing to do (:351-367)
    assert not tum.request_queue


@pytest.mark.gpu
def test_streaming_pipeline_matches_oracle(orc, noise):
    """GPU: Pipeline(streaming=True) follows a moving camera one slab per frame; the device window equals
    expected_window for the reported offset, and the frame equals the oracle's for that window and lr."""
    import raytrace_b200 as R
    from raytrace_b200 import _native as N
    from raytrace_b200.terrain_upload import expected_window
    st = _storage(3, orc=orc)
    win = (128, 128, 128)
    game = R.Game(args=[0.5, -60.0, 40.25, 1.5707963, -0.3, 0.7], world=st)
    pipe = R.create_instance(game, width=96, height=64, window=win, blue_noise=noise, streaming=True,
                             flags=N.FLAG_DENOISE | N.FLAG_FINALIZE | N.FLAG_HIT_BUFFERS)
    try:
        offsets = []
        for k in range(10):
            game.camera.origin[0] += 9.0             # drifts +x faster than one slab per two frames
            img = pipe.draw_frame(game).copy()
            lr = pipe.render_offset
            offsets.append(lr)
            gm, gf = pipe.renderer.read_world()
            em, ef = expected_window(st, win, lr)
            assert np.array_equal(gm, em) and np.array_equal(gf, ef), (k, lr)
            blk = np.frombuffer(bytes(pipe.uniform_data), np.uint8)
            ref = orc.render(em, ef, noise, blk, 2, 96, 64, denoise=True, finalize=True)
            assert np.array_equal(pipe.renderer.read_buffer(N.BUF_HIT_VOXEL), ref.hit_voxel), k
            assert np.array_equal(pipe.renderer.read_buffer(N.BUF_HIT_WORD), ref.hit_word), k
            assert np.max(np.abs(img.astype(int) - ref.output.astype(int))) <= 1, k
        assert offsets[-1][0] >= 48 and offsets[-1][1] == 0     # y is pinned (pipeline.rs:174-178); z follows the camera too
    finally:
        pipe.close()


@pytest.mark.gpu
def test_cpp_host_streams_across_a_region_boundary(orc, noise, tmp_path):
    """The compiled host (csrc/host/vxrt_host.hpp): Pipeline::draw_frame drives its own TerrainUploadManager as
    pipeline.rs:174-189 does.  The headless driver flies +x for 12 frames over a world twice as wide as the window, one slab
    per frame, through a region roll-over; the device window must equal expected_window for the render offset the driver
    put into its last uniform block, occupancy field included (built on the device: the chunks carry none)."""
    import os, subprocess
    import raytrace_b200 as R
    from raytrace_b200.terrain_upload import expected_window
    exe = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "raytrace_b200", "vxrt_main")
    assert os.path.exists(exe), "vxrt_main not built (raytrace_b200/build.py)"
    wd, win = (256, 128, 128), (128, 128, 128)
    m = orc.generate_world(*wd, 0, 11)
    (tmp_path / "w.u32").write_bytes(np.ascontiguousarray(m).tobytes())
    (tmp_path / "n.rgba").write_bytes(np.ascontiguousarray(noise).tobytes())
    cmd = [exe, "--world", str(tmp_path / "w.u32"), "--world-dims", *map(str, wd), "--dims", *map(str, win), "--noise", str(tmp_path / "n.rgba"),
           "--size", "96", "64", "--frames", "12", "--hold", "right", "--dt", "0.4", "--mode", "2", "--stream",
           "--camera", "0.5", "-60.0", "45.25", "1.5707963", "-0.3", "0.7",
           "--block", str(tmp_path / "b.bin"), "--dump-world", str(tmp_path / "d.bin")]
    p = subprocess.run(cmd, capture_output=True, text=True, timeout=180)
    assert p.returncode == 0, p.stderr
    blk = np.frombuffer((tmp_path / "b.bin").read_bytes(), np.uint8)
    lr = tuple(int(v) for v in blk[160:172].view(np.int32))
    assert lr[0] >= 128 + 16 and lr[1] == 0, lr                     # past the first region (8 slabs of 16) on x; y pinned
    raw = (tmp_path / "d.bin").read_bytes()
    V = win[0] * win[1] * win[2]
    gm = np.frombuffer(raw[:4 * V], np.uint32).reshape(win[2], win[1], win[0])
    gf = np.frombuffer(raw[4 * V:], np.uint8).reshape(win[2], win[1], win[0])
    st = R.ChunkStorage.from_window(m)                              # the same chunk cut as ChunkStorage::store_window
    em, _ = expected_window(st, win, lr)
    assert np.array_equal(gm, em)
    _, ef = orc.build_minefield(em)                                 # pack_into per 64^3 chunk: window chunks are world chunks
    assert np.array_equal(gf, ef)
