"""Temporal reprojection (VXRT_FLAG_REPROJECT; extension).

The reference only prepares this pass's inputs (old_origin / old_transform_c0..2, pipeline.rs:213-227; MAX_SAMPLES,
raytrace.comp:58; completed_buffer, :19), so the CPU tests pin (i) those inputs against their definition and (ii) the
properties the rule in include/vxrt.h must have; the GPU tests compare libvxrt with the oracle: sample counts and hit
records bit for bit, history and images bit for bit when both sides start from the same lighting image, and within
the colour tolerance end to end."""
import math

import numpy as np
import pytest

W, H = 160, 96
C16_TOL = 257


def _blocks(orc, dims, path):
    """Frame blocks of a camera path [(k, n)...] with old_* chained as draw_frame does (pipeline.rs:213-227)."""
    from raytrace_b200 import scenes
    out, prev = [], None
    for i, (k, n) in enumerate(path):
        cam = scenes.camera(dims, k, n)
        blk = orc.fill_frame(cam["origin"], cam["heading"], cam["pitch"], cam["sun_angle"], i + 1)
        if prev is not None:
            orc.fill_old_transform(blk, prev)
        out.append(blk)
        prev = blk
    return out


def _f3(blk, off):
    return np.frombuffer(blk.tobytes()[off:off + 12], np.float32).astype(np.float64)


def test_old_transform_is_the_inverse_camera_matrix(orc):
    """pipeline.rs:215-226: old_transform * (right*0.4*sx + up*0.4*sy + forward) * depth == (sx*d, sy*d, d), i.e. the
    matrix maps a world offset back to screen coordinates; the first frame's block has it all zero
    (render_data.rs:143-146)."""
    from raytrace_b200 import scenes
    dims = scenes.WORLDS["c1"][0]
    b0, b1 = _blocks(orc, dims, [(2, 8), (3, 8)])
    assert not b0[80:144].any()
    assert np.array_equal(b1[80:92], b0[16:28])                       # old_origin = previous origin
    fwd, up, right = _f3(b0, 32), _f3(b0, 48), _f3(b0, 64)
    Minv = np.stack([_f3(b1, 96), _f3(b1, 112), _f3(b1, 128)], axis=1)   # columns c0, c1, c2
    M = np.stack([right, up, fwd], axis=1)
    assert np.allclose(Minv @ M, np.eye(3), atol=2e-6)
    for sx, sy, d in [(-0.5, 0.25, 10.0), (0.9, -0.9, 80.0)]:
        v = Minv @ ((right * sx + up * sy + fwd) * d)
        assert np.allclose(v, [sx * d, sy * d, d], rtol=1e-5, atol=1e-5)


def test_host_mirror_fills_old_transform_like_the_oracle(orc):
    """raytrace_b200.render.fill_old_transform (what Pipeline.draw_frame runs) against orc_fill_old_transform."""
    import raytrace_b200 as R
    from raytrace_b200 import _native as N, scenes
    import ctypes as C
    dims = scenes.WORLDS["c1"][0]
    b0, b1 = _blocks(orc, dims, [(1, 8), (2, 8)])
    u = N.VxrtFrameParams.from_buffer_copy(b0.tobytes())
    R.render.fill_old_transform(u)
    got = np.frombuffer(C.string_at(C.addressof(u), 192), np.uint8)
    assert np.array_equal(got[80:144], b1[80:144])


def test_static_camera_counts_saturate_and_mean_is_running_mean(orc, world, noise):
    """Same camera, new seed every frame: every hit pixel re-finds its own voxel face, the count climbs 1..8 and stays
    (MAX_SAMPLES), sky pixels stay at 1, and the history equals the running mean of the lighting images."""
    from raytrace_b200 import scenes
    dims = scenes.WORLDS["c1"][0]
    m, f = world("c1")
    hist = orc.History(W, H)
    blocks = _blocks(orc, dims, [(0, 1)] * 10)
    raws = []
    for i, blk in enumerate(blocks):
        fr = orc.render(m, f, noise, blk, orc.MODE_FULL, W, H, history=hist)
        hit = fr.hit_pos.view(np.int32)[..., 3] < 16
        assert hit.any() and (~hit).any()
        cnt = hist.hist[..., 3]
        assert (cnt[hit] == min(i + 1, 8)).all() and (cnt[~hit] == 1).all()
        raws.append(fr.lighting_raw[..., :3].astype(np.float64) / 65535 * 16)
        if i < 8:
            assert np.allclose(hist.hist[..., :3][hit], np.mean(raws, axis=0)[hit], rtol=1e-5, atol=1e-5)
        assert np.array_equal(hist.completed[..., 3][hit], np.full(hit.sum(), round(min(i + 1, 8) / 8 * 65535), np.uint16))
        # the lighting image the denoiser sees is the history, quantised
        assert np.array_equal(hist.completed[..., :3], fr.lighting[..., :3])


def test_moving_camera_reuses_matching_faces_only(orc, world, noise):
    """Under camera motion a pixel keeps history only if the reprojected pixel saw the same voxel face; every pixel whose
    count is n > 1 therefore has a predecessor with count >= n - 1 holding the same key."""
    from raytrace_b200 import scenes
    dims = scenes.WORLDS["c1"][0]
    m, f = world("c1")
    hist = orc.History(W, H)
    path = [(0, 64)] * 3 + [(k, 64) for k in range(1, 5)]
    reused = []
    for i, blk in enumerate(_blocks(orc, dims, path)):
        prev_pos, prev_cnt = hist.hit_pos.copy(), hist.hist[..., 3].copy()
        fr = orc.render(m, f, noise, blk, orc.MODE_FULL, W, H, history=hist)
        cnt = hist.hist[..., 3]
        face = fr.hit_pos.view(np.int32)[..., 3]
        assert (cnt[face == 16] == 1).all()
        if i >= 3:
            again = cnt > 1
            reused.append(again.mean())
            keys_prev = {tuple(np.floor(p[:3]).astype(int)) + (int(np.float32(p[3]).view(np.int32)),)
                         for p, c in zip(prev_pos.reshape(-1, 4), prev_cnt.ravel()) if c >= 1}
            for p in fr.hit_pos[again][:200]:
                assert tuple(np.floor(p[:3]).astype(int)) + (int(np.float32(p[3]).view(np.int32)),) in keys_prev
    assert all(0.05 < r < 1.0 for r in reused), reused


def test_flag_combinations_are_rejected_without_a_device():
    """vxrt_create validates REPROJECT's restrictions before it looks for a device."""
    import ctypes as C
    from raytrace_b200 import _native as N
    for flags, rows in [(N.FLAG_REPROJECT | N.FLAG_ACCUMULATE, (0, 0)), (N.FLAG_REPROJECT, (0, 48))]:
        cfg = N.VxrtConfig(N.ABI_VERSION, W, H, 64, 64, 64, N.MODE_FULL, flags, 0, 0, rows[0], rows[1])
        h = C.c_void_p()
        assert N.lib.vxrt_create(C.byref(cfg), C.byref(h)) == N.ERR_INVALID
        assert b"REPROJECT" in N.lib.vxrt_last_error(None)


# ---- GPU ------------------------------------------------------------------------------------------------------------

@pytest.mark.gpu
@pytest.mark.parametrize("general", [False, True], ids=["fast", "general"])
def test_reprojection_matches_oracle(orc, world, noise, general):
    import raytrace_b200 as R
    from raytrace_b200 import _native as N, scenes
    dims = scenes.WORLDS["c1"][0]
    m, f = world("c1")
    path = [(0, 64)] * 3 + [(k, 64) for k in range(1, 6)] + [(5, 64)] * 2
    flags = N.FLAG_REPROJECT | N.FLAG_HIT_BUFFERS | N.FLAG_DENOISE | N.FLAG_FINALIZE | (N.FLAG_GENERAL_MARCH if general else 0)
    hist = orc.History(W, H)           # end to end on the CPU
    iso = orc.History(W, H)            # fed with the GPU's own trace output: isolates k_reproject, bit-exact
    d = lambda a, b: int(np.max(np.abs(a.astype(np.int64) - b.astype(np.int64)), initial=0))
    with R.Renderer(W, H, dims, N.MODE_FULL, flags) as r:
        r.upload_blue_noise(noise)
        r.upload_world(m, f)
        for i, blk in enumerate(_blocks(orc, dims, path)):
            ref = orc.render(m, f, noise, blk, orc.MODE_FULL, W, H, denoise=True, finalize=True, history=hist)
            r.set_frame(blk)
            r.render_frame()
            pos = r.read_buffer(N.BUF_HIT_POS)
            assert np.array_equal(pos.view(np.uint32), ref.hit_pos.view(np.uint32)), "hit records"
            g_hist = r.read_buffer(N.BUF_HISTORY)
            assert np.array_equal(g_hist[..., 3], hist.hist[..., 3]), "sample counts"
            assert np.max(np.abs(g_hist[..., :3] - hist.hist[..., :3])) <= 16.0 * C16_TOL / 65535, "history (colour tolerance)"
            assert d(r.read_buffer(N.BUF_COMPLETED), hist.completed) <= C16_TOL
            assert d(r.read_buffer(N.BUF_LIGHTING), ref.lighting) <= C16_TOL
            assert d(r.read_output(), ref.output) <= 1
            # isolated: the oracle's pass over the GPU's trace-pass images must give the GPU's history bit for bit
            fr = orc.Frame(W, H, hit_pos=True)
            fr.lighting[...] = r.read_buffer(N.BUF_LIGHTING_RAW)
            fr.hit_pos[...] = pos
            orc.reproject(blk, fr, iso)
            assert np.array_equal(g_hist.view(np.uint32), iso.hist.view(np.uint32)), "history, bit for bit"
            assert np.array_equal(r.read_buffer(N.BUF_COMPLETED), iso.completed)
            if i >= 4:
                assert 0 < (g_hist[..., 3] > 1).mean() < 1
        # a camera cut: history forgotten, every count back to 1
        r.reset_history()
        r.set_frame(blk)
        r.render_frame()
        assert (r.read_buffer(N.BUF_HISTORY)[..., 3] == 1).all()


@pytest.mark.gpu
def test_pipeline_draw_frame_reprojects(orc, world, noise):
    """Through the reference-shaped interface: Pipeline.draw_frame chains old_* itself (pipeline.rs:213-227)."""
    import raytrace_b200 as R
    from raytrace_b200 import _native as N, scenes
    dims = scenes.WORLDS["c1"][0]
    m, f = world("c1")
    cam = scenes.camera(dims)
    g = R.Game(args=[*cam["origin"], cam["heading"], cam["pitch"], cam["sun_angle"]], world=R.ChunkStorage.from_window(m, f))
    p = R.create_instance(g, width=W, height=H, window=dims, flags=N.FLAG_REPROJECT | N.FLAG_DENOISE | N.FLAG_FINALIZE, blue_noise=noise)
    try:
        for i in range(4):
            p.draw_frame(g)
            cnt = p.renderer.read_buffer(N.BUF_HISTORY)[..., 3]
            hit = p.renderer.read_buffer(N.BUF_NORMAL) < 16
            assert hit.any() and (cnt[hit] == i + 1).all()
        g.borrow_controls_mut().on_pressed("forward"); g.tick(0.02)
        p.draw_frame(g)
        cnt = p.renderer.read_buffer(N.BUF_HISTORY)[..., 3]
        assert 0.2 < (cnt[hit] == 5).mean() and (cnt <= 5).all()
    finally:
        p.close()
