"""GPU parity: libvxrt (through the C ABI) against the CPU oracle on the same seeded inputs.
Integer outputs (hit voxel, face, material word, step and ray counts, G-buffer codes) are compared
bit for bit; hit distance within 1e-4; colour within 1/255 (8-bit) or 1/255 of the 16-bit range."""
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

T_TOL = 1e-4          # north_star: max abs error on hit t
C8_TOL = 1            # north_star: <= 1/255 per 8-bit channel
C16_TOL = 257         # the same tolerance on a 16-bit UNORM channel


def _frame_block(orc, dims, k=0, n=1, seed=1, lr=(0, 0, 0)):
    from raytrace_b200 import scenes
    cam = scenes.camera(dims, k, n)
    return orc.fill_frame(cam["origin"], cam["heading"], cam["pitch"], cam["sun_angle"], seed, lr)


def _compare(orc_frame, R, N, *, denoise, finalize, check_steps=True):
    B = lambda i: R.read_buffer(i)
    assert np.array_equal(B(N.BUF_HIT_VOXEL), orc_frame.hit_voxel), "hit voxel"
    assert np.array_equal(B(N.BUF_HIT_FACE), orc_frame.hit_face), "hit face"
    assert np.array_equal(B(N.BUF_HIT_WORD), orc_frame.hit_word), "material word"
    if check_steps:
        assert np.array_equal(B(N.BUF_STEPS), orc_frame.steps), "march iterations"
        assert np.array_equal(B(N.BUF_RAYS), orc_frame.rays), "rays per pixel"
    t_gpu, t_cpu = B(N.BUF_HIT_T), orc_frame.hit_t
    assert np.array_equal(np.isnan(t_gpu), np.isnan(t_cpu))
    ok = ~np.isnan(t_cpu)
    assert np.max(np.abs(t_gpu[ok] - t_cpu[ok]), initial=0.0) <= T_TOL
    assert np.array_equal(B(N.BUF_DEPTH), orc_frame.depth), "depth"
    assert np.array_equal(B(N.BUF_NORMAL), orc_frame.normal), "normal"
    assert np.array_equal(B(N.BUF_ALBEDO), orc_frame.albedo), "albedo"
    assert np.array_equal(B(N.BUF_EMISSION), orc_frame.emission), "emission"
    d = lambda a, b: int(np.max(np.abs(a.astype(np.int64) - b.astype(np.int64)), initial=0))
    assert d(B(N.BUF_FOG), orc_frame.fog) <= C8_TOL, "fog"
    assert d(B(N.BUF_LIGHTING_RAW), orc_frame.lighting_raw) <= C16_TOL, "lighting (trace)"
    assert d(B(N.BUF_LIGHTING), orc_frame.lighting) <= C16_TOL, "lighting (after denoise)"
    if finalize:
        assert d(B(N.BUF_OUTPUT), orc_frame.output) <= C8_TOL, "output image"


@pytest.mark.parametrize("scene,width,height,mode", [
    ("c1", 256, 256, 0),         # BASELINE config 1: 256x256 primary-ray frame, 128^3 heightmap
    ("c1", 256, 256, 2),
    ("small", 72, 40, 2),        # ragged frame (not a multiple of the 8x8 group or the 16x8 block)
    ("mid", 320, 200, 1),        # reference window size, shadow mode
    ("mid", 320, 200, 2),
])
def test_frame_matches_oracle(orc, world, noise, scene, width, height, mode):
    import raytrace_b200 as R
    from raytrace_b200 import _native as N, scenes
    dims = scenes.WORLDS[scene][0]
    m, f = world(scene)
    blk = _frame_block(orc, dims, seed=7)
    ref = orc.render(m, f, noise, blk, mode, width, height, denoise=True, finalize=True)
    with R.Renderer(width, height, dims, mode, N.FLAG_HIT_BUFFERS | N.FLAG_DENOISE | N.FLAG_FINALIZE) as r:
        r.upload_blue_noise(noise)
        r.upload_world(m, f)
        r.set_frame(blk)
        r.render_frame()
        _compare(ref, r, N, denoise=True, finalize=True)
        s = r.frame_stats()
        assert (s.rays, s.steps, s.hits, s.noise_fetches) == (ref.stats.rays, ref.stats.steps, ref.stats.hits, ref.stats.noise_fetches)
        assert s.limit_exhausted == ref.stats.ub_limit_exhausted


def test_camera_path_and_render_offset(orc, world, noise):
    """A few frames of the config-2 camera path + a non-zero window centre (toroidal addressing)."""
    import raytrace_b200 as R
    from raytrace_b200 import _native as N, scenes
    dims = scenes.WORLDS["mid"][0]
    m, f = world("mid")
    with R.Renderer(192, 108, dims, N.MODE_FULL, N.FLAG_HIT_BUFFERS | N.FLAG_DENOISE | N.FLAG_FINALIZE) as r:
        r.upload_blue_noise(noise)
        r.upload_world(m, f)
        for k, lr in [(0, (0, 0, 0)), (3, (0, 0, 0)), (7, (16, 0, -16))]:
            blk = _frame_block(orc, dims, k, 8, seed=100 + k, lr=lr)
            ref = orc.render(m, f, noise, blk, N.MODE_FULL, 192, 108, denoise=True, finalize=True)
            r.set_frame(blk)
            r.render_frame()
            _compare(ref, r, N, denoise=True, finalize=True)


@pytest.mark.parametrize("general", [False, True], ids=["fast", "general"])
@pytest.mark.parametrize("case", ["underground", "few_steps", "above_window", "lr_far", "lr_neg", "axis_aligned"])
def test_edge_rays_both_kernels(orc, world, noise, case, general):
    """The reference's undefined corners as DESIGN.md D2 defines them, on both march kernels (trace_fast.cu and
    the general form in trace.cu): camera inside a solid (E4), loop-limit exhaustion, start slid to the -Y
    face (:312-315), window centres away from 0 (toroidal addressing + (E8) thresholds), axis-aligned rays
    (a direction component of exactly 0: lpa = inf, (E9))."""
    import math
    import raytrace_b200 as R
    from raytrace_b200 import _native as N, scenes
    dims = scenes.WORLDS["mid"][0]
    m, f = world("mid")
    cam = scenes.camera(dims)
    origin, heading, pitch, lr, max_steps = cam["origin"], cam["heading"], cam["pitch"], (0, 0, 0), 2048
    if case == "underground":
        origin = (0.5, 0.5, -100.25)
    elif case == "few_steps":
        max_steps = 6
    elif case == "above_window":
        origin = (3.3, -400.0, 40.7)
    elif case == "lr_far":
        lr = (1024, -2048, 512); origin = (origin[0] + 1024, origin[1] - 2048, origin[2] + 512)
    elif case == "lr_neg":
        lr = (-48, 16, -32); origin = (origin[0] - 48, origin[1] + 16, origin[2] - 32)
    elif case == "axis_aligned":
        heading, pitch = math.pi / 2, 0.0
    W, H = 136, 72
    blk = orc.fill_frame(origin, heading, pitch, cam["sun_angle"], 11, lr)
    ref = orc.render(m, f, noise, blk, 2, W, H, denoise=True, finalize=True, max_steps=max_steps)
    flags = N.FLAG_HIT_BUFFERS | N.FLAG_DENOISE | N.FLAG_FINALIZE | (N.FLAG_GENERAL_MARCH if general else 0)
    with R.Renderer(W, H, dims, 2, flags, max_steps=max_steps) as r:
        r.upload_blue_noise(noise); r.upload_world(m, f); r.set_frame(blk); r.render_frame()
        _compare(ref, r, N, denoise=True, finalize=True)
        s = r.frame_stats()
        assert (s.rays, s.steps, s.hits) == (ref.stats.rays, ref.stats.steps, ref.stats.hits)
        assert s.limit_exhausted == ref.stats.ub_limit_exhausted
        if case == "few_steps":
            assert s.limit_exhausted > 0
        if case == "underground":
            assert ref.stats.ub_start_in_solid > 0


def test_general_kernel_chosen_far_from_origin(orc, world, noise):
    """Coordinates beyond 2^21 violate (E5)'s precondition: vxrt_render_frame must fall back to the general
    kernel by itself, and the result still matches the oracle."""
    import raytrace_b200 as R
    from raytrace_b200 import _native as N, scenes
    dims = scenes.WORLDS["small"][0]
    m, f = world("small")
    cam = scenes.camera(dims)
    off = (1 << 22)
    lr = (off, 0, 0)
    origin = (cam["origin"][0] + off, cam["origin"][1], cam["origin"][2])
    blk = orc.fill_frame(origin, cam["heading"], cam["pitch"], cam["sun_angle"], 5, lr)
    ref = orc.render(m, f, noise, blk, 2, 64, 48, denoise=True, finalize=True)
    with R.Renderer(64, 48, dims, 2, N.FLAG_HIT_BUFFERS | N.FLAG_DENOISE | N.FLAG_FINALIZE) as r:
        r.upload_blue_noise(noise); r.upload_world(m, f); r.set_frame(blk); r.render_frame()
        _compare(ref, r, N, denoise=True, finalize=True)


def test_async_readback_matches_sync(orc, world, noise):
    """vxrt_read_output_async / vxrt_wait_output: three frames in flight through the two device images land in
    page-locked host buffers identical to the synchronous read of the same frames."""
    import raytrace_b200 as R
    from raytrace_b200 import _native as N, scenes
    dims = scenes.WORLDS["c1"][0]
    m, f = world("c1")
    W, H = 160, 96
    blks = [_frame_block(orc, dims, k, 4, seed=20 + k) for k in range(4)]
    with R.Renderer(W, H, dims, 1, N.FLAG_DENOISE | N.FLAG_FINALIZE) as r:
        r.upload_blue_noise(noise); r.upload_world(m, f)
        want = []
        for b in blks:
            r.set_frame(b); r.render_frame(); want.append(r.read_output().copy())
        got = [np.zeros((H, W, 4), np.uint8) for _ in blks]
        for g in got:
            r.host_register(g)
        tickets = []
        for b, g in zip(blks, got):
            r.set_frame(b); r.render_frame(); tickets.append(r.read_output_async(g))
            if len(tickets) >= 2:
                r.wait_output(tickets[-2])
        r.wait_output(tickets[-1])
        for g, w in zip(got, want):
            assert np.array_equal(g, w)
        for g in got:
            r.host_unregister(g)
        with pytest.raises(R.VxrtError):
            r.wait_output(tickets[-1] + 5)


def test_pong_swap_flag(orc, world, noise):
    """descriptor_sets.rs:37-39: odd denoise passes read swapped depth/normal images."""
    import raytrace_b200 as R
    from raytrace_b200 import _native as N, scenes
    dims = scenes.WORLDS["c1"][0]
    m, f = world("c1")
    blk = _frame_block(orc, dims, seed=3)
    ref = orc.render(m, f, noise, blk, 2, 160, 96, denoise=True, finalize=True, pong_swap=True)
    with R.Renderer(160, 96, dims, 2, N.FLAG_HIT_BUFFERS | N.FLAG_DENOISE | N.FLAG_FINALIZE | N.FLAG_PONG_SWAP) as r:
        r.upload_blue_noise(noise); r.upload_world(m, f); r.set_frame(blk); r.render_frame()
        _compare(ref, r, N, denoise=True, finalize=True)


def test_accumulation(orc, world, noise):
    """Temporal accumulation (extension): running mean of `light` over seeds 1..6."""
    import raytrace_b200 as R
    from raytrace_b200 import _native as N, scenes
    dims = scenes.WORLDS["c1"][0]
    m, f = world("c1")
    W, H = 128, 96
    fr = orc.Frame(W, H, accumulate=True)
    with R.Renderer(W, H, dims, 2, N.FLAG_HIT_BUFFERS | N.FLAG_DENOISE | N.FLAG_FINALIZE | N.FLAG_ACCUMULATE) as r:
        r.upload_blue_noise(noise); r.upload_world(m, f)
        for seed in range(1, 7):
            blk = _frame_block(orc, dims, seed=seed)
            ref = orc.render(m, f, noise, blk, 2, W, H, denoise=True, finalize=True, frame=fr)
            r.set_frame(blk); r.render_frame()
        _compare(ref, r, N, denoise=True, finalize=True)
        acc = r.read_buffer(N.BUF_ACCUM)
        assert np.array_equal(acc[..., 3], fr.accum[..., 3])
        assert np.allclose(acc[..., :3], fr.accum[..., :3], rtol=1e-5, atol=1e-5)
        r.reset_accumulation()
        r.render_frame()
        assert np.all(r.read_buffer(N.BUF_ACCUM)[..., 3] == 1.0)


@pytest.mark.parametrize("scene", ["small", "c1", "mid"])
def test_generator_and_minefield_match_oracle(orc, world, scene):
    """k_generate == orc_generate_world and k_build_minefield == pack_into (chunk.rs:125-184), bit exact."""
    import raytrace_b200 as R
    from raytrace_b200 import _native as N, scenes
    dims, kind, seed = scenes.WORLDS[scene]
    m, f = world(scene)
    with R.Renderer(64, 64, dims, 0, 0) as r:
        r.generate_world(kind, seed)
        gm, gf = r.read_world()
        assert np.array_equal(gm, m)
        assert np.array_equal(gf, f)
        # chunk-wise upload without minefield -> built on the device
        st = R.ChunkStorage.from_window(m)
        for (cx, cy, cz), data in st.chunks.items():
            r.upload_chunk(cx + dims[0] // 128, cy + dims[1] // 128, cz + dims[2] // 128, data.materials, None)
        gm, gf = r.read_world()
        assert np.array_equal(gm, m) and np.array_equal(gf, f)


def test_pack_into_edge_chunks(orc):
    """Known-answer chunks of chunk.rs:125-184 through the device builder."""
    import raytrace_b200 as R
    solid = 0x1FC07F
    cases = {"empty": np.zeros((64, 64, 64), np.uint32), "full": np.full((64, 64, 64), solid, np.uint32)}
    one = np.zeros((64, 64, 64), np.uint32); one[0, 0, 0] = solid; cases["corner"] = one
    nonsolid = np.full((64, 64, 64), 0x7F, np.uint32); cases["non_solid_words"] = nonsolid   # bit 15 clear -> chunk wiped
    rng = np.random.default_rng(5)
    sparse = np.where(rng.random((64, 64, 64)) < 0.001, solid, 0).astype(np.uint32); cases["sparse"] = sparse
    with R.Renderer(64, 64, (64, 64, 64), 0, 0) as r:
        for name, words in cases.items():
            em, ef = orc.pack_chunk(words)
            r.upload_chunk(0, 0, 0, words, None)
            gm, gf = r.read_world()
            assert np.array_equal(gm, em), name
            assert np.array_equal(gf, ef), name


def test_edits_rebuild_matches_oracle(orc, world, noise):
    """BASELINE config 5 in small: seeded unique edits -> touched chunks rebuilt -> re-render."""
    import raytrace_b200 as R
    from raytrace_b200 import _native as N, scenes
    dims = scenes.WORLDS["mid"][0]
    m, f = world("mid")
    m = m.copy()
    with R.Renderer(160, 90, dims, 2, N.FLAG_HIT_BUFFERS | N.FLAG_DENOISE | N.FLAG_FINALIZE) as r:
        r.upload_blue_noise(noise); r.upload_world(m, f)
        for frame in range(3):
            e = orc.generate_edits(2000, *dims, 9, frame)
            assert len(np.unique(e[:, :3], axis=0)) == len(e)
            orc.apply_edits(m, e)
            r.apply_edits(e)
        em, ef = orc.build_minefield(m)
        gm, gf = r.read_world()
        assert np.array_equal(gm, em) and np.array_equal(gf, ef)
        blk = _frame_block(orc, dims, seed=11)
        ref = orc.render(em, ef, noise, blk, 2, 160, 90, denoise=True, finalize=True)
        r.set_frame(blk); r.render_frame()
        _compare(ref, r, N, denoise=True, finalize=True)


def test_slice_upload(orc, world):
    """vxrt_upload_slice == writing a 16-thick slab into the window (terrain_upload.rs:212-267)."""
    import raytrace_b200 as R
    from raytrace_b200 import scenes
    dims = scenes.WORLDS["c1"][0]
    m, f = world("c1")
    rng = np.random.default_rng(1)
    with R.Renderer(64, 64, dims, 0, 0) as r:
        r.upload_world(m, f)
        em, ef = m.copy(), f.copy()
        for axis, off in [(0, 16), (1, 112), (2, 48)]:
            shape = [dims[2], dims[1], dims[0]]; shape[2 - axis] = 16
            sm = rng.integers(0, 2 ** 21, size=shape, dtype=np.uint32); sf = rng.integers(0, 7, size=shape, dtype=np.uint8)
            r.upload_slice(axis, off, sm, sf)
            sl = [slice(None)] * 3; sl[2 - axis] = slice(off, off + 16)
            em[tuple(sl)] = sm; ef[tuple(sl)] = sf
        gm, gf = r.read_world()
        assert np.array_equal(gm, em) and np.array_equal(gf, ef)


def test_row_band_matches_full_frame(orc, world, noise):
    """Image-space sharding: a handle that renders rows [r0,r1) reproduces those rows of the full frame."""
    import raytrace_b200 as R
    from raytrace_b200 import _native as N, scenes
    dims = scenes.WORLDS["c1"][0]
    m, f = world("c1")
    W, H = 160, 400
    blk = _frame_block(orc, dims, seed=5)
    flags = N.FLAG_DENOISE | N.FLAG_FINALIZE
    with R.Renderer(W, H, dims, 2, flags) as full:
        full.upload_blue_noise(noise); full.upload_world(m, f); full.set_frame(blk); full.render_frame()
        img = full.read_output()
    for r0, r1 in [(0, 136), (136, 264), (264, 400)]:
        with R.Renderer(W, H, dims, 2, flags, rows=(r0, r1)) as band:
            band.upload_blue_noise(noise); band.upload_world(m, f); band.set_frame(blk); band.render_frame()
            out = band.read_output()
            assert np.array_equal(out[H - r1:H - r0], img[H - r1:H - r0])


def test_pipeline_interface(orc, world, noise):
    """The reference-shaped call sequence: Game -> create_instance -> draw_frame (src/bin/main.rs:8-54)."""
    import raytrace_b200 as R
    from raytrace_b200 import _native as N, scenes
    dims = scenes.WORLDS["c1"][0]
    m, f = world("c1")
    cam = scenes.camera(dims)
    game = R.Game(args=[*cam["origin"], cam["heading"], cam["pitch"], cam["sun_angle"]], world=R.ChunkStorage.from_window(m, f))
    pipe = R.create_instance(game, width=128, height=128, window=dims, blue_noise=noise)
    img1 = pipe.draw_frame(game).copy()
    assert pipe.seed == 1 and img1.shape == (128, 128, 4) and img1[..., 3].min() == 255
    blk = np.frombuffer(bytes(pipe.uniform_data), np.uint8)
    ref = orc.render(m, f, noise, blk, 2, 128, 128, denoise=True, finalize=True)
    assert np.max(np.abs(img1.astype(int) - ref.output.astype(int))) <= C8_TOL
    game.borrow_controls_mut().on_pressed("forward"); game.tick(0.1)
    img2 = pipe.draw_frame(game)
    assert pipe.seed == 2 and not np.array_equal(img1, img2)
    pipe.close()


def _golden():
    import glob, os
    return sorted(glob.glob(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "*.npz")))


@pytest.mark.parametrize("path", _golden(), ids=lambda p: p.split("/")[-1])
def test_cuda_matches_reference_golden(orc, world, noise, path):
    """libvxrt against fixtures written by the reference's own shader text (tests/golden/make_golden.py):
    integer images bit-exact, colour images within the stated tolerance."""
    import raytrace_b200 as R
    from raytrace_b200 import _native as N, scenes
    g = np.load(path)
    dims = scenes.WORLDS["mid"][0]
    m, f = world("mid")
    H, W = g["depth"].shape
    flags = N.FLAG_DENOISE | N.FLAG_FINALIZE | N.FLAG_HIT_BUFFERS | (N.FLAG_PONG_SWAP if bool(g["pong_swap"]) else 0)
    with R.Renderer(W, H, dims, N.MODE_FULL, flags) as r:
        r.upload_blue_noise(noise); r.upload_world(m, f); r.set_frame(g["frame_block"]); r.render_frame()
        for buf, name in [(N.BUF_DEPTH, "depth"), (N.BUF_NORMAL, "normal"), (N.BUF_ALBEDO, "albedo"), (N.BUF_EMISSION, "emission")]:
            assert np.array_equal(r.read_buffer(buf), g[name]), name
        d = lambda a, b: int(np.max(np.abs(a.astype(np.int64) - b.astype(np.int64))))
        assert d(r.read_buffer(N.BUF_FOG), g["fog"]) <= C8_TOL
        assert d(r.read_buffer(N.BUF_LIGHTING_RAW), g["lighting_raw"]) <= C16_TOL
        assert d(r.read_buffer(N.BUF_LIGHTING), g["lighting"]) <= C16_TOL
        assert d(r.read_output(), g["output"]) <= C8_TOL


def test_cpp_host_driver(orc, world, noise, tmp_path):
    """The C++ host side (csrc/host/: Game -> create_instance -> draw_frame, a headless src/bin/main.rs)
    uploads the world chunk by chunk with no minefield (built on the device), renders, and its frame is
    checked against the oracle fed with the uniform block the driver itself filled."""
    import os, subprocess
    from raytrace_b200 import scenes
    exe = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "raytrace_b200", "vxrt_main")
    assert os.path.exists(exe), "vxrt_main not built (raytrace_b200/build.py)"
    dims = scenes.WORLDS["c1"][0]
    m, f = world("c1")
    cam = scenes.camera(dims)
    (tmp_path / "w.u32").write_bytes(np.ascontiguousarray(m).tobytes())
    (tmp_path / "n.rgba").write_bytes(np.ascontiguousarray(noise).tobytes())
    W, H = 192, 136
    cmd = [exe, "--world", str(tmp_path / "w.u32"), "--noise", str(tmp_path / "n.rgba"), "--dims", *map(str, dims),
           "--size", str(W), str(H), "--frames", "3", "--hold", "forward", "--dt", "0.05", "--mode", "2",
           "--camera", *[repr(float(v)) for v in (*cam["origin"], cam["heading"], cam["pitch"], cam["sun_angle"])],
           "--out", str(tmp_path / "o.ppm"), "--block", str(tmp_path / "b.bin")]
    p = subprocess.run(cmd, capture_output=True, text=True, timeout=120)
    assert p.returncode == 0, p.stderr
    assert "Creating renderer" in p.stdout and "ms /" in p.stdout
    blk = np.frombuffer((tmp_path / "b.bin").read_bytes(), np.uint8)
    assert blk.size == 192 and int(blk[4:8].view(np.uint32)[0]) == 3          # seed pre-incremented per frame
    raw = (tmp_path / "o.ppm").read_bytes()
    hdr = f"P6\n{W} {H}\n255\n".encode()
    assert raw.startswith(hdr)
    img = np.frombuffer(raw[len(hdr):], np.uint8).reshape(H, W, 3)
    ref = orc.render(m, f, noise, blk, 2, W, H, denoise=True, finalize=True)
    assert np.max(np.abs(img.astype(int) - ref.output[..., :3].astype(int))) <= C8_TOL


@pytest.mark.parametrize("bands", [[(0, 136), (136, 264), (264, 400)], [(0, 56), (56, 112), (112, 200), (200, 400)]],
                         ids=["3_bands", "4_bands_thinner_than_halo"])
def test_peer_memory_frame_matches_single_handle(orc, world, noise, bands):
    """The multi-GPU frame of peer.cu (own-rows trace, halo rows pushed into the neighbours' images, finalize
    storing straight into the root's output, flag-ordered) with every band as a handle of THIS process on one GPU:
    the root's image must equal the single-handle frame bit for bit, over several frames (both output parities,
    back-pressure flags) and through the pipelined read-back."""
    import raytrace_b200 as R
    from raytrace_b200 import _native as N, scenes
    dims = scenes.WORLDS["c1"][0]
    m, f = world("c1")
    W, H = 160, 400
    flags = N.FLAG_DENOISE | N.FLAG_FINALIZE
    blocks = [_frame_block(orc, dims, seed=s) for s in (5, 6, 7, 8)]
    want = []
    with R.Renderer(W, H, dims, 2, flags) as full:
        full.upload_blue_noise(noise); full.upload_world(m, f)
        for b in blocks:
            full.set_frame(b); full.render_frame(); want.append(full.read_output().copy())
    rs = [R.Renderer(W, H, dims, 2, flags, rows=b) for b in bands]
    try:
        for r in rs:
            r.upload_blue_noise(noise); r.upload_world(m, f)
        blobs = [r.peer_export() for r in rs]
        root = 1
        for k, r in enumerate(rs):
            r.peer_attach(blobs, k, root=root)
        host = [np.zeros((H, W, 4), np.uint8) for _ in range(2)]
        for b in host:
            rs[root].host_register(b)
        tickets = []
        for i, b in enumerate(blocks):
            for r in rs:                                   # enqueue only: no rank may block the host before all have queued
                r.set_frame(b); r.render_frame()
            if i < 2:
                got = rs[root].read_output()
                bad = np.nonzero((got != want[i]).any(axis=(1, 2)))[0]
                assert bad.size == 0, f"frame {i}: {bad.size} output rows differ, first {bad[:8]} last {bad[-8:]} (G-buffer row = {H - 1} - output row)"
            else:                                          # pipelined read-back of frames 2 and 3
                tickets.append(rs[root].read_output_async(host[i & 1]))
        for t in tickets:
            rs[root].wait_output(t)
        assert np.array_equal(host[0], want[2]) and np.array_equal(host[1], want[3])
        for b in host:
            rs[root].host_unregister(b)
        for r in rs:
            r.sync(); r.frame_stats()                      # raises VXRT_ERR_COMM if a flag wait timed out
        rays = sum(r.frame_stats().rays for r in rs)
        with R.Renderer(W, H, dims, 2, flags) as full:
            full.upload_blue_noise(noise); full.upload_world(m, f); full.set_frame(blocks[-1]); full.render_frame()
            assert rays == full.frame_stats().rays         # no band traced a halo row
    finally:
        for r in rs:
            r.close()


def test_row_costs_and_balanced_bands(orc, world, noise):
    """vxrt_row_costs against numpy on the hit buffers; the weighted plan gives the sky-heavy part of the frame the
    taller band, and a peer-memory frame over those bands still equals the single-handle frame."""
    import raytrace_b200 as R
    from raytrace_b200 import _native as N, scenes
    dims = scenes.WORLDS["c1"][0]
    m, f = world("c1")
    W, H = 160, 400
    flags = N.FLAG_DENOISE | N.FLAG_FINALIZE
    blk = _frame_block(orc, dims, seed=5)
    with R.Renderer(W, H, dims, 2, flags | N.FLAG_HIT_BUFFERS) as full:
        full.upload_blue_noise(noise); full.upload_world(m, f); full.set_frame(blk); full.render_frame()
        want = full.read_output().copy()
        t, d = full.row_costs()
        steps = full.read_buffer(N.BUF_STEPS).astype(np.float64).sum(1)
        rays = full.read_buffer(N.BUF_RAYS).astype(np.float64).sum(1)
        filt = (full.read_buffer(N.BUF_NORMAL) < 16).sum(1)
    assert np.allclose(t, steps + 7 * rays + 16.7 * W, rtol=1e-5) and np.allclose(d, 15 * filt + 0.45 * W, rtol=1e-5)
    bands = R.plan_bands_weighted(H, 3, t, d, flags)
    cost = [t[a:b].sum() + 6 * d[a:b].sum() for a, b in bands]
    eq = [t[a:b].sum() + 6 * d[a:b].sum() for a, b in R.plan_bands(H, 3)]
    assert max(cost) < max(eq)                                 # better balanced than equal-height bands
    rs = [R.Renderer(W, H, dims, 2, flags, rows=b) for b in bands]
    try:
        for r in rs:
            r.upload_blue_noise(noise); r.upload_world(m, f)
        blobs = [r.peer_export() for r in rs]
        for k, r in enumerate(rs):
            r.peer_attach(blobs, k, root=0)
        for r in rs:
            r.set_frame(blk); r.render_frame()
        assert np.array_equal(rs[0].read_output(), want)
    finally:
        for r in rs:
            r.close()


def test_refined_division_is_ieee_division():
    """trace_fast.cu (E13): the shared-reciprocal quotient sequence equals __fdiv_rn bit for bit on 2^27 operand
    sets inside the guarded range (three normalize quotients, one free quotient and three reciprocals each)."""
    import ctypes as C
    from raytrace_b200 import _native as N
    lib = N.load()
    for seed in (1, 0x9E3779B97F4A7C15):
        bad = C.c_uint64(123)
        assert lib.vxrt_selftest_division(0, 1 << 27, seed, C.byref(bad)) == N.OK
        assert bad.value == 0


def test_shared_world_handles_and_frames_in_flight(orc, world, noise):
    """vxrt_create_shared: a borrower renders from the owner's world (no second copy) on its own stream; frames
    submitted alternately (FramePipeline) equal the frames of a standalone handle bit for bit; world writes go through
    the owner and reach the borrower's next frame; a borrower rejects them."""
    import raytrace_b200 as R
    from raytrace_b200 import _native as N, scenes
    dims = scenes.WORLDS["mid"][0]
    m, f = world("mid")
    W, H = 200, 120
    flags = N.FLAG_HIT_BUFFERS | N.FLAG_DENOISE | N.FLAG_FINALIZE
    blocks = [_frame_block(orc, dims, k, 6, seed=20 + k) for k in range(6)]
    with R.Renderer(W, H, dims, N.MODE_FULL, flags) as solo:
        solo.upload_blue_noise(noise); solo.upload_world(m, f)
        want = []
        for b in blocks:
            solo.set_frame(b); solo.render_frame()
            want.append((solo.read_output().copy(), solo.read_buffer(N.BUF_HIT_VOXEL).copy(), solo.frame_stats().steps))
        edits = orc.generate_edits(2000, *dims, 7, 0)
        solo.apply_edits(edits)
        solo.set_frame(blocks[0]); solo.render_frame()
        want_edit = solo.read_output().copy()
    with R.FramePipeline(W, H, dims, N.MODE_FULL, flags, depth=3) as fp:
        fp.owner.upload_blue_noise(noise); fp.owner.upload_world(m, f)
        used = [fp.submit(b) for b in blocks[:3]]                      # three frames in flight, one per handle
        for r, (img, vox, steps) in zip(used, want[:3]):
            assert np.array_equal(r.read_output(), img) and np.array_equal(r.read_buffer(N.BUF_HIT_VOXEL), vox)
            assert r.frame_stats().steps == steps
        used = [fp.submit(b) for b in blocks[3:]]
        for r, (img, vox, steps) in zip(used, want[3:]):
            assert np.array_equal(r.read_output(), img)
        borrower = fp.renderers[1]
        with pytest.raises(R.VxrtError) as e:
            borrower.apply_edits(edits)
        assert e.value.code == N.ERR_STATE
        with pytest.raises(R.VxrtError):
            borrower.generate_world(0, 1)
        fp.owner.apply_edits(edits)                                   # ordered against the borrowers' frames
        borrower.set_frame(blocks[0]); borrower.render_frame()
        assert np.array_equal(borrower.read_output(), want_edit)
    # a handle on another window cannot borrow
    with R.Renderer(64, 64, (64, 64, 64)) as a:
        with pytest.raises(R.VxrtError):
            R.Renderer(64, 64, (128, 64, 64), share_world=a)


def test_shared_accumulation_chain(orc, world, noise):
    """VXRT_FLAG_ACCUMULATE on handles that share a world: ONE running sum (the owner's).  Frames dealt alternately
    to owner and borrower accumulate exactly as on a single handle - each trace pass waits for the previous one, the
    passes after it overlap the next frame's trace pass - and every frame's image equals the single-handle frame."""
    import raytrace_b200 as R
    from raytrace_b200 import _native as N, scenes
    dims = scenes.WORLDS["c1"][0]
    m, f = world("c1")
    W, H = 160, 96
    flags = N.FLAG_DENOISE | N.FLAG_FINALIZE | N.FLAG_ACCUMULATE
    blocks = [_frame_block(orc, dims, seed=s) for s in range(1, 8)]
    with R.Renderer(W, H, dims, N.MODE_FULL, flags) as solo:
        solo.upload_blue_noise(noise); solo.upload_world(m, f)
        want = []
        for b in blocks:
            solo.set_frame(b); solo.render_frame()
            want.append(solo.read_output().copy())
        want_acc = solo.read_buffer(N.BUF_ACCUM).copy()
        solo.reset_accumulation(); solo.set_frame(blocks[2]); solo.render_frame()
        want_reset = solo.read_output().copy()
    with R.FramePipeline(W, H, dims, N.MODE_FULL, flags, depth=2) as fp:
        fp.owner.upload_blue_noise(noise); fp.owner.upload_world(m, f)
        used = [fp.submit(b) for b in blocks]                      # nothing synchronised in between
        got = {}
        for k in (5, 6):                                               # the last frame of each handle is still in its image
            got[k] = used[k].read_output()
        assert np.array_equal(got[5], want[5]) and np.array_equal(got[6], want[6])
        for r in fp.renderers:                                         # either handle reads the one running sum
            assert np.array_equal(r.read_buffer(N.BUF_ACCUM), want_acc)
        fp.renderers[1].reset_accumulation()                           # through the borrower: the chain's sum
        r = fp.submit(blocks[2])
        assert np.array_equal(r.read_output(), want_reset)
        assert np.all(fp.owner.read_buffer(N.BUF_ACCUM)[..., 3] == 1.0)
    with R.Renderer(W, H, dims, N.MODE_FULL, N.FLAG_DENOISE | N.FLAG_FINALIZE) as plain:   # owner without a running sum
        with pytest.raises(R.VxrtError):
            R.Renderer(W, H, dims, N.MODE_FULL, flags, share_world=plain)


@pytest.mark.parametrize("scene,width,height", [("c1", 256, 256), ("mid", 320, 200), ("small", 72, 40)])
def test_second_bounce_handoff_is_invisible(orc, world, noise, scene, width, height):
    """trace_fast.cu (E16), VXRT_FLAG_RAY_HANDOFF: a FULL-lighting frame gives each second-bounce diffuse ray to a finished
    lane of its warp.  Every ray and every operation of a pixel stays the same, so the images are those of the
    sequential form (a handle with hit buffers) bit for bit, the frame counters too, and both match the oracle."""
    import raytrace_b200 as R
    from raytrace_b200 import _native as N, scenes
    dims = scenes.WORLDS[scene][0]
    m, f = world(scene)
    frames = [_frame_block(orc, dims, k, 3, seed=40 + k) for k in range(3)]
    with R.Renderer(width, height, dims, N.MODE_FULL, N.FLAG_HIT_BUFFERS | N.FLAG_ACCUMULATE) as seq, \
         R.Renderer(width, height, dims, N.MODE_FULL, N.FLAG_ACCUMULATE | N.FLAG_RAY_HANDOFF) as hand:
        for r in (seq, hand):
            r.upload_blue_noise(noise); r.upload_world(m, f)
        fr = orc.Frame(width, height, accumulate=True)
        for blk in frames:
            ref = orc.render(m, f, noise, blk, N.MODE_FULL, width, height, frame=fr)
            for r in (seq, hand):
                r.set_frame(blk); r.render_frame()
            for b in (N.BUF_LIGHTING, N.BUF_DEPTH, N.BUF_NORMAL, N.BUF_ALBEDO, N.BUF_FOG, N.BUF_ACCUM):
                assert np.array_equal(seq.read_buffer(b), hand.read_buffer(b)), b
            a, b = seq.frame_stats(), hand.frame_stats()
            assert (a.rays, a.steps, a.hits, a.noise_fetches, a.limit_exhausted) == (b.rays, b.steps, b.hits, b.noise_fetches, b.limit_exhausted)
            assert (b.rays, b.steps, b.hits) == (ref.stats.rays, ref.stats.steps, ref.stats.hits)
            assert np.array_equal(hand.read_buffer(N.BUF_DEPTH), ref.depth) and np.array_equal(hand.read_buffer(N.BUF_NORMAL), ref.normal)
            d = np.abs(hand.read_buffer(N.BUF_LIGHTING).astype(np.int64) - ref.lighting.astype(np.int64)).max()
            assert d <= C16_TOL, d


def test_duplicate_edits_last_entry_wins(orc, world):
    """An edit list that names voxels several times: set_block applied in list order (chunk.rs:115-117) leaves the LAST word;
    k_apply_edits resolves it the same way on every run (one thread per edit used to race)."""
    import raytrace_b200 as R
    from raytrace_b200 import scenes
    dims = scenes.WORLDS["c1"][0]
    m, f = world("c1")
    rng = np.random.default_rng(3)
    base = orc.generate_edits(3000, *dims, 9, 0)
    pick = rng.integers(0, 200, size=9000)                         # 200 voxels named ~45 times each, interleaved
    e = base[pick].copy()
    e[:, 3] = rng.integers(0, 2, size=len(e)) * 0xCD9B3
    e = np.concatenate([base, e, base[:50]]).astype(np.int32)
    want = m.copy(); orc.apply_edits(want, e)                      # sequential, list order
    want, wf = orc.build_minefield(want)
    for _ in range(3):
        with R.Renderer(64, 64, dims, 0, 0) as r:
            r.upload_world(m, f); r.apply_edits(e)
            gm, gf = r.read_world()
            assert np.array_equal(gm, want) and np.array_equal(gf, wf)


@pytest.mark.parametrize("scene,width,height,lr", [("c1", 256, 256, (0, 0, 0)), ("mid", 320, 200, (0, 0, 0)), ("small", 72, 40, (0, 0, 0)),
                                                   ("mid", 200, 120, (16, 0, -16))])
def test_ray_pool_is_invisible(orc, world, noise, scene, width, height, lr):
    """trace_pool.cu (vxrt_set_trace_form POOL): a FULL-lighting handle without hit buffers runs k_trace_pool (rays decoupled from lanes, refilled from a
    per-warp queue).  Every ray and every fp32 operation of a pixel is that of the sequential form (k_trace_fast, which a handle
    WITH hit buffers runs), so the images, the running sum and the frame counters are equal bit for bit, and both match the
    oracle."""
    import raytrace_b200 as R
    from raytrace_b200 import _native as N, scenes
    dims = scenes.WORLDS[scene][0]
    m, f = world(scene)
    frames = []
    for k in range(3):
        cam = scenes.camera(dims, k, 3)
        o = tuple(c + l for c, l in zip(cam["origin"], lr))
        frames.append(orc.fill_frame(o, cam["heading"], cam["pitch"], cam["sun_angle"], 40 + k, lr))
    with R.Renderer(width, height, dims, N.MODE_FULL, N.FLAG_HIT_BUFFERS | N.FLAG_ACCUMULATE) as seq, \
         R.Renderer(width, height, dims, N.MODE_FULL, N.FLAG_ACCUMULATE) as pool:
        pool.set_trace_form(N.TRACE_POOL)
        for r in (seq, pool):
            r.upload_blue_noise(noise); r.upload_world(m, f)
        fr = orc.Frame(width, height, accumulate=True)
        for blk in frames:
            ref = orc.render(m, f, noise, blk, N.MODE_FULL, width, height, frame=fr)
            for r in (seq, pool):
                r.set_frame(blk); r.render_frame()
            a, b = seq.frame_stats(), pool.frame_stats()
            assert (a.rays, a.steps, a.hits, a.noise_fetches, a.limit_exhausted) == (b.rays, b.steps, b.hits, b.noise_fetches, b.limit_exhausted)
            assert (b.rays, b.steps, b.hits) == (ref.stats.rays, ref.stats.steps, ref.stats.hits)
            for buf in (N.BUF_DEPTH, N.BUF_NORMAL, N.BUF_ALBEDO, N.BUF_EMISSION, N.BUF_FOG, N.BUF_LIGHTING, N.BUF_ACCUM):
                assert np.array_equal(seq.read_buffer(buf), pool.read_buffer(buf)), buf
            assert np.array_equal(pool.read_buffer(N.BUF_DEPTH), ref.depth) and np.array_equal(pool.read_buffer(N.BUF_NORMAL), ref.normal)
            d = np.abs(pool.read_buffer(N.BUF_LIGHTING).astype(np.int64) - ref.lighting.astype(np.int64)).max()
            assert d <= C16_TOL, d


@pytest.mark.parametrize("scene,width,height,lr", [("c1", 256, 256, (0, 0, 0)), ("mid", 320, 200, (0, 0, 0)), ("small", 72, 40, (0, 0, 0)),
                                                   ("mid", 200, 120, (16, 0, -16))])
def test_wavefront_is_invisible(orc, world, noise, scene, width, height, lr):
    """trace_wave.cu (vxrt_set_trace_form WAVEFRONT): the FULL-lighting tree cut at its rays - primary stage, then per level a
    ray-queue march with lane refill and a resolve pass.  Every ray and every fp32 operation of a pixel is that of the sequential
    form (which a handle WITH hit buffers runs), so the images, the running sum and the frame counters are equal bit for bit, and
    both match the oracle."""
    import raytrace_b200 as R
    from raytrace_b200 import _native as N, scenes
    dims = scenes.WORLDS[scene][0]
    m, f = world(scene)
    frames = []
    for k in range(3):
        cam = scenes.camera(dims, k, 3)
        o = tuple(c + l for c, l in zip(cam["origin"], lr))
        frames.append(orc.fill_frame(o, cam["heading"], cam["pitch"], cam["sun_angle"], 40 + k, lr))
    with R.Renderer(width, height, dims, N.MODE_FULL, N.FLAG_HIT_BUFFERS | N.FLAG_ACCUMULATE) as seq, \
         R.Renderer(width, height, dims, N.MODE_FULL, N.FLAG_ACCUMULATE) as wave:
        wave.set_trace_form(N.TRACE_WAVEFRONT)
        for r in (seq, wave):
            r.upload_blue_noise(noise); r.upload_world(m, f)
        fr = orc.Frame(width, height, accumulate=True)
        for blk in frames:
            ref = orc.render(m, f, noise, blk, N.MODE_FULL, width, height, frame=fr)
            for r in (seq, wave):
                r.set_frame(blk); r.render_frame()
            assert wave.launches_per_frame() == 5, "the wavefront form did not run"
            a, b = seq.frame_stats(), wave.frame_stats()
            assert (a.rays, a.steps, a.hits, a.noise_fetches, a.limit_exhausted) == (b.rays, b.steps, b.hits, b.noise_fetches, b.limit_exhausted)
            assert (b.rays, b.steps, b.hits) == (ref.stats.rays, ref.stats.steps, ref.stats.hits)
            for buf in (N.BUF_DEPTH, N.BUF_NORMAL, N.BUF_ALBEDO, N.BUF_EMISSION, N.BUF_FOG, N.BUF_LIGHTING, N.BUF_ACCUM):
                assert np.array_equal(seq.read_buffer(buf), wave.read_buffer(buf)), buf
            assert np.array_equal(wave.read_buffer(N.BUF_DEPTH), ref.depth) and np.array_equal(wave.read_buffer(N.BUF_NORMAL), ref.normal)
            d = np.abs(wave.read_buffer(N.BUF_LIGHTING).astype(np.int64) - ref.lighting.astype(np.int64)).max()
            assert d <= C16_TOL, d


@pytest.mark.parametrize("scene,width,height", [("mid", 192, 112), ("c1", 160, 96)])
def test_trace_forms_agree_on_random_cameras(orc, world, noise, scene, width, height):
    """Differential test of the four forms of the FULL-lighting trace pass (sequential with the parity buffers, sequential without,
    per-warp pool, wavefront, second-bounce hand-off) on seeded random cameras, including the awkward ones: inside a solid (E4),
    looking straight down and up, outside the window above the terrain, at dusk.  The G-buffer, the lighting image, the running sum
    and the frame counters must be equal bit for bit on every frame; the first frame of each scene is also held against the oracle."""
    import raytrace_b200 as R
    from raytrace_b200 import _native as N, scenes
    dims = scenes.WORLDS[scene][0]
    m, f = world(scene)
    rng = np.random.default_rng(20261105)
    nx, ny, nz = dims
    cams = []
    for k in range(14):
        o = (float(rng.uniform(-0.45, 0.45) * nx), float(rng.uniform(-0.45, 0.45) * ny), float(rng.uniform(-0.45, 0.6) * nz))
        cams.append(dict(origin=o, heading=float(rng.uniform(0, 6.283)), pitch=float(rng.uniform(-1.5, 1.0)), sun_angle=float(rng.uniform(0.0, 3.3))))
    cams.append(dict(origin=(0.3, 0.2, 0.49 * nz), heading=1.0, pitch=-1.5707, sun_angle=0.7))          # straight down from the top
    cams.append(dict(origin=(0.3, 0.2, -0.45 * nz), heading=2.0, pitch=1.5, sun_angle=1.2))             # from inside the ground, up
    forms = [("sequential", N.TRACE_SEQUENTIAL, 0), ("pool", N.TRACE_POOL, 0), ("wavefront", N.TRACE_WAVEFRONT, 0), ("handoff", N.TRACE_SEQUENTIAL, N.FLAG_RAY_HANDOFF)]
    with R.Renderer(width, height, dims, N.MODE_FULL, N.FLAG_HIT_BUFFERS | N.FLAG_ACCUMULATE) as ref_r:
        others = [R.Renderer(width, height, dims, N.MODE_FULL, N.FLAG_ACCUMULATE | fl) for _, _, fl in forms]
        try:
            for r, (_, form, _) in zip(others, forms):
                r.set_trace_form(form)
            for r in [ref_r] + others:
                r.upload_blue_noise(noise); r.upload_world(m, f)
            for k, cam in enumerate(cams):
                blk = orc.fill_frame(cam["origin"], cam["heading"], cam["pitch"], cam["sun_angle"], 100 + k)
                for r in [ref_r] + others:
                    r.set_frame(blk); r.render_frame()
                a = ref_r.frame_stats()
                want = {buf: ref_r.read_buffer(buf) for buf in (N.BUF_DEPTH, N.BUF_NORMAL, N.BUF_ALBEDO, N.BUF_EMISSION, N.BUF_FOG, N.BUF_LIGHTING, N.BUF_ACCUM)}
                for r, (name, _, _) in zip(others, forms):
                    b = r.frame_stats()
                    assert (a.rays, a.steps, a.hits, a.noise_fetches, a.limit_exhausted) == (b.rays, b.steps, b.hits, b.noise_fetches, b.limit_exhausted), (name, k)
                    for buf, w in want.items():
                        got = r.read_buffer(buf)
                        assert np.array_equal(got, w, equal_nan=True), (name, k, buf)
                if k == 0:
                    o = orc.render(m, f, noise, blk, N.MODE_FULL, width, height)
                    assert (a.rays, a.steps, a.hits) == (o.stats.rays, o.stats.steps, o.stats.hits)
                    assert np.array_equal(want[N.BUF_DEPTH], o.depth) and np.array_equal(want[N.BUF_NORMAL], o.normal)
        finally:
            for r in others:
                r.close()


def test_auto_handoff_on_long_rays_is_invisible(orc, world, noise):
    """AUTO form policy (vxrt_render_frame): once vxrt_frame_stats has shown more than 28 march steps per ray - a heavily edited
    world - a FULL-lighting handle hands second bounce rays to finished lanes.  Frames before and after the switch equal those of a
    handle with parity buffers (always the plain sequential form) bit for bit."""
    import raytrace_b200 as R
    from raytrace_b200 import _native as N, scenes
    dims = scenes.WORLDS["mid"][0]
    m, f = world("mid")
    width, height = 256, 144
    e = orc.generate_edits(60000, *dims, 5, 0)           # solid specks all over the window: short steps, long rays
    with R.Renderer(width, height, dims, N.MODE_FULL, N.FLAG_HIT_BUFFERS | N.FLAG_ACCUMULATE) as seq, \
         R.Renderer(width, height, dims, N.MODE_FULL, N.FLAG_ACCUMULATE) as auto:
        for r in (seq, auto):
            r.upload_blue_noise(noise); r.upload_world(m, f); r.apply_edits(e)
        ratios = []
        for k in range(4):
            cam = scenes.camera(dims, k, 4)
            blk = orc.fill_frame(cam["origin"], cam["heading"], cam["pitch"], cam["sun_angle"], 70 + k)
            for r in (seq, auto):
                r.set_frame(blk); r.render_frame()
            a, b = seq.frame_stats(), auto.frame_stats()
            assert (a.rays, a.steps, a.hits, a.noise_fetches, a.limit_exhausted) == (b.rays, b.steps, b.hits, b.noise_fetches, b.limit_exhausted), k
            ratios.append(b.steps / max(b.rays, 1))
            for buf in (N.BUF_DEPTH, N.BUF_NORMAL, N.BUF_ALBEDO, N.BUF_FOG, N.BUF_LIGHTING, N.BUF_ACCUM):
                assert np.array_equal(seq.read_buffer(buf), auto.read_buffer(buf), equal_nan=True), (k, buf)
        assert ratios[0] > 28.0 and max(ratios[:-1]) > 28.0, ("the scene is meant to cross the policy's threshold after its first frame", ratios)
