"""The oracle against (i) known answers derived from reference code, (ii) the golden fixtures produced by the
reference's own shader text (tests/golden/make_golden.py), (iii) oracle/_ref live when it has been built."""
import glob
import os

import numpy as np
import pytest

GOLDEN = sorted(glob.glob(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "*.npz")))
G_BUFFERS = ["lighting_raw", "depth", "normal", "albedo", "emission", "fog", "lighting", "output"]


def test_material_words_known_answers(orc):
    """GEN_MATERIALS.rs:44-51 + build.rs:50-79 over misc/materials.csv."""
    import raytrace_b200 as R
    want = [0x0, 0x1FC07F, 0x9F73D, 0xCD319, 0xCD9B3, 0xF8D96, 0x1BBA73]
    assert [m.pack() for m in R.MATERIALS] == want
    csv = [(0, 0, 0), (255, 0, 255), (79, 221, 122), (102, 77, 51), (102, 102, 102), (124, 54, 44), (221, 233, 231)]
    assert [orc.lib().orc_material_word_from_csv(i, *c) for i, c in enumerate(csv)] == want
    # bit 15 (solid) overlaps red bit 1: Material::black() decodes red = 2
    assert R.Material.unpack(orc.lib().orc_pack_word(0, 0, 0, 1)).albedo == (2, 0, 0)
    assert R.MATERIALS[3].emission == (160 * 4 // 2, 77 * 4 // 2, 38 * 4 // 2)


def test_pack_into_known_answers(orc):
    """chunk.rs:125-184: smallest k whose 2^k-aligned cube holds a solid; empty chunk -> all 6, words 0."""
    solid = 0x1FC07F
    m, f = orc.pack_chunk(np.zeros((64, 64, 64), np.uint32))
    assert (f == 6).all() and (m == 0).all()
    m, f = orc.pack_chunk(np.full((64, 64, 64), solid, np.uint32))
    assert (f == 0).all() and (m == solid).all()
    w = np.zeros((64, 64, 64), np.uint32); w[0, 0, 0] = solid
    m, f = orc.pack_chunk(w)
    z, y, x = np.indices((64, 64, 64))
    d = np.maximum(np.maximum(x, y), z)
    want = np.where(d == 0, 0, np.ceil(np.log2(d + 1)).astype(int)).astype(np.uint8)   # 1 -> 1, 2..3 -> 2, 4..7 -> 3 ...
    assert np.array_equal(f, want) and f.max() == 6
    # a chunk whose words never set bit 15 is "empty": wiped (chunk.rs:155-161)
    m, f = orc.pack_chunk(np.full((64, 64, 64), 0x7F, np.uint32))
    assert (f == 6).all() and (m == 0).all()
    # LODs never span chunks
    big = np.zeros((64, 64, 128), np.uint32); big[0, 0, 63] = solid
    m, f = orc.build_minefield(big)
    assert (f[:, :, 64:] == 6).all() and f[0, 0, 62] == 1


def test_pixel_workgroup_map(orc):
    """raytrace.comp:291-294 is a bijection on 128x128 super-tiles; the reference grid (W/8, H/8+1) leaves
    24 rows of a 1080p frame unwritten."""
    import ctypes as C
    L = orc.lib()
    px, py = C.c_int(), C.c_int()
    seen = set()
    for wg in range(32):
        for l in range(8):
            L.orc_wg_to_pixel(wg, 0, l, 0, C.byref(px), C.byref(py))
            seen.add(px.value)
            wx, wy = C.c_int(), C.c_int()
            L.orc_pixel_to_wg(px.value, 0, C.byref(wx), C.byref(wy))
            assert wx.value == wg
    assert seen == set(range(256))
    rows = set()
    for wgy in range(1080 // 8 + 1):
        for l in range(8):
            L.orc_wg_to_pixel(0, wgy, 0, l, C.byref(px), C.byref(py))
            rows.add(py.value)
    missing = sorted(set(range(1080)) - rows)
    assert len(missing) == 24 and missing[:8] == list(range(1032, 1040)) and missing[-8:] == list(range(1064, 1072))


def test_analytic_rays(orc, noise):
    """Empty 256^3 window: origin (0.5,0.5,0.5), dir +x -> 4 iterations of step 32, exits through face code 1."""
    f = np.full((256, 256, 256), 6, np.uint8); m = np.zeros((256, 256, 256), np.uint32)
    r = orc.trace_ray(m, f, noise, (0, 0, 0), (0.5, 0.5, 0.5), (1, 0, 0))
    assert r["air"] and r["steps"] == 4 and r["face"] == 1 and r["voxel"][0] == 128
    # a solid slab at texel x = 200 (world x = 72): hit on its -x face (code 1 = moving +x), word returned
    m[:, :, 200] = 0xCD9B3; m, f = orc.build_minefield(m)
    r = orc.trace_ray(m, f, noise, (0, 0, 0), (0.5, 0.5, 0.5), (1, 0, 0))
    assert not r["air"] and r["face"] == 1 and r["voxel"] == (72, 0, 0) and r["word"] == 0xCD9B3
    assert abs(r["t"] - 71.5) < 1e-3
    # toroidal addressing: the same slab seen with the window centre moved by 16
    r2 = orc.trace_ray(m, f, noise, (16, 0, 0), (0.5, 0.5, 0.5), (1, 0, 0))
    assert not r2["air"] and r2["voxel"] == (72, 0, 0)
    # start inside a solid: defined outcome (NaN position, "hit" of word 0 after one iteration)
    r3 = orc.trace_ray(m, f, noise, (0, 0, 0), (72.5, 0.5, 0.5), (0, 1, 0))
    assert not r3["air"] and r3["steps"] == 1 and r3["word"] == 0 and np.isnan(r3["t"])


def test_frame_block_layout(orc):
    blk = orc.fill_frame((1.0, 2.0, 3.0), np.pi / 2, 0.0, 0.25, 9, (16, 0, -16))
    f = blk.view(np.float32); i = blk.view(np.int32)
    assert f[0] == np.float32(0.25) and i[1] == 9
    assert tuple(f[4:7]) == (1.0, 2.0, 3.0)
    assert abs(f[9] - 1.0) < 1e-6 and abs(f[8]) < 1e-6                # forward = +Y at heading pi/2
    assert abs(f[14] - 0.4) < 1e-6                                    # up * 0.4 = (0,0,0.4)
    assert abs(f[16] - 0.4) < 1e-6                                    # right * 0.4 = (0.4,0,0)
    assert tuple(i[40:43]) == (16, 0, -16) and tuple(i[44:47]) == (16, 0, -16)


@pytest.mark.parametrize("path", GOLDEN, ids=[os.path.basename(p) for p in GOLDEN])
def test_oracle_matches_reference_golden(orc, world, noise, path):
    """Bit equality of every image the reference shaders write (fixtures made from the reference text)."""
    g = np.load(path)
    m, f = world("mid")
    H, W = g["depth"].shape
    fr = orc.render(m, f, noise, g["frame_block"], orc.MODE_FULL, W, H, denoise=True, finalize=True,
                    pong_swap=bool(g["pong_swap"]))
    for name in G_BUFFERS:
        assert np.array_equal(getattr(fr, name), g[name]), name
    assert fr.stats.ub_limit_exhausted == 0 and fr.stats.ub_coord_eq_n == 0


def test_oracle_matches_reference_live(orc, world, noise):
    """oracle/_ref (the reference's shader text compiled here) against the hand restatement, more cameras."""
    if not orc.ref_available():
        pytest.skip("oracle/_ref not built (needs /root/reference: python oracle/make_ref.py)")
    from raytrace_b200 import scenes
    m, f = world("mid")
    dims = scenes.WORLDS["mid"][0]
    for k, (W, H, lr, swap) in enumerate([(64, 64, (0, 0, 0), False), (136, 72, (-32, 16, 0), False), (48, 200, (0, 0, 16), True)]):
        cam = scenes.camera(dims, k * 3, 8)
        blk = orc.fill_frame(cam["origin"], cam["heading"], cam["pitch"], 0.3 + 0.4 * k, 40 + k, lr)
        a = orc.render(m, f, noise, blk, orc.MODE_FULL, W, H, denoise=True, finalize=True, pong_swap=swap)
        b = orc.ref_render(m, f, noise, blk, W, H, pong_swap=swap)
        for name in G_BUFFERS:
            assert np.array_equal(getattr(a, name), getattr(b, name)), (k, name)
    # the reference's own grid (W/8, H/8+1) leaves rows unwritten for heights that are not multiples of 128
    blk = orc.fill_frame(cam["origin"], cam["heading"], cam["pitch"], 0.5, 3, (0, 0, 0))
    b = orc.ref_render(m, f, noise, blk, 128, 200, denoise=False, finalize=False, cover_all_rows=False)
    a = orc.render(m, f, noise, blk, orc.MODE_FULL, 128, 200)
    written = np.all(b.lighting_raw[..., 3] != 0, axis=1)      # alpha = 1/16 wherever a store happened
    # 26 group rows: rows 0..127 complete; in 128..199 only rows with (row - 128) % 16 < 10 are produced
    expect = np.array([r < 128 or (r - 128) % 16 < 10 for r in range(200)])
    assert np.array_equal(written, expect)
    assert np.array_equal(a.lighting_raw[written], b.lighting_raw[written])


def test_modes_are_truncations(orc, world, noise):
    """PRIMARY / SHADOW / FULL share the primary hit; ray counts per pixel are 1 / <=2 / <=5."""
    from raytrace_b200 import scenes
    m, f = world("c1")
    dims = scenes.WORLDS["c1"][0]
    cam = scenes.camera(dims)
    blk = orc.fill_frame(cam["origin"], cam["heading"], cam["pitch"], cam["sun_angle"], 2)
    fr = [orc.render(m, f, noise, blk, mode, 96, 64) for mode in (0, 1, 2)]
    for a in fr[1:]:
        assert np.array_equal(a.hit_voxel, fr[0].hit_voxel) and np.array_equal(a.hit_word, fr[0].hit_word)
    assert fr[0].rays.max() == 1 and fr[1].rays.max() == 2 and fr[2].rays.max() == 5
    hit = fr[0].hit_face < 16
    assert (fr[1].rays[hit] == 2).all() and (fr[1].rays[~hit] == 1).all()
    assert fr[0].stats.rays == 96 * 64


def test_generator_is_deterministic_and_edits_unique(orc):
    a = orc.generate_world(64, 64, 64, 1, 9); b = orc.generate_world(64, 64, 64, 1, 9)
    assert np.array_equal(a, b) and not np.array_equal(a, orc.generate_world(64, 64, 64, 1, 10))
    e = orc.generate_edits(10000, 512, 512, 512, 3, 0)
    assert len(np.unique(e[:, :3], axis=0)) == 10000
    assert (np.abs(e[:, :3]) <= 256).all()
    assert set(np.unique(e[:, 3])) == {0, 0xCD9B3}


def test_ray_log_agrees_with_the_pixel_counters(orc):
    """orc_set_ray_log (the hook tests/analysis_divergence.py costs the march loop with): per pixel, the five logged
    march lengths add up to the steps counter and the non-zero ones count the rays."""
    from raytrace_b200 import scenes
    dims, kind, seed = scenes.WORLDS["small"]
    m, f = orc.build_minefield(orc.generate_world(*dims, kind, seed))
    cam = scenes.camera(dims)
    blk = orc.fill_frame(cam["origin"], cam["heading"], cam["pitch"], cam["sun_angle"], 3)
    W, H = 96, 64
    log = np.zeros((H, W, 5), np.uint16)
    L = orc.lib()
    L.orc_set_ray_log.argtypes = [orc.C.c_void_p]
    L.orc_set_ray_log(orc._p(log))
    try:
        fr = orc.render(m, f, orc.blue_noise(), blk, orc.MODE_FULL, W, H)
    finally:
        L.orc_set_ray_log(None)
    assert np.array_equal(log.sum(axis=2, dtype=np.uint32), fr.steps)
    assert np.array_equal((log > 0).sum(axis=2).astype(np.uint8), fr.rays)
    assert np.all(log[..., 0] > 0)                                   # every pixel has a primary ray
    assert np.all((log[..., 3] > 0) == (log[..., 4] > 0))            # sun2 and dif2 come together
    assert np.all((log[..., 3] == 0) | (log[..., 2] > 0))            # and only after dif1


def test_scene_edit_stream_and_frame_blocks_match_oracle(orc):
    """raytrace_b200/scenes.py defines the benchmark's inputs without the oracle: its integer edit stream must equal
    orc_generate_edits bit for bit, and its frame blocks orc_fill_frame's up to the last bit of host trigonometry
    (blocks are inputs of the boundary; both bench arms use scenes.py's)."""
    from raytrace_b200 import scenes
    for dims in [(512, 512, 512), (256, 256, 256), (128, 128, 128), (2048, 2048, 256)]:
        for n, fr in [(10000, 0), (2000, 3), (7, 1), (1, 0)]:
            a, b = orc.generate_edits(n, *dims, 7, fr), scenes.edit_stream(n, dims, 7, fr)
            assert np.array_equal(a, b), (dims, n, fr)
            assert len(np.unique(b[:, :3], axis=0)) == n
    with pytest.raises(ValueError):
        scenes.edit_stream(4, (64, 64, 64), 7, 0)
    for i in (0, 17, 63):
        cam = scenes.camera((512, 512, 512), i, 64)
        a = orc.fill_frame(cam["origin"], cam["heading"], cam["pitch"], cam["sun_angle"], i + 1, (16, 0, -32))
        b = np.frombuffer(scenes.frame_block(cam, i + 1, (16, 0, -32)), np.uint8)
        assert b.size == 192 and np.array_equal(a[:32], b[:32]) and np.array_equal(a[80:], b[80:])
        assert np.abs(a[32:80].view(np.float32) - b[32:80].view(np.float32)).max() <= 2 ** -23
