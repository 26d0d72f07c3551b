"""Chunk files in the reference's format (src/world/chunk_storage.rs:37-68): name = three {:016X}
fields of the signed chunk coordinates, contents = LZ4 frame of materials (u32 LE) || minefield (u8).  The reference
links liblz4 through the `lz4` crate; liblz4 is reachable here through pyarrow's "lz4" (frame) codec, so both
directions are checked against the real library."""
import numpy as np
import pytest


def _terrain_chunk(orc, seed):
    rng = np.random.default_rng(seed)
    z = np.arange(64)[:, None, None]
    h = 20 + rng.integers(0, 12, size=(64, 64))
    words = np.where(z < h[None], np.uint32(0x9F73D), np.uint32(0)).astype(np.uint32)
    return orc.pack_chunk(words)


def test_file_name_known_answers():
    from raytrace_b200 import world as W
    assert W.chunk_file_name((0, 0, 0)) == "0" * 48
    assert W.chunk_file_name((1, 2, 255)) == "0000000000000001" "0000000000000002" "00000000000000FF"
    assert W.chunk_file_name((-2, 0, 5)) == "FFFFFFFFFFFFFFFE" "0000000000000000" "0000000000000005"   # isize as two's complement


def test_round_trip_and_liblz4_interop(orc):
    pa = pytest.importorskip("pyarrow")
    import raytrace_b200 as R
    from raytrace_b200 import world as W
    m, f = _terrain_chunk(orc, 0)
    raw = m.astype("<u4").tobytes() + f.tobytes()
    blob = W.encode_chunk(R.PackedChunkData(m, f))
    assert len(blob) < len(raw) // 4                                   # terrain compresses
    back = W.decode_chunk(blob)
    assert np.array_equal(back.materials, m) and np.array_equal(back.minefield, f)
    assert pa.decompress(blob, decompressed_size=len(raw), codec="lz4", asbytes=True) == raw      # liblz4 reads our frame
    theirs = pa.compress(raw, codec="lz4", asbytes=True)              # linked 64 KiB blocks, as the lz4 crate writes
    back = W.decode_chunk(theirs)
    assert np.array_equal(back.materials, m) and np.array_equal(back.minefield, f)
    rng = np.random.default_rng(1)                                     # incompressible -> stored block
    m2 = rng.integers(0, 2 ** 32, size=(64, 64, 64), dtype=np.uint32); f2 = rng.integers(0, 256, size=(64, 64, 64), dtype=np.uint8)
    b2 = W.encode_chunk(R.PackedChunkData(m2, f2))
    back = W.decode_chunk(b2)
    assert np.array_equal(back.materials, m2) and np.array_equal(back.minefield, f2)
    assert pa.decompress(b2, decompressed_size=len(raw), codec="lz4", asbytes=True) == m2.astype("<u4").tobytes() + f2.tobytes()


def test_bad_files_are_rejected(orc):
    import raytrace_b200 as R
    from raytrace_b200 import world as W
    m, f = _terrain_chunk(orc, 2)
    blob = W.encode_chunk(R.PackedChunkData(m, f))
    for bad in (blob[:-3], blob[:200], b"\x00" * 64, blob[:4] + b"\xff" + blob[5:]):
        with pytest.raises(ValueError):
            W.decode_chunk(bad)
    # content checksum catches payload damage (a stored block: in run-length terrain a damaged match offset can still
    # reproduce the same bytes, which is not an error)
    rng = np.random.default_rng(5)
    noisy = W.encode_chunk(R.PackedChunkData(rng.integers(0, 2 ** 32, size=(64, 64, 64), dtype=np.uint32),
                                             rng.integers(0, 256, size=(64, 64, 64), dtype=np.uint8)))
    flipped = bytearray(noisy); flipped[len(noisy) // 2] ^= 0x10
    with pytest.raises(ValueError):
        W.decode_chunk(bytes(flipped))
    with pytest.raises(ValueError):
        W.encode_chunk(R.PackedChunkData(m, None))


def test_disk_chunk_storage(orc, tmp_path):
    """save -> new storage over the same directory -> borrow_packed_chunk_data reads the file; unknown chunks are air."""
    import raytrace_b200 as R
    from raytrace_b200 import world as W
    m, f = _terrain_chunk(orc, 3)
    st = W.DiskChunkStorage(str(tmp_path))
    st.save((-1, 0, 2), R.PackedChunkData(m, f))
    assert (tmp_path / ("F" * 16 + "0" * 16 + "0" * 15 + "2")).exists()
    st2 = W.DiskChunkStorage(str(tmp_path))
    got = st2.borrow_packed_chunk_data((-1, 0, 2))
    assert np.array_equal(got.materials, m) and np.array_equal(got.minefield, f)
    assert not st2.borrow_packed_chunk_data((9, 9, 9)).materials.any()


@pytest.mark.gpu
def test_stored_chunks_feed_the_device_and_validate(orc, tmp_path):
    """SURVEY 8(f)2 end to end on the GPU: chunks written in the reference's file format (chunk_storage.rs:37-68), read back
    through DiskChunkStorage -> vxrt_chunk_decode -> vxrt_upload_chunk(materials, NULL): the occupancy field the device builds
    from the stored words equals the stored one (vxrt_validate_chunk, compared on the device), a corrupted file is caught with
    the offset of the damage, and the whole window equals the world the files were cut from."""
    import raytrace_b200 as R
    from raytrace_b200 import scenes
    from raytrace_b200.world import DiskChunkStorage
    dims, kind, seed = scenes.WORLDS["c1"]
    m = orc.generate_world(*dims, kind, seed)
    m, f = orc.build_minefield(m)
    src = R.ChunkStorage.from_window(m, f)
    disk = DiskChunkStorage(str(tmp_path))
    for coord, data in src.chunks.items():
        disk.save(coord, data)
    fresh = DiskChunkStorage(str(tmp_path))                     # nothing in memory: every chunk comes off the disk
    n = dims[0] // 64
    with R.Renderer(64, 64, dims, 0, 0) as r:
        for cz in range(n):
            for cy in range(n):
                for cx in range(n):
                    data = fresh.borrow_packed_chunk_data((cx - n // 2, cy - n // 2, cz - n // 2))
                    r.upload_chunk(cx, cy, cz, data.materials, None)       # occupancy built on the device
                    assert r.validate_chunk(cx, cy, cz, data.minefield) == (0, -1)
        gm, gf = r.read_world()
        assert np.array_equal(gm, m) and np.array_equal(gf, f)
        bad = fresh.borrow_packed_chunk_data((0, 0, 0)).minefield.copy()
        bad[5, 6, 7] ^= 1; bad[40, 1, 2] ^= 3
        assert r.validate_chunk(n // 2, n // 2, n // 2, bad) == (2, (5 * 64 + 6) * 64 + 7)
        with pytest.raises(R.VxrtError):
            r.validate_chunk(n, 0, 0, bad)
