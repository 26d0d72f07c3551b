"""The C-ABI library loads on a machine without a GPU, exports every symbol include/vxrt.h declares, keeps
the 192-byte frame block layout of the reference, and FAILS LOUDLY instead of falling back when no CUDA
device exists."""
import ctypes as C
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _header_symbols():
    text = open(os.path.join(ROOT, "include", "vxrt.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(vxrt_[a-z_0-9]+)\s*\(", text)))


def test_header_and_binding_agree():
    from raytrace_b200 import _native as N
    names = _header_symbols()
    assert len(names) >= 25
    assert set(names) == set(N.SYMBOLS), set(names) ^ set(N.SYMBOLS)


def test_library_exports_every_declared_symbol():
    from raytrace_b200 import _native as N
    lib = C.CDLL(N.LIB_PATH)
    for name in _header_symbols():
        assert getattr(lib, name) is not None
    assert lib.vxrt_abi_version() == N.ABI_VERSION


def test_struct_layouts():
    from raytrace_b200 import _native as N
    F = N.VxrtFrameParams
    # structs.rs:3-31 (#[repr(C)], std140-compatible): byte offsets of the fields the shaders read
    assert C.sizeof(F) == 192
    assert (F.sun_angle.offset, F.seed.offset, F.origin.offset, F.forward.offset, F.up.offset, F.right.offset) == (0, 4, 16, 32, 48, 64)
    assert (F.old_origin.offset, F.region_offset.offset, F.rotation.offset, F.space_offset.offset) == (80, 144, 160, 176)
    assert C.sizeof(N.VxrtEdit) == 16
    assert C.sizeof(N.VxrtConfig) == 64


def test_plan_bands_and_passes():
    import raytrace_b200 as R
    from raytrace_b200 import _native as N
    for H in (1080, 2160, 256, 100):
        for n in (1, 2, 3, 4, 8):
            b = R.plan_bands(H, n)
            assert b[0][0] == 0 and b[-1][1] == H
            assert all(b[i][1] == b[i + 1][0] for i in range(n - 1))
            assert all(e % 8 == 0 for _, e in b[:-1])
    trace, passes = R.plan_passes(1080, 136, 272, N.FLAG_DENOISE)
    assert trace == (136 - 117, 272 + 117)
    # pass k may only read rows the previous pass (or the trace) produced: +-3*size
    prev = trace
    for (r0, r1), size in zip(passes, (1, 2, 4, 8, 8, 16)):
        assert r0 - 3 * size >= prev[0] and r1 + 3 * size <= prev[1]
        prev = (r0, r1)
    assert passes[-1] == (136, 272)
    trace, passes = R.plan_passes(1080, 0, 136, N.FLAG_DENOISE)
    assert trace == (0, 253) and passes[-1] == (0, 136)
    assert R.plan_passes(1080, 136, 272, 0)[0] == (136, 272)


@pytest.mark.skipif(os.path.exists("/dev/nvidiactl"), reason="needs a machine WITHOUT a GPU")
def test_no_cpu_fallback():
    import raytrace_b200 as R
    with pytest.raises(R.VxrtError) as e:
        R.Renderer(64, 64, (64, 64, 64))
    assert e.value.code == -2 and "no CUDA device" in str(e.value)


def test_argument_validation_without_device():
    from raytrace_b200 import _native as N
    h = C.c_void_p()
    cfg = N.VxrtConfig()
    cfg.abi_version = 99
    assert N.lib.vxrt_create(C.byref(cfg), C.byref(h)) == N.ERR_INVALID
    assert b"abi_version" in N.lib.vxrt_last_error(None)
    cfg.abi_version = N.ABI_VERSION
    cfg.width, cfg.height, cfg.nx, cfg.ny, cfg.nz = 64, 64, 100, 64, 64      # not a power of two
    assert N.lib.vxrt_create(C.byref(cfg), C.byref(h)) == N.ERR_INVALID
    assert N.lib.vxrt_create(None, C.byref(h)) == N.ERR_INVALID
    assert N.lib.vxrt_render_frame(None) == N.ERR_INVALID
    N.lib.vxrt_destroy(None)


@pytest.mark.skipif(os.path.exists("/dev/nvidiactl"), reason="needs a machine WITHOUT a GPU")
def test_cpp_host_driver_fails_loudly_without_device(tmp_path):
    """csrc/host/vxrt_main (the C++ caller of the ABI) must surface VXRT_ERR_CUDA, not render on the CPU."""
    import subprocess
    import numpy as np
    exe = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "raytrace_b200", "vxrt_main")
    assert os.path.exists(exe), "vxrt_main not built (raytrace_b200/build.py)"
    (tmp_path / "w.u32").write_bytes(np.zeros(64 ** 3, np.uint32).tobytes())
    (tmp_path / "n.rgba").write_bytes(np.zeros(512 * 512 * 4, np.uint8).tobytes())
    p = subprocess.run([exe, "--world", str(tmp_path / "w.u32"), "--noise", str(tmp_path / "n.rgba"), "--dims", "64", "64", "64",
                        "--size", "64", "64"], capture_output=True, text=True, timeout=60)
    assert p.returncode == 1 and "libvxrt error -2" in p.stderr and "no CUDA device" in p.stderr


def test_plan_bands_weighted():
    """Pure host arithmetic: bands tile the frame on 8-row edges; cheap (sky) rows make wider bands; uniform cost
    gives the edge ranks (one denoise halo instead of two) the taller bands."""
    import numpy as np
    import raytrace_b200 as R
    H = 1080
    t = np.full(H, 100.0, np.float32); d = np.full(H, 10.0, np.float32)
    for n in (1, 2, 3, 4, 8):
        b = R.plan_bands_weighted(H, n, t, d)
        assert b[0][0] == 0 and b[-1][1] == H and all(b[i][1] == b[i + 1][0] for i in range(n - 1))
        assert all(lo % 8 == 0 for lo, _ in b) and all(hi > lo for lo, hi in b)
    b = R.plan_bands_weighted(H, 8, t, d)
    assert b[0][1] - b[0][0] > b[3][1] - b[3][0] and b[7][1] - b[7][0] > b[4][1] - b[4][0]
    t[:400] = 10.0; d[:400] = 0.5                                 # rows 0..399 nearly free
    b = R.plan_bands_weighted(H, 2, t, d)
    assert b[0][1] > 650
    assert R.plan_bands_weighted(64, 8, t[:64], d[:64]) == [(8 * i, 8 * i + 8) for i in range(8)]
    with pytest.raises(R.VxrtError):
        R.plan_bands_weighted(32, 8, t[:32], d[:32])              # fewer than 8 rows per rank
