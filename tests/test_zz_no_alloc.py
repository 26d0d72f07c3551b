"""vxrt_apply_edits_static (include/vxrt.h): the edit path for callers that may not allocate.  Its scratch is a file-scope
device array; the caller passes its own window arrays and a device edit list.  Checked through the C ABI against the existing
entry point (vxrt_apply_edits on a handle + vxrt_read_world): voxel words and occupancy field bit for bit (integer work, the
tolerance every edit test here uses), on two calls in a row - the second call finds the first one's table and flags in the shared
array and must clear them.  Collected last on purpose."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def test_static_scratch_edits_match_handle_edits(orc, world):
    import torch
    import raytrace_b200 as R
    from raytrace_b200 import _native as N, scenes
    dims = scenes.WORLDS["c1"][0]
    nx, ny, nz = dims
    m, f = world("c1")
    rng = np.random.default_rng(11)
    lists = []
    for frame in range(2):
        base = scenes.edit_stream(3000, dims, 9, frame)
        dup = base[rng.integers(0, 150, size=2000)].copy()               # voxels named many times: last entry wins
        dup[:, 3] = rng.integers(0, 2, size=len(dup)) * scenes.SOLID_EDIT_WORD
        lists.append(np.ascontiguousarray(np.concatenate([base, dup]), np.int32))
    assert all(len(e) <= N.STATIC_MAX_EDITS for e in lists)
    dev = torch.device("cuda", 0)
    tm = torch.from_numpy(m.view(np.int32).copy()).to(dev)               # the caller's own window arrays
    tf = torch.from_numpy(f.copy()).to(dev)
    lib = N.lib
    with R.Renderer(64, 64, dims, 0, 0) as r:
        r.upload_world(m, f)
        for e in lists:                                                  # two calls in a row on the shared scratch
            r.apply_edits(e)                                             # existing entry point
            want_m, want_f = r.read_world()
            te = torch.from_numpy(e).to(dev)
            torch.cuda.synchronize()
            rc = lib.vxrt_apply_edits_static(0, tm.data_ptr(), tf.data_ptr(), nx, ny, nz, te.data_ptr(), len(e), None)
            assert rc == N.OK, lib.vxrt_last_error(None).decode()
            torch.cuda.synchronize()
            got_m = tm.cpu().numpy().view(np.uint32)
            got_f = tf.cpu().numpy()
            assert np.array_equal(got_m, want_m), "voxel words"
            assert np.array_equal(got_f, want_f), "occupancy field"
    # also against the oracle: set_block in list order, then pack_into
    ref = m.copy()
    for e in lists:
        orc.apply_edits(ref, e)
    ref_m, ref_f = orc.build_minefield(ref)
    assert np.array_equal(got_f, ref_f) and np.array_equal(got_m, ref_m)
    # limits: a clear error, nothing written
    rc = lib.vxrt_apply_edits_static(0, tm.data_ptr(), tf.data_ptr(), nx, ny, nz, te.data_ptr(), N.STATIC_MAX_EDITS + 1, None)
    assert rc == N.ERR_SIZE and b"VXRT_STATIC_MAX_EDITS" in lib.vxrt_last_error(None)
    rc = lib.vxrt_apply_edits_static(0, tm.data_ptr(), tf.data_ptr(), nx, ny, 100, te.data_ptr(), 4, None)
    assert rc == N.ERR_INVALID
    rc = lib.vxrt_apply_edits_static(0, None, tf.data_ptr(), nx, ny, nz, te.data_ptr(), 4, None)
    assert rc == N.ERR_INVALID
