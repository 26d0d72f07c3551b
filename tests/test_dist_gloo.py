"""world_size-2 `gloo` test of the multi-GPU HOST logic on CPU: the band plan of the C ABI
(vxrt_plan_bands / vxrt_plan_passes) drives two ranks that each produce their row band with the oracle
(trace rows incl. the 117-row halo, denoise passes on the shrinking row ranges, finalize on the band);
gathering the bands must reproduce the single-process frame bit for bit, and a broadcast edit list must
leave every replica with the same world."""
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _worker(rank, world_size, port, q):
    sys.path.insert(0, ROOT)
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world_size))
    import torch
    import torch.distributed as dist
    import raytrace_b200 as R
    from raytrace_b200 import _native as N, scenes
    from oracle import orc
    dist.init_process_group("gloo", rank=rank, world_size=world_size)
    try:
        dims, kind, seed = scenes.WORLDS["c1"]
        m, f = orc.build_minefield(orc.generate_world(*dims, kind, seed))      # replicated world
        noise = orc.blue_noise()
        m_path, f_path = m.copy(), f.copy()        # the world before the edit broadcast below
        W, H = 96, 320
        cam = scenes.camera(dims)
        blk = orc.fill_frame(cam["origin"], cam["heading"], cam["pitch"], cam["sun_angle"], 4)
        r0, r1 = R.plan_bands(H, world_size)[rank]
        trace_rows, pass_rows = R.plan_passes(H, r0, r1, N.FLAG_DENOISE)
        fr = orc.Frame(W, H)
        L = orc.lib()
        import ctypes as C
        sc = orc.OrcScene(dims[0], dims[1], dims[2], orc._p(m), orc._p(f), orc._p(noise), 2048)
        b = fr.buffers()
        L.orc_trace(C.byref(sc), orc._p(blk), 2, C.byref(b), trace_rows[0], trace_rows[1], C.byref(fr.stats), 2)
        pong = np.zeros_like(fr.lighting)
        for k, (size, (p0, p1)) in enumerate(zip(orc.DENOISE_SIZES, pass_rows)):
            src, dst = (fr.lighting, pong) if k % 2 == 0 else (pong, fr.lighting)
            L.orc_denoise_pass(orc._p(src), orc._p(dst), orc._p(fr.depth), orc._p(fr.normal), W, H, size, 0, p0, p1, 2)
        L.orc_finalize(C.byref(b), orc._p(fr.lighting), orc._p(noise), orc._p(fr.output), r0, r1, 2)
        # gather the bands on rank 0: G-buffer rows [r0,r1) are output rows [H-r1, H-r0)
        mine = torch.from_numpy(fr.output[H - r1:H - r0].copy())
        bands = R.plan_bands(H, world_size)
        if rank == 0:
            parts = [torch.empty((e - s, W, 4), dtype=torch.uint8) for s, e in bands]
            dist.gather(mine, parts, dst=0)
            img = np.zeros((H, W, 4), np.uint8)
            for (s, e), p in zip(bands, parts):
                img[H - e:H - s] = p.numpy()
            full = orc.render(m, f, noise, blk, 2, W, H, denoise=True, finalize=True)
            ok_img = bool(np.array_equal(img, full.output))
        else:
            dist.gather(mine, None, dst=0)
            ok_img = True
        # edit broadcast: rank 0 owns the list, every replica applies it
        n = 500
        e = torch.from_numpy(orc.generate_edits(n, *dims, 7, 0)) if rank == 0 else torch.zeros((n, 4), dtype=torch.int32)
        dist.broadcast(e, src=0)
        orc.apply_edits(m, e.numpy())
        m2, f2 = orc.build_minefield(m)
        digest = torch.tensor([int(m2.astype(np.uint64).sum() % (1 << 31)), int(f2.astype(np.uint64).sum() % (1 << 31))])
        both = [torch.zeros_like(digest) for _ in range(world_size)]
        dist.all_gather(both, digest)
        ok_world = all(bool(torch.equal(both[0], x)) for x in both)
        # frames of a camera path dealt round-robin (plan_frames): the ranks' frames together are the path, in order
        n_path = 5
        mine_idx = R.plan_frames(n_path, world_size)[rank]
        sums = torch.zeros(n_path, dtype=torch.int64)
        for i in mine_idx:
            c = scenes.camera(dims, i, n_path)
            bi = orc.fill_frame(c["origin"], c["heading"], c["pitch"], c["sun_angle"], i + 1)
            sums[i] = int(orc.render(m_path, f_path, noise, bi, 1, 64, 40, denoise=True, finalize=True).output.astype(np.int64).sum())
        dist.all_reduce(sums)
        ok_frames = True
        if rank == 0:
            for i in range(n_path):
                c = scenes.camera(dims, i, n_path)
                bi = orc.fill_frame(c["origin"], c["heading"], c["pitch"], c["sun_angle"], i + 1)
                ok_frames &= int(orc.render(m_path, f_path, noise, bi, 1, 64, 40, denoise=True, finalize=True).output.astype(np.int64).sum()) == int(sums[i])
        q.put((rank, ok_img, ok_world and ok_frames))
    finally:
        dist.destroy_process_group()


def test_two_rank_band_render_and_edit_broadcast():
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29600 + os.getpid() % 300
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = [q.get(timeout=240) for _ in procs]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert all(ok_img and ok_world for _, ok_img, ok_world in res), res
