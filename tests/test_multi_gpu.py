"""Multi-process, multi-GPU tests (need >= 2 GPUs; skipped otherwise): one process per GPU, the peer-memory frame
of peer.cu over CUDA IPC, and the NCCL band gather / edit broadcast of comm.cu.  Rank 0's gathered image must
equal the single-handle frame bit for bit."""
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _ngpus():
    try:
        import torch
        return torch.cuda.device_count()
    except Exception:  # noqa: BLE001
        return 0


def _worker(rank, world_size, port, comm, q):
    sys.path.insert(0, ROOT)
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world_size))
    import ctypes as C
    import torch
    import torch.distributed as dist
    import raytrace_b200 as R
    from raytrace_b200 import _native as N, scenes
    from oracle import orc
    torch.cuda.set_device(rank)
    dist.init_process_group("gloo", rank=rank, world_size=world_size)     # host channel only (blobs / ids)
    try:
        dims, kind, seed = scenes.WORLDS["c1"]
        W, H = 256, 512
        flags = N.FLAG_DENOISE | N.FLAG_FINALIZE
        noise = orc.blue_noise()
        blocks = []
        for s in (3, 4, 5):
            cam = scenes.camera(dims, s, 8)
            blocks.append(orc.fill_frame(cam["origin"], cam["heading"], cam["pitch"], cam["sun_angle"], s))
        band = R.plan_bands(H, world_size)[rank]
        r = R.Renderer(W, H, dims, N.MODE_FULL, flags, device=rank, rows=band)
        r.upload_blue_noise(noise); r.generate_world(kind, seed)
        if comm == "peer":
            blobs = [None] * world_size
            dist.all_gather_object(blobs, r.peer_export())
            r.peer_attach(blobs, rank, root=0)
        else:
            ident = [None]
            if rank == 0:
                buf = (C.c_uint8 * N.COMM_ID_BYTES)()
                assert N.lib.vxrt_comm_unique_id(buf, N.COMM_ID_BYTES) == 0
                ident[0] = bytes(buf)
            dist.broadcast_object_list(ident, src=0)
            r._check(N.lib.vxrt_comm_init(r.handle, ident[0], N.COMM_ID_BYTES, rank, world_size))
        imgs = []
        for b in blocks:
            r.set_frame(b); r.render_frame()
            if comm == "nccl":
                r._check(N.lib.vxrt_comm_gather_output(r.handle, 0))
            if rank == 0:
                imgs.append(r.read_output().copy())
        r.sync(); r.frame_stats()
        ok = True
        if rank == 0:
            with R.Renderer(W, H, dims, N.MODE_FULL, flags, device=0) as full:
                full.upload_blue_noise(noise); full.generate_world(kind, seed)
                for b, img in zip(blocks, imgs):
                    full.set_frame(b); full.render_frame()
                    ok = ok and bool(np.array_equal(full.read_output(), img))
        # edit broadcast (NCCL): every replica ends with the same world
        ok_world = True
        if comm == "nccl":
            n = 2000
            e = orc.generate_edits(n, *dims, 9, 0) if rank == 0 else np.zeros((n, 4), np.int32)
            e = np.ascontiguousarray(e, np.int32)
            r._check(N.lib.vxrt_comm_broadcast_edits(r.handle, e.ctypes.data, n, 0))
            m, f = r.read_world()
            digest = torch.tensor([int(m.astype(np.uint64).sum() % (1 << 31)), int(f.astype(np.uint64).sum() % (1 << 31))])
            both = [torch.zeros_like(digest) for _ in range(world_size)]
            dist.all_gather(both, digest)
            ok_world = all(bool(torch.equal(both[0], x)) for x in both)
            if rank == 0:
                m0 = orc.generate_world(*dims, kind, seed)
                orc.apply_edits(m0, e)
                m1, f1 = orc.build_minefield(m0)
                ok_world = ok_world and bool(np.array_equal(m1, m) and np.array_equal(f1, f))
        dist.barrier()
        r.close()
        q.put((rank, ok, ok_world))
    finally:
        dist.destroy_process_group()


@pytest.mark.gpu
@pytest.mark.parametrize("comm", ["peer", "nccl"])
def test_two_process_frame(comm):
    n = _ngpus()
    if n < 2:
        pytest.skip("needs >= 2 GPUs")
    import torch.multiprocessing as mp
    ws = 2 if n < 4 else 4
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29900 + os.getpid() % 300 + (0 if comm == "peer" else 1)
    procs = [ctx.Process(target=_worker, args=(r, ws, port, comm, q)) for r in range(ws)]
    for p in procs:
        p.start()
    res = [q.get(timeout=300) for _ in procs]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert all(a and b for _, a, b in res), res
