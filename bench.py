#!/usr/bin/env python
"""bench.py — headline benchmark of the voxel ray-tracing hot path (BASELINE.json metric).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl vxrt|reference] [--workload c2|c3|c1|c4|c5]

A "step" is one frame of the workload's camera path per GPU: trace -> 6x denoise -> finalize.  Default workload =
BASELINE config 2: 1920x1080, primary + 1 shadow ray per pixel, 512^3 seeded biome world, 64-frame camera path.
One JSON line is printed by rank 0.

At N > 1 the frames of a camera path are independent units, so they are dealt round-robin to the GPUs (frame i on
rank i % N; world replicated; each rank reads its frames back itself; no collective on the data path; "weak"
scaling: a job-wide step is N frames).  The same run then also times the image-space split north_star describes — ONE
frame cut into load-balanced row bands, halo rows and finished bands stored over NVLink peer memory — and reports it
under `image_bands` ("strong" scaling: the latency of a frame), with the root's image compared against a single-GPU
frame of the same block.  Workloads whose frame k+1 depends on frame k (c3: temporal accumulation, c5: edit stream)
only have the band form.  --shard frames|bands forces one of them.

  value     : Mrays/s over the whole step, world + frame block already resident in HBM, CUDA-event timed.
  e2e       : the same through the public call sequence with HOST buffers: vxrt_set_frame(192-byte host block)
              + vxrt_render_frame + vxrt_read_output(host RGBA8), copies inside the timed region.
  sustained : the device-resident loop again for >= 2 s, with its own clock record (the burst loop lasts milliseconds).
  parity    : the first frames of the timed path rendered once more with hit buffers and compared with the CPU oracle
              (the cpu_baseline leg renders them anyway): mismatch counts, max |dt|, max |dcolour|, UB counters.
  --impl reference : the reference path's CPU form (the oracle: C++ restatement of the GLSL passes, OpenMP
              over all host cores; the reference itself has no runnable form here — DESIGN.md §2).

The product arm never touches oracle/: frame blocks, edit streams and worlds come from raytrace_b200 (scenes.py,
vxrt_generate_world); only the cpu_baseline + parity leg, after all timing, loads the oracle as the checker.
"""
import argparse
import importlib.util
import json
import os
import statistics
import sys
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

import numpy as np  # noqa: E402

WORKLOADS = {
    # name: (scene, width, height, mode, n_path_frames, accumulate, description)
    "c1": ("c1", 256, 256, 0, 1, False, "256x256 primary-ray frame, 128^3 heightmap world (BASELINE config 1)"),
    "c2": ("c2", 1920, 1080, 1, 64, False,
           "1920x1080 primary + 1 shadow ray/pixel, 512^3 biome world, 64-frame camera path (BASELINE config 2)"),
    "c3": ("c2", 1920, 1080, 2, 1, True,
           "1080p full lighting (primary + 2 bounces + sky), temporal accumulation, 512^3 biome world (BASELINE config 3)"),
    "c4": ("c4", 3840, 2160, 2, 16, False, "3840x2160 full lighting on a 2048x2048x256 world (BASELINE config 4)"),
    "c5": ("c2", 1920, 1080, 2, 64, False,
           "interactive edit stream: 10k voxel edits/frame + occupancy rebuild of the touched chunks (+ NCCL broadcast at N>1), "
           "then 1080p full-lighting re-render, 512^3 biome world (BASELINE config 5)"),
}
EDITS_PER_FRAME = {"c5": 10000}
EDIT_SEED = 7
IN_FLIGHT = {"c1": 4, "c2": 2, "c3": 1, "c4": 1}   # frames in flight per GPU (--in-flight 0; measured 1..4, misc/r2/run_z.sh: C2 2939 / 3529 / 3516 / 3516 frames/s,
                                                    # C1 12.6k / 23.9k / 31.2k / 38.7k, C4 315 / 286 / 286 / 287); c3 can chain two handles through the
                                                    # shared running sum (--in-flight 2) but measures no gain: 0.672 against 0.660 ms
MODE_NAMES = {0: "primary", 1: "primary+shadow", 2: "full"}


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=64)
    ap.add_argument("--warmup", type=int, default=4)
    ap.add_argument("--impl", default="vxrt", choices=["vxrt", "reference"])
    ap.add_argument("--workload", default="c2", choices=sorted(WORKLOADS))
    ap.add_argument("--comm", default="peer", choices=["peer", "nccl"],
                    help="N > 1: peer = halo rows and output bands stored over NVLink peer memory by our own kernels (peer.cu); "
                         "nccl = every band re-traces its denoise halo, NCCL send/recv gathers the bands (comm.cu)")
    ap.add_argument("--bands", default="balanced", choices=["balanced", "equal"],
                    help="N > 1: balanced = row bands from per-row work counters of three calibration frames "
                         "(vxrt_plan_bands_weighted); equal = equal-height bands")
    ap.add_argument("--shard", default="auto", choices=["auto", "frames", "bands"],
                    help="N > 1: frames = the frames of the camera path are dealt round-robin to the GPUs, each renders whole "
                         "frames on its replica of the world and reads them back itself (throughput; weak scaling, no "
                         "collective on the data path); bands = ONE frame is split into row bands (latency; strong scaling). "
                         "auto = frames where the path's frames are independent (c1, c2, c4), bands where frame k+1 depends on "
                         "frame k (c3: temporal accumulation, c5: edit stream); an auto run at N > 1 also times the band "
                         "split and reports it under `image_bands`")
    ap.add_argument("--in-flight", type=int, default=0,
                    help="frames in flight per GPU (handles sharing one world, vxrt_create_shared) where frames are "
                         "independent (c1, c2, c4 on one GPU or sharded by frames); 1 = the reference's one frame in flight; "
                         "0 = the workload's measured best: 2 for c1/c2, 1 for c4 (a 4K frame over the 5 GiB world fills the GPU on "
                         "its own: two in flight measured 237 frames/s against 264)")
    ap.add_argument("--no-cpu-baseline", action="store_true", help="skip the bounded oracle timing + parity check on rank 0")
    ap.add_argument("--no-l2-flush", action="store_true", help="do not flush L2 between timed steps")
    ap.add_argument("--sustained-seconds", type=float, default=2.0,
                    help="length of the sustained leg (device-resident loop repeated; 0 = skip)")
    return ap.parse_args()


def load_scenes():
    """raytrace_b200/scenes.py loaded by path: pure Python, so the reference arm never imports the package (and never
    maps libvxrt.so)."""
    spec = importlib.util.spec_from_file_location("vxrt_scenes", os.path.join(ROOT, "raytrace_b200", "scenes.py"))
    m = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(m)
    return m


def resolve_shard(args):
    chained = WORKLOADS[args.workload][5] or args.workload in EDITS_PER_FRAME
    return args.shard if args.shard != "auto" else ("bands" if chained else "frames")


def make_config(args, scenes, world_size, shard):
    """The workload both arms run, as one dict with the same keys and values in both (the driver compares them)."""
    scene, W, H, mode, n_path, accumulate, desc = WORKLOADS[args.workload]
    dims = scenes.WORLDS[scene][0]
    by_frames = shard == "frames" and world_size > 1
    chained = args.workload in EDITS_PER_FRAME
    S = (args.in_flight or IN_FLIGHT.get(args.workload, 1)) if (world_size == 1 or by_frames) and not chained else 1
    world_mib = 5 * dims[0] * dims[1] * dims[2] >> 20
    if world_size == 1:
        par = "one GPU"
    elif by_frames:
        par = (f"camera-path frames dealt round-robin to {world_size} GPUs (frame i on rank i % {world_size}), each renders whole "
               f"frames and reads them back itself; world replicated; a step = {world_size} frames")
    else:
        par = f"image row bands x{world_size}, world replicated"
    return {"workload": desc, "mode": MODE_NAMES[mode], "width": W, "height": H, "world": list(dims),
            "camera_path_frames": n_path, "edits_per_frame": EDITS_PER_FRAME.get(args.workload, 0), "accumulate": bool(accumulate),
            "passes": "trace + 6 denoise + finalize", "parallelism": par, "frames_in_flight": S,
            "l2": ((f"GPU arm: flushed before the timed loop; inside it frames overlap and the inputs (world {world_mib} MiB) are larger than L2"
                    if world_mib > 126 else
                    f"GPU arm: flushed before the timed loop; inside it frames overlap and are not flushed (world {world_mib} MiB fits L2); "
                    "`one_frame_in_flight` is the flushed figure")
                   if S > 1 else
                   ("GPU arm: not flushed (inputs larger than L2)" if args.no_l2_flush else "GPU arm: flushed between steps (256 MiB fill, untimed)"))}


class ClockSampler:
    """SM clock + throttle reasons sampled DURING the timed region (the B200_PROFILING.md clocks line, read
    through NVML in a thread every 2 ms because the timed loop is only tens of milliseconds long)."""

    def __init__(self, device):
        import threading
        self.samples, self.max_mhz, self.reasons, self._stop = [], None, set(), threading.Event()
        self.err = None
        try:
            import pynvml
            pynvml.nvmlInit()
            # NVML indexes physical GPUs; honour CUDA_VISIBLE_DEVICES when it lists ordinals
            vis = os.environ.get("CUDA_VISIBLE_DEVICES")
            idx = device
            if vis:
                ids = [v.strip() for v in vis.split(",") if v.strip()]
                if device < len(ids) and ids[device].isdigit():
                    idx = int(ids[device])
            self.nv, self.h = pynvml, pynvml.nvmlDeviceGetHandleByIndex(idx)
            self.max_mhz = float(pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM))
        except Exception as e:  # noqa: BLE001
            self.err = f"nvml unavailable: {e}"
            return
        self.t = threading.Thread(target=self._run, daemon=True)
        self.t.start()

    def _run(self):
        nv = self.nv
        names = {"hw_slowdown": nv.nvmlClocksEventReasonHwSlowdown if hasattr(nv, "nvmlClocksEventReasonHwSlowdown") else 0x8,
                 "hw_thermal_slowdown": 0x40, "sw_thermal_slowdown": 0x20, "sw_power_cap": 0x4}
        while not self._stop.is_set():
            try:
                self.samples.append(float(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM)))
                try:
                    r = nv.nvmlDeviceGetCurrentClocksEventReasons(self.h)
                except Exception:  # noqa: BLE001
                    r = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                for k, bit in names.items():
                    if r & bit:
                        self.reasons.add(k)
            except Exception as e:  # noqa: BLE001
                self.err = str(e)
                break
            self._stop.wait(0.002)

    def stop(self):
        if self.err and not self.samples:
            return {"sm_mhz": None, "sm_max_mhz": self.max_mhz, "samples": 0, "reasons": [self.err]}
        self._stop.set()
        self.t.join(timeout=2)
        return {"sm_mhz": statistics.median(self.samples) if self.samples else None, "sm_max_mhz": self.max_mhz,
                "samples": len(self.samples), "reasons": sorted(self.reasons)}


def algo_bytes_trace(stats, pixels):
    """per ray (1 + S_r) minefield bytes + 4 B if it ends in a hit; per pixel 23 B of G-buffer
    stores; 4 B per blue-noise fetch."""
    return (stats.rays + stats.steps) + 4 * stats.hits + 23 * pixels + 4 * stats.noise_fetches


def host_threads():
    try:
        return len(os.sched_getaffinity(0))
    except AttributeError:
        return os.cpu_count() or 1


def cpu_apply_edits(orc, m, f, edits):
    """The reference's edit path on the CPU: set_block per voxel (chunk.rs:115), then pack_into of every touched
    64^3 chunk (chunk.rs:125-184)."""
    nz, ny, nx = m.shape
    orc.apply_edits(m, edits)
    tex = (edits[:, :3] + np.array([nx // 2, ny // 2, nz // 2], np.int32)) % np.array([nx, ny, nz], np.int32)
    for cx, cy, cz in {tuple(c) for c in (tex // 64).tolist()}:
        sl = (slice(cz * 64, cz * 64 + 64), slice(cy * 64, cy * 64 + 64), slice(cx * 64, cx * 64 + 64))
        m[sl], f[sl] = orc.pack_chunk(m[sl])


def time_oracle(orc, world, noise, blocks, mode, W, H, steps, warmup, accumulate, edits=None, threads=0, keep=None):
    """warmup + steps oracle frames of the path (blocks[i], edits[i] applied first); returns (seconds, rays) of the timed
    ones.  keep: list that receives a copy of every frame's buffers (the parity leg compares them with the device's)."""
    m, f = world
    fr = orc.Frame(W, H, accumulate=accumulate)
    times, rays = [], 0
    for i in range(warmup + steps):
        t0 = time.perf_counter()
        if edits is not None:
            cpu_apply_edits(orc, m, f, edits[i % len(edits)])
        orc.render(m, f, noise, np.frombuffer(blocks[i % len(blocks)], np.uint8), mode, W, H, denoise=True, finalize=True, frame=fr,
                   threads=threads)
        dt = time.perf_counter() - t0
        if i >= warmup:
            times.append(dt); rays += fr.stats.rays
        if keep is not None:
            keep.append(dict(hit_voxel=fr.hit_voxel.copy(), hit_face=fr.hit_face.copy(), hit_word=fr.hit_word.copy(),
                             hit_t=fr.hit_t.copy(), output=fr.output.copy(), limit=int(fr.stats.ub_limit_exhausted),
                             start_in_solid=int(fr.stats.ub_start_in_solid), nan_dir=int(fr.stats.ub_nan_dir),
                             coord_eq_n=int(fr.stats.ub_coord_eq_n), rays=int(fr.stats.rays), steps=int(fr.stats.steps)))
    return sum(times), rays


def run_reference(args, rank, world_size):
    """--impl reference: the reference path's CPU form (oracle) on the host cores, rank 0 only.  torchrun exports
    OMP_NUM_THREADS=1 to its workers; the oracle is told its thread count explicitly instead."""
    if rank != 0:
        return
    scenes = load_scenes()
    from oracle import orc
    scene, W, H, mode, n_path, accumulate, desc = WORKLOADS[args.workload]
    dims, kind, seed = scenes.WORLDS[scene]
    threads = host_threads()
    m = orc.generate_world(*dims, kind, seed)
    world = orc.build_minefield(m)
    noise = orc.blue_noise()
    blocks = scenes.path_blocks(dims, n_path, args.warmup + args.steps)
    ne = EDITS_PER_FRAME.get(args.workload, 0)
    edits = [scenes.edit_stream(ne, dims, EDIT_SEED, i) for i in range(args.warmup + args.steps)] if ne else None
    world = (np.ascontiguousarray(world[0]), np.ascontiguousarray(world[1]))
    total, rays = time_oracle(orc, world, noise, blocks, mode, W, H, args.steps, args.warmup, accumulate, edits, threads)
    v = rays / total / 1e6
    shard = resolve_shard(args)
    line = {"metric": "Mrays/s", "value": v, "unit": "Mrays/s", "impl": "reference", "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": total / args.steps * 1e3, "frames_per_s": args.steps / total,
            "higher_is_better": True, "scaling": "weak" if shard == "frames" else "strong",   # the label of the vxrt arm's line
            "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": make_config(args, scenes, args.gpus, shard),
            "cpu_baseline": {"value": v, "unit": "Mrays/s", "cores": threads, "kind": "port",
                             "sample": f"{args.steps} full frames (trace + 6 denoise + finalize) of the workload on rank 0's host cores"},
            "e2e": {"value": v, "unit": "Mrays/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0,
            "note": "reference GLSL/Vulkan path cannot run on this box; this is the CPU restatement (oracle/) with OpenMP"}
    _emit(line)


def parity_stats(dev, ref):
    """Device buffers of one frame against the oracle's: mismatch counts on the integer buffers, max abs errors on the rest."""
    tg, tc = dev["hit_t"], ref["hit_t"]
    ok = ~np.isnan(tc) & ~np.isnan(tg)
    return {"hit_voxel_mismatch": int((dev["hit_voxel"] != ref["hit_voxel"]).any(axis=-1).sum()),
            "face_mismatch": int((dev["hit_face"] != ref["hit_face"]).sum()),
            "word_mismatch": int((dev["hit_word"] != ref["hit_word"]).sum()),
            "nan_t_mismatch": int((np.isnan(tg) != np.isnan(tc)).sum()),
            "max_abs_dt": float(np.max(np.abs(tg[ok] - tc[ok]), initial=0.0)),
            "max_abs_dcolour": int(np.max(np.abs(dev["output"].astype(np.int16) - ref["output"].astype(np.int16)), initial=0))}


def run_vxrt(args, rank, world_size, local_rank, shard="bands", dist=None):
    """One measurement.  shard = "bands": one frame split into row bands over the ranks; "frames": every rank renders
    whole frames, frame i of the path on rank i % world_size.  Returns the JSON line on rank 0 (None elsewhere)."""
    import ctypes as C
    import torch
    import raytrace_b200 as R
    from raytrace_b200 import _native as N, scenes
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; libvxrt has no CPU path (use --impl reference for the CPU baseline)")
    by_frames = shard == "frames" and world_size > 1
    scene, W, H, mode, n_path, accumulate, desc = WORKLOADS[args.workload]
    dims, kind, seed = scenes.WORLDS[scene]
    flags = N.FLAG_DENOISE | N.FLAG_FINALIZE | (N.FLAG_ACCUMULATE if accumulate else 0)
    noise = R.world.default_blue_noise()
    config = make_config(args, scenes, world_size, shard)
    bands = [(0, H)]
    if by_frames:
        pass                                      # every rank renders whole frames
    elif world_size > 1 and args.bands == "balanced":
        # Load-balanced bands (setup, untimed): every rank renders three frames of the path on a full-frame handle,
        # sums the per-row work counters and runs the same deterministic plan, so no rank needs to be told its rows.
        with R.Renderer(W, H, dims, mode, (flags | N.FLAG_HIT_BUFFERS) & ~N.FLAG_ACCUMULATE, device=local_rank) as cal:
            cal.upload_blue_noise(noise); cal.generate_world(kind, seed)
            tc, dc = np.zeros(H, np.float32), np.zeros(H, np.float32)
            cb = scenes.path_blocks(dims, n_path, n_path)
            for k in sorted({0, n_path // 2, n_path - 1}):
                cal.set_frame(np.frombuffer(cb[k], np.uint8)); cal.render_frame()
                t, d = cal.row_costs(); tc += t; dc += d
        bands = R.plan_bands_weighted(H, world_size, tc, dc, flags)
    elif world_size > 1:
        bands = R.plan_bands(H, world_size)
    band = bands[0] if by_frames else bands[rank]
    r = R.Renderer(W, H, dims, mode, flags, device=local_rank, rows=band if world_size > 1 and not by_frames else None)
    r.upload_blue_noise(noise)
    r.generate_world(kind, seed)                 # world replicated: every rank generates the same seeded world
    lib, h = N.lib, r.handle
    S = config["frames_in_flight"]
    # S handles on this GPU share r's world (one copy in HBM); frames go to them round-robin, so S frames are in flight
    rs = [r] + [R.Renderer(W, H, dims, mode, flags, device=local_rank, share_world=r) for _ in range(S - 1)]
    if by_frames:
        pass                                      # independent replicas: nothing to attach
    elif world_size > 1 and args.comm == "peer":
        blobs = [None] * world_size
        dist.all_gather_object(blobs, r.peer_export())       # 512-byte CUDA IPC blobs over the host channel, once
        r.peer_attach(blobs, rank, root=0)
    elif world_size > 1:
        ident = [None]
        if rank == 0:
            buf = (C.c_uint8 * N.COMM_ID_BYTES)()
            assert lib.vxrt_comm_unique_id(buf, N.COMM_ID_BYTES) == 0
            ident[0] = bytes(buf)
        dist.broadcast_object_list(ident, src=0)
        r._check(lib.vxrt_comm_init(h, ident[0], N.COMM_ID_BYTES, rank, world_size))

    # frame blocks: host-side parameter fill (pipeline.rs:191-207)
    n_frames = args.warmup + args.steps
    # frames mode: job-wide step k is frames k*N .. k*N+N-1 of the path, this rank renders frame k*N + rank
    if by_frames:
        mine_idx = R.plan_frames(n_frames * world_size, world_size)[rank]
        assert mine_idx == list(range(rank, n_frames * world_size, world_size))
    raw_blocks = scenes.path_blocks(dims, n_path, n_frames, rank, world_size) if by_frames else scenes.path_blocks(dims, n_path, n_frames)
    blocks = [np.frombuffer(b, np.uint8) for b in raw_blocks]
    stream = torch.cuda.ExternalStream(r.cuda_stream(), device=torch.device("cuda", local_rank))
    flush = None if args.no_l2_flush else torch.empty(256 << 20, dtype=torch.uint8, device=f"cuda:{local_rank}")

    def barrier():
        torch.cuda.synchronize()
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()

    ne = EDITS_PER_FRAME.get(args.workload, 0)
    edits = [np.ascontiguousarray(scenes.edit_stream(ne, dims, EDIT_SEED, i), np.int32) for i in range(n_frames)] if ne else None
    if ne and by_frames:
        raise SystemExit("bench.py: the edit stream (c5) makes frame k+1 depend on frame k; use --shard bands")
    if ne and world_size > 1 and args.comm == "peer":      # edit lists travel by NCCL broadcast in either frame mode
        ident = [None]
        if rank == 0:
            buf = (C.c_uint8 * N.COMM_ID_BYTES)()
            assert lib.vxrt_comm_unique_id(buf, N.COMM_ID_BYTES) == 0
            ident[0] = bytes(buf)
        dist.broadcast_object_list(ident, src=0)
        r._check(lib.vxrt_comm_init(h, ident[0], N.COMM_ID_BYTES, rank, world_size))

    def step(i):
        if edits is not None:                    # host edit list -> device -> every replica rebuilds its touched chunks
            if world_size > 1:
                r._check(lib.vxrt_comm_broadcast_edits(h, edits[i].ctypes.data, ne, 0))
            else:
                r.apply_edits(edits[i])
        r.set_frame(blocks[i])
        r.render_frame()
        if world_size > 1 and args.comm == "nccl" and not by_frames:
            r._check(lib.vxrt_comm_gather_output(h, 0))

    # ---- device-resident timing: per-step CUDA events on the handle's stream, L2 flushed between steps ----
    for i in range(args.warmup):
        step(i); r.sync()
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(args.steps)]
    tot = dict(rays=0, steps=0, hits=0, noise=0, limit=0, trace_ms=0.0, denoise_ms=0.0, finalize_ms=0.0, algo=0)
    barrier()
    clocks = ClockSampler(local_rank)
    wall0 = time.perf_counter()
    for k in range(args.steps):
        if flush is not None:
            with torch.cuda.stream(stream):
                flush.fill_(k & 0xFF)            # evicts L2 (126 MB) — untimed, before the step's start event
        if world_size > 1 and not by_frames:     # align the ranks (untimed): a rank that starts late would otherwise
            if args.comm == "peer":              # put its L2 flush inside its neighbours' timed region
                r.peer_barrier()                 # device-side barrier on the handle's stream (peer.cu)
            else:
                torch.cuda.synchronize(); dist.barrier()
        ev[k][0].record(stream)
        step(args.warmup + k)
        ev[k][1].record(stream)
        s = r.frame_stats()                      # syncs the stream; per-kernel event times of this step
        tot["rays"] += s.rays; tot["steps"] += s.steps; tot["hits"] += s.hits; tot["noise"] += s.noise_fetches
        tot["limit"] += s.limit_exhausted
        tot["trace_ms"] += s.trace_ms; tot["denoise_ms"] += s.denoise_ms; tot["finalize_ms"] += s.finalize_ms
        tot["algo"] += algo_bytes_trace(s, (band[1] - band[0]) * W)   # own rows only (halo rows are extra, uncounted work)
    barrier()
    wall = time.perf_counter() - wall0
    dev_ms = serial_ms = sum(a.elapsed_time(b) for a, b in ev)

    # ---- S frames in flight: the same K frames dealt round-robin to the S handles, one event pair around the loop ----
    # (no L2 flush inside the loop: frames overlap; the inputs - 640 MiB world and up - are larger than the 126 MB L2)
    streams = [torch.cuda.ExternalStream(x.cuda_stream(), device=torch.device("cuda", local_rank)) for x in rs]

    def overlapped_loop(n):
        """n frames of the path dealt round-robin to the S handles; returns the device time in ms (events on stream 0,
        every other stream starting after the first event and joined before the second)."""
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        done = [torch.cuda.Event() for _ in rs]
        e0.record(streams[0])
        for j in range(1, S):
            streams[j].wait_event(e0)
        for k in range(n):
            x = rs[k % S]
            x.set_frame(blocks[args.warmup + k % args.steps]); x.render_frame()
        for j in range(1, S):
            done[j].record(streams[j]); streams[0].wait_event(done[j])
        e1.record(streams[0])
        torch.cuda.synchronize()
        return e0.elapsed_time(e1)

    if S > 1:
        for k in range(2 * S):                   # untimed: every handle has rendered before
            rs[k % S].set_frame(blocks[k % n_frames]); rs[k % S].render_frame()
        for x in rs:
            x.sync()
        if flush is not None:
            with torch.cuda.stream(stream):
                flush.fill_(1)
        barrier()
        wall0 = time.perf_counter()
        dev_ms = overlapped_loop(args.steps)
        barrier()
        wall = time.perf_counter() - wall0
    clk = clocks.stop()

    # ---- sustained leg: the same device-resident loop for >= sustained-seconds, with its own clock record ----
    sustained = None
    if args.sustained_seconds > 0 and edits is None and (world_size == 1 or by_frames):
        n_sus = int(min(200000, max(args.steps, args.sustained_seconds * 1e3 / (dev_ms / args.steps) * 1.05)))
        barrier()
        c2 = ClockSampler(local_rank)
        sus_ms = overlapped_loop(n_sus)
        barrier()
        sustained = {"frames": n_sus, "seconds": sus_ms / 1e3, "ms_per_step": sus_ms / n_sus, "clocks": c2.stop()}

    # ---- end to end through the public call sequence with host buffers -------------------------------------
    # Per frame: vxrt_set_frame(192-byte host block) -> vxrt_render_frame -> vxrt_read_output_async into one of two
    # page-locked host images -> vxrt_wait_output of the PREVIOUS frame (the double-buffered read-back of
    # include/vxrt.h: frame k's D2H copy overlaps frame k+1's kernels).  Every frame's 8.3 MB lands in host memory
    # inside the timed region.
    if edits is not None:                        # the edit stream changes the world: replay it from the state the
        r.generate_world(kind, seed)             # device-timed loop started from (untimed), so that both loops
        for i in range(args.warmup):             # render the same sequence of worlds
            step(i)
        r.sync()
    reads = rank == 0 or by_frames               # frames mode: every rank delivers the frames it rendered to host memory
    out_host = [np.empty((H, W, 4), np.uint8) for _ in range(2 * S)]
    if reads:
        for b in out_host:
            r.host_register(b)
    barrier()
    t0 = time.perf_counter()
    prev = [None] * S
    for k in range(args.steps):
        j = k % S
        if S == 1:
            step(args.warmup + k)
        else:
            rs[j].set_frame(blocks[args.warmup + k]); rs[j].render_frame()
        if reads:
            t = rs[j].read_output_async(out_host[2 * j + ((k // S) & 1)])
            if prev[j] is not None:
                rs[j].wait_output(prev[j])
            prev[j] = t
    for j in range(S):
        if reads and prev[j] is not None:
            rs[j].wait_output(prev[j])
        rs[j].sync()
    barrier()
    e2e_s = time.perf_counter() - t0
    if reads:
        for b in out_host:
            r.host_unregister(b)
    # What the host side of this box can take: every rank that delivers frames copies a device buffer of one image's size to
    # page-locked host memory with plain cudaMemcpyAsync calls (torch), all ranks at once, untimed work outside the metric.  Beside
    # the rate the end-to-end loop reached it says whether the library or the host is the limit at N GPUs.
    d2h_probe = None
    nb = W * H * 4
    n_probe = max(8, min(256, int(2e9 // nb)))
    if reads:
        src = torch.empty(nb, dtype=torch.uint8, device=f"cuda:{local_rank}")
        dst = torch.empty(nb, dtype=torch.uint8).pin_memory()
        for _ in range(2):
            dst.copy_(src, non_blocking=True)
        torch.cuda.synchronize()
    barrier()                                    # every rank, reading or not
    if reads:
        p0 = time.perf_counter()
        for _ in range(n_probe):
            dst.copy_(src, non_blocking=True)
        torch.cuda.synchronize()
        p1 = time.perf_counter()
        d2h_probe = {"plain_copy_gbs_rank0": nb * n_probe / (p1 - p0) / 1e9, "e2e_loop_gbs_rank0": nb * args.steps / e2e_s / 1e9,
                     "copies": n_probe, "bytes": nb,
                     "note": "plain pinned cudaMemcpyAsync of one image per call, every delivering rank at once, against the D2H rate of the end-to-end loop"}
    barrier()

    # ---- band mode: the root's image against a single-GPU frame of the same block (untimed) -----------------
    bands_match = None
    if world_size > 1 and not by_frames and edits is None:
        if accumulate:
            r.reset_accumulation()
        chk = blocks[args.warmup]
        r.set_frame(chk); r.render_frame()
        if args.comm == "nccl":
            r._check(lib.vxrt_comm_gather_output(h, 0))
        r.sync()
        barrier()
        if rank == 0:
            got = r.read_output()
            with R.Renderer(W, H, dims, mode, flags, device=local_rank) as one:
                one.upload_blue_noise(noise); one.generate_world(kind, seed)
                one.set_frame(chk); one.render_frame()
                want = one.read_output()
            bad_rows = int((got != want).any(axis=(1, 2)).sum())
            bands_match = {"matches_single_gpu": bad_rows == 0, "rows_differing": bad_rows}
        barrier()

    # ---- reduce over ranks: time = max, work = sum ---------------------------------------------------------
    vals = torch.tensor([dev_ms, e2e_s, wall, serial_ms, sustained["seconds"] if sustained else 0.0], dtype=torch.float64,
                        device=f"cuda:{local_rank}")
    sums = torch.tensor([tot["rays"], tot["steps"], tot["hits"], tot["algo"], tot["limit"]], dtype=torch.float64,
                        device=f"cuda:{local_rank}")
    if dist is not None:
        dist.all_reduce(vals, op=dist.ReduceOp.MAX)
        dist.all_reduce(sums, op=dist.ReduceOp.SUM)
    dev_ms, e2e_s, wall, serial_ms, sus_s = vals.tolist()
    rays, steps_total, hits, algo, limit = sums.tolist()
    mine = {"trace": tot["trace_ms"] / args.steps, "denoise": tot["denoise_ms"] / args.steps, "finalize": tot["finalize_ms"] / args.steps,
            "rows": list(band)}
    per_rank = [mine]
    if dist is not None:
        per_rank = [None] * world_size
        dist.all_gather_object(per_rank, mine)
    line = None
    nf = world_size if by_frames else 1           # frames per job-wide step
    if rank == 0:
        peaks = {}
        try:
            peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        except OSError:
            pass
        peak = float(peaks.get("hbm_gbs", 6650.0))
        prof = {}                                # figures of the committed ncu capture of this workload's trace pass, if any
        try:
            prof = json.load(open(os.path.join(ROOT, "profiles", "ncu_traffic.json"))).get(args.workload, {})
        except (OSError, ValueError):
            pass
        own_px = (band[1] - band[0]) * W
        trace_s = tot["trace_ms"] / 1e3
        achieved = tot["algo"] / trace_s / 1e9
        launches_per_frame = 8                    # k_trace_fast, 6 k_denoise, k_finalize
        if hasattr(lib, "vxrt_launches_per_frame"):
            launches_per_frame = int(lib.vxrt_launches_per_frame(h))

        def kroof(name, bytes_per_frame, ms_total, n_launch, what):
            t = ms_total / args.steps / 1e3
            a = bytes_per_frame / t / 1e9 if t > 0 else 0.0
            return {"kernel": name, "bound": "hbm", "achieved": a, "peak": peak, "unit": "GB/s", "frac": a / peak,
                    "algo_bytes_per_frame": bytes_per_frame, "ms_per_frame": t * 1e3, "launches_per_frame": n_launch, "algorithmic": what}
        line = {
            "metric": "Mrays/s", "value": rays / (dev_ms / 1e3) / 1e6, "unit": "Mrays/s", "n_gpus": world_size,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": dev_ms / args.steps,
            "frames_per_s": args.steps * nf / (dev_ms / 1e3), "higher_is_better": True, "scaling": "weak" if shard == "frames" else "strong",   # frames dealt to the GPUs: per-GPU work fixed as N grows
            "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": config,
            "run": {"bands": [list(b) for b in bands], "comm": None if world_size == 1 or by_frames else args.comm,
                    "band_passes": None if world_size == 1 or by_frames else
                                   ("bands re-trace their denoise halo, NCCL send/recv gathers the bands" if args.comm == "nccl" else
                                    "halo rows pushed / finished bands stored into the root image over NVLink peer memory")},
            # one frame in flight, L2 flushed before every frame, an event pair per frame: where the per-kernel times and
            # the roofline below come from (with frames overlapping, kernels of two frames share the SMs)
            "one_frame_in_flight": {"ms_per_step": serial_ms / args.steps, "value": rays / (serial_ms / 1e3) / 1e6,
                                    "frames_per_s": args.steps * nf / (serial_ms / 1e3),
                                    "l2": "not flushed" if flush is None else "flushed between steps (256 MiB fill, untimed)"},
            "rays_per_frame": rays / (args.steps * nf), "mean_steps_per_ray": steps_total / max(rays, 1.0),
            "limit_exhausted": limit,
            "kernel_ms_per_step_rank0": {"trace": tot["trace_ms"] / args.steps, "denoise": tot["denoise_ms"] / args.steps,
                                          "finalize": tot["finalize_ms"] / args.steps},
            "kernel_ms_per_step_by_rank": per_rank,
            "trace_mrays_per_s_rank0": tot["rays"] / trace_s / 1e6,
            "e2e": {"value": rays / e2e_s / 1e6, "unit": "Mrays/s", "frames_per_s": args.steps * nf / e2e_s,
                    "h2d_bytes_per_step": (192 + 16 * ne) * nf, "d2h_bytes_per_step": W * H * 4 * nf, "d2h_probe": d2h_probe},
            # per frame and rank: the handle's kernels (vxrt_launches_per_frame); peer mode adds k_halo_push + 3 k_peer_sync (+1 on the root)
            "gpu_launches": args.steps * (launches_per_frame * world_size + ((4 * world_size + 1) if world_size > 1 and args.comm == "peer" and not by_frames else 0)),
            "roofline": {"kernel": "k_trace_fast", "bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s",
                         "frac": achieved / peak, "peak_source": "MEASURED_PEAKS.json hbm_gbs" if peaks else "fallback 6650",
                         "algo_bytes_per_launch": tot["algo"] / args.steps, "launch_ms": tot["trace_ms"] / args.steps,
                         "traffic": prof.get("dram_bytes_per_launch"),
                         "launch_timing": "CUDA events around the launch, one frame in flight, L2 flushed before the frame",
                         "ncu": {k: prof.get(k) for k in ("issue_active_pct", "lanes_per_instruction", "warp_instructions", "l2_hit_pct", "source")} if prof else None,
                         "note": "k_trace_fast is bound by instruction issue, not by memory (the ncu figures beside this note; the touched part of the "
                                 "occupancy field is L2-resident on the 512^3 world); frac is reported against HBM as the contract asks"},
            "roofline_kernels": (
                [kroof("k_denoise (6 passes, finalize.comp folded into the last)", (19 * 6 + 22) * own_px, tot["denoise_ms"] + tot["finalize_ms"], 6,
                       "19 B/pixel/pass: read 8+2+1, write 8; the last pass also reads 4+4+4+4 (albedo, emission, fog, noise) and writes the 4 B output (SURVEY 8d minus finalize's 8 B lighting + 2 B depth re-read)")]
                if launches_per_frame == 7 else
                [kroof("k_denoise (6 passes)", 19 * 6 * own_px, tot["denoise_ms"], 6, "19 B/pixel/pass: read 8+2+1, write 8 (SURVEY 8d)"),
                 kroof("k_finalize", 30 * own_px, tot["finalize_ms"], 1, "30 B/pixel: read 4+4+4+8+2+4, write 4 (SURVEY 8d)")]),
            "clocks": clk, "wall_s_timed_loop": wall,
        }
        if sustained:
            n_sus = sustained["frames"]
            line["sustained"] = {"seconds": sus_s, "frames": n_sus * nf, "ms_per_step": sus_s * 1e3 / n_sus,
                                 "frames_per_s": n_sus * nf / sus_s, "value": rays / args.steps * n_sus / sus_s / 1e6, "unit": "Mrays/s",
                                 "clocks": sustained["clocks"],
                                 "note": "the device-resident loop of `value` repeated back to back (host submission included; no L2 flush: frames overlap)"}
        if bands_match is not None:
            line["matches_single_gpu"] = bands_match["matches_single_gpu"]
            line["rows_differing"] = bands_match["rows_differing"]
        if not args.no_cpu_baseline and world_size == 1:
            line.update(cpu_baseline_and_parity(args, R, N, r, noise, raw_blocks, edits, kind, seed))
    for x in reversed(rs):
        x.close()
    return line


def cpu_baseline_and_parity(args, R, N, r, noise, raw_blocks, edits, kind, seed):
    """After all timing: the oracle (the checker; the only use of oracle/ in this arm) renders the first frames of the
    timed path on the host cores — that is `cpu_baseline` — and the same frames are rendered once more on the device with
    hit buffers and compared — that is `parity`."""
    from oracle import orc
    scene, W, H, mode, n_path, accumulate, desc = WORKLOADS[args.workload]
    nb, wu = 2, 1
    seq = raw_blocks[args.warmup:]
    eseq = None if edits is None else edits[args.warmup:]
    if edits is not None:
        r.generate_world(kind, seed)             # the edit stream starts from the generated world in both legs
    m, f = r.read_world()
    threads = host_threads()
    kept = []
    t, cr = time_oracle(orc, (m.copy(), f.copy()) if edits is not None else (m, f), noise, seq, mode, W, H, nb, wu, accumulate, eseq, threads, kept)
    out = {"cpu_baseline": {"value": cr / t / 1e6, "unit": "Mrays/s", "frames_per_s": nb / t, "cores": threads,
                            "kind": "port", "sample": f"{nb} full frames of the same workload through the oracle (OpenMP), after 1 warm-up frame"}}
    # the same wu + nb frames on the device, hit buffers on, sharing r's world
    pflags = N.FLAG_DENOISE | N.FLAG_FINALIZE | N.FLAG_HIT_BUFFERS | (N.FLAG_ACCUMULATE if accumulate else 0)
    worst, dev_limit, ref_ub = {}, 0, dict(limit=0, start_in_solid=0, nan_dir=0, coord_eq_n=0)
    with R.Renderer(W, H, r.window, mode, pflags, device=r.device) as pr:
        pr.upload_blue_noise(noise); pr.upload_world(m, f)
        for i in range(wu + nb):
            if eseq is not None:
                pr.apply_edits(eseq[i % len(eseq)])
            pr.set_frame(np.frombuffer(seq[i % len(seq)], np.uint8)); pr.render_frame()
            dev = dict(hit_voxel=pr.read_buffer(N.BUF_HIT_VOXEL), hit_face=pr.read_buffer(N.BUF_HIT_FACE),
                       hit_word=pr.read_buffer(N.BUF_HIT_WORD), hit_t=pr.read_buffer(N.BUF_HIT_T), output=pr.read_output())
            st = pr.frame_stats()
            ps = parity_stats(dev, kept[i])
            ps["ray_count_mismatch"] = int(st.rays != kept[i]["rays"]); ps["step_count_mismatch"] = int(st.steps != kept[i]["steps"])
            for k, v in ps.items():
                worst[k] = max(worst.get(k, 0), v)
            dev_limit += int(st.limit_exhausted)
            for k in ref_ub:
                ref_ub[k] += kept[i][k]
    worst.update({"frames_compared": wu + nb, "ub_limit_exhausted": dev_limit, "ub_limit_exhausted_oracle": ref_ub["limit"],
                  "ub_start_in_solid": ref_ub["start_in_solid"], "ub_nan_dir": ref_ub["nan_dir"], "ub_coord_eq_n": ref_ub["coord_eq_n"],
                  "tolerance": "integer buffers bit-exact; |dt| <= 1e-4; |dcolour| <= 1 (of 255)"})
    out["parity"] = worst
    return out


def _emit(line):
    """The ONE JSON line goes to the real stdout; everything else libraries print (e.g. NCCL's version banner)
    was redirected to stderr in main()."""
    os.write(_REAL_STDOUT, (json.dumps(line) + "\n").encode())


_REAL_STDOUT = 1


def main():
    global _REAL_STDOUT
    sys.stdout.flush()
    _REAL_STDOUT = os.dup(1)
    os.dup2(2, 1)
    args = parse_args()
    rank = int(os.environ.get("RANK", "0"))
    world_size = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if args.impl == "reference":
        run_reference(args, rank, world_size)
        return
    dist = None
    if world_size > 1:
        import torch
        import torch.distributed as dist
        torch.cuda.set_device(local_rank)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
        try:                                     # run (and first-touch the page-locked frame buffers) on the GPU's own NUMA node
            import pynvml
            pynvml.nvmlInit()
            pynvml.nvmlDeviceSetCpuAffinity(pynvml.nvmlDeviceGetHandleByIndex(local_rank))
        except Exception as e:  # noqa: BLE001
            print(f"rank {rank}: cpu affinity not set: {e}", file=sys.stderr)
    # frames of a camera path are independent units (c1, c2, c4): deal them to the GPUs; c3 (accumulation) and c5 (edit
    # stream) chain frame k+1 to frame k, so only the image-space split applies
    shard = resolve_shard(args)
    line = run_vxrt(args, rank, world_size, local_rank, shard, dist)
    if world_size > 1 and args.shard == "auto" and shard == "frames":
        # the same workload with ONE frame split into row bands (latency of a frame instead of throughput of the path)
        b = run_vxrt(args, rank, world_size, local_rank, "bands", dist)
        if rank == 0:
            line["image_bands"] = {k: b[k] for k in ("value", "unit", "ms_per_step", "frames_per_s", "scaling", "e2e", "gpu_launches",
                                                      "kernel_ms_per_step_by_rank", "run") if k in b}
            line["image_bands"]["parallelism"] = b["config"]["parallelism"]
            for k in ("matches_single_gpu", "rows_differing"):
                if k in b:
                    line["image_bands"][k] = b[k]
            line["gpu_launches"] += b["gpu_launches"]
    if rank == 0:
        _emit(line)
    if dist is not None:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
