"""Where the FULL-lighting trace pass loses its lanes — a CPU model, not a test (pytest does not collect it).

    python tests/analysis_divergence.py [c3|c4]

The oracle logs the march length of each of a pixel's five rays (orc_set_ray_log).  The trace kernel (trace_fast.cu)
runs a warp of 8x4 pixels through five rounds (primary, sun1, dif1, sun2, dif2); a round's march loop runs until the
warp's longest ray of that round ends, so a round costs max(len) warp-steps for sum(len) useful lane-steps.  This
script reproduces that cost from the logged lengths (the total must land on ncu's lanes-per-instruction of the loop:
13.9 of 32 on C3) and then re-costs the same rays under the regroupings DESIGN.md §4.1 discusses, so that the next
kernel is chosen from numbers.  Only lane-steps of the march loop are modelled (70 % of the pass's instructions).
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import orc  # noqa: E402
from raytrace_b200 import scenes  # noqa: E402

NAMES = ["primary", "sun1", "dif1", "sun2", "dif2"]


def ray_lengths(workload):
    scene, W, H = {"c3": ("c2", 1920, 1080), "c4": ("c4", 3840, 2160)}[workload]
    dims, kind, seed = scenes.WORLDS[scene]
    m = orc.generate_world(*dims, kind, seed)
    m, f = orc.build_minefield(m)
    cam = scenes.camera(dims)
    blk = orc.fill_frame(cam["origin"], cam["heading"], cam["pitch"], cam["sun_angle"], 1)
    log = np.zeros((H, W, 5), np.uint16)
    orc.lib().orc_set_ray_log.argtypes = [orc.C.c_void_p]
    orc.lib().orc_set_ray_log(orc._p(log))
    try:
        orc.render(m, f, orc.blue_noise(), blk, orc.MODE_FULL, W, H)
    finally:
        orc.lib().orc_set_ray_log(None)
    return log


def warps(log):
    """[n_warps, 32, 5]: the kernel's warp = 8x4 pixels."""
    H, W, _ = log.shape
    Hp, Wp = (H + 3) // 4 * 4, (W + 7) // 8 * 8
    pad = np.zeros((Hp, Wp, 5), log.dtype)
    pad[:H, :W] = log
    return pad.reshape(Hp // 4, 4, Wp // 8, 8, 5).transpose(0, 2, 1, 3, 4).reshape(-1, 32, 5).astype(np.int64)


def packed_rounds(lengths):
    """Rays of one group (any number) dealt 32 at a time, longest first: warp-steps = sum of each batch's max."""
    s = -np.sort(-lengths, axis=1)
    return s[:, ::32].sum(axis=1)


def main():
    wl = sys.argv[1] if len(sys.argv) > 1 else "c3"
    log = ray_lengths(wl)
    w = warps(log)
    useful = w.sum()
    print(f"{wl}: {np.count_nonzero(log)} rays, {useful} march steps, mean {useful / np.count_nonzero(log):.1f} per ray")
    for k, n in enumerate(NAMES):
        v = log[..., k][log[..., k] > 0]
        print(f"  {n:8s} rays {v.size:8d}  mean {v.mean():6.1f}  p50 {np.percentile(v, 50):5.0f}  p90 {np.percentile(v, 90):5.0f}  "
              f"p99 {np.percentile(v, 99):5.0f}  max {v.max()}")

    def report(name, warp_steps):
        print(f"  {name:72s} {warp_steps / 1e6:8.2f} M warp-steps  {useful / warp_steps:5.1f} lanes  x{base / warp_steps:4.2f}")

    per_round = w.max(axis=1)                       # [n_warps, 5]
    base = per_round.sum()
    print("march-loop cost (warp-steps), lanes = useful lane-steps / warp-steps:")
    report("as built: five rounds per 8x4-pixel warp", base)
    for k, n in enumerate(NAMES):
        lanes = w[..., k].sum() / max(per_round[:, k].sum(), 1)
        print(f"      round {n:8s} {per_round[:, k].sum() / 1e6:8.2f} M warp-steps ({100 * per_round[:, k].sum() / base:4.1f} %), {lanes:5.1f} lanes")
    # (a) level-2 rays of a warp packed into its idle lanes, longest first (needs the lengths): sun2 and dif2 of the
    #     survivors share one round when they fit
    lvl2 = np.concatenate([w[..., 3], w[..., 4]], axis=1)
    report("(a) sun2 + dif2 of a warp packed into full rounds, longest rays first (needs the lengths)", per_round[:, :3].sum() + packed_rounds(lvl2).sum())
    # (a') what trace_fast.cu (E16, VXRT_FLAG_RAY_HANDOFF) does: owners in lane order pair with the finished lanes; the
    #      owners left without a partner still march dif2 in a fifth round
    own = w[..., 3] > 0
    n_own = own.sum(axis=1)
    pairs = np.minimum(n_own, 32 - n_own)
    rank = np.cumsum(own, axis=1) - 1
    helped = own & (rank < pairs[:, None])
    r4 = np.maximum(w[..., 3].max(axis=1), np.where(helped, w[..., 4], 0).max(axis=1))
    r5 = np.where(own & ~helped, w[..., 4], 0).max(axis=1)
    report("(a') as implemented (E16): lane-order pairing, unpaired owners march dif2 in a fifth round", per_round[:, :3].sum() + r4.sum() + r5.sum())
    # (b) both levels packed: sun1 + dif1 too (only helps warps with <= 16 ground pixels)
    lvl1 = np.concatenate([w[..., 1], w[..., 2]], axis=1)
    report("(b) (a) and sun1 + dif1 packed the same way", per_round[:, 0].sum() + packed_rounds(lvl1).sum() + packed_rounds(lvl2).sum())
    # (c) CTA-wide pool (4 warps = 16x8 pixels): every round's rays sorted by length and dealt to the CTA's warps
    n4 = w.shape[0] // 4 * 4
    # CTA = 2x2 warps; warps() orders warps row-major over (H/4, W/8): pair rows and columns
    H4, W8 = (log.shape[0] + 3) // 4, (log.shape[1] + 7) // 8
    g = w[: H4 * W8].reshape(H4, W8, 32, 5)
    g = g[: H4 // 2 * 2, : W8 // 2 * 2].reshape(H4 // 2, 2, W8 // 2, 2, 32, 5).transpose(0, 2, 1, 3, 4, 5).reshape(-1, 128, 5)
    scale = w.shape[0] / (g.shape[0] * 4)
    cta = sum(packed_rounds(g[..., k]).sum() for k in range(5)) * scale
    report("(c) per round, a CTA's 128 rays sorted by length (oracle knowledge: upper bound of any binning)", cta)
    cta2 = (packed_rounds(g[..., 0]).sum() + packed_rounds(np.concatenate([g[..., 1], g[..., 2]], axis=1)).sum() +
            packed_rounds(np.concatenate([g[..., 3], g[..., 4]], axis=1)).sum()) * scale
    report("(d) (c) with both rays of a level in one pool", cta2)
    # (g), (h): what a CTA pool can do WITHOUT knowing the lengths: live rays of a level compacted in pixel order and dealt
    #     32 at a time (g); the same after every ray has marched its first K steps in place (h) - half of the sun and
    #     second-bounce rays end within two steps, so the peel removes them before the pool is formed
    def unsorted_rounds(lengths):                  # [n, m] -> live rays compacted in order, batches of 32, max per batch
        tot = 0
        for row in lengths:
            v = row[row > 0]
            if v.size:
                padn = (-v.size) % 32
                tot += np.pad(v, (0, padn)).reshape(-1, 32).max(axis=1).sum()
        return tot
    pools = [g[..., 0], np.concatenate([g[..., 1], g[..., 2]], axis=1), np.concatenate([g[..., 3], g[..., 4]], axis=1)]
    sample = slice(None, None, max(1, g.shape[0] // 3000))
    sc = scale * g.shape[0] / len(range(*sample.indices(g.shape[0])))
    report("(g) per level, a CTA's live rays compacted in pixel order (no knowledge of lengths), primary as built",
           per_round[:, 0].sum() + sum(unsorted_rounds(p_[sample]) for p_ in pools[1:]) * sc)
    for K in (2, 4):
        tot = per_round[:, 0].sum()
        for lv in ((1, 2), (3, 4)):
            for k in lv:                            # the peel: K steps (or fewer) of the round as built
                tot += np.minimum(per_round[:, k], K).sum()
            rest = np.concatenate([np.maximum(g[..., k] - K, 0) for k in lv], axis=1)
            tot += unsorted_rounds(rest[sample]) * sc
        report(f"(h) (g) after the first {K} steps of every secondary ray are marched in place", tot)
    # (e) lane refill inside the round (persistent lanes): the bound is one warp-step per 32 lane-steps
    report("(e) perfect refill (every warp-step carries 32 rays): the bound", useful / 32)
    # (f) refill at a threshold: a warp leaves the loop when fewer than T lanes march, refills, and re-enters;
    #     modelled per warp over its whole ray list: rays sorted longest first run in batches, a batch ends when
    #     fewer than T of its rays are still marching and the survivors join the next batch
    for T in (8, 16, 24):
        tot = 0
        rays = w.reshape(w.shape[0], -1)
        for r in rays[:: max(1, rays.shape[0] // 4000)]:               # a sample of warps
            live = []
            queue = sorted((int(x) for x in r if x > 0), reverse=True)
            steps = 0
            while queue or live:
                while queue and len(live) < 32:
                    live.append(queue.pop())
                if not live:
                    break
                live.sort()
                # march until fewer than T remain (or nothing left to refill with)
                k = max(0, len(live) - T) if queue else len(live) - 1
                d = live[k]
                steps += d
                live = [x - d for x in live if x - d > 0]
            tot += steps
        tot *= rays.shape[0] / len(rays[:: max(1, rays.shape[0] // 4000)])
        report(f"(f) refill when fewer than {T} lanes march (ray order ignored; each refill costs a set-up at partial lanes)", tot)


if __name__ == "__main__":
    main()
