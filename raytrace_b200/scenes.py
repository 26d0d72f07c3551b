"""Seeded scenes shared by tests/, bench.py and __graft_entry__.smoke(): BASELINE.json configs C1..C5
scaled by a size argument.  World + camera only; no reference or oracle code is needed to define them."""
import math

# name -> (window dims, generator kind, world seed)
WORLDS = {
    "c1": ((128, 128, 128), 0, 1),            # 128^3 heightmap (BASELINE config 1)
    "c2": ((512, 512, 512), 1, 2),            # 512^3 biome world (configs 2, 3, 5)
    "c4": ((2048, 2048, 256), 0, 4),          # 2048x2048x256 multi-chunk world (config 4)
    "small": ((64, 64, 64), 0, 3),            # single chunk, unit tests
    "mid": ((256, 256, 256), 1, 5),           # reference window size
}


def camera(window, k=0, n=1):
    """Camera k of an n-frame path: stands just inside the -Y face, looks along +Y and down;
    the heading sweeps +-0.5 rad over the path (BASELINE config 2: 64-frame camera path)."""
    nx, ny, nz = window
    sweep = 0.0 if n <= 1 else (k / (n - 1) - 0.5)
    origin = (0.3 + 0.2 * nx * sweep, -ny / 2 + 1.5, 0.3 * nz + 0.25)
    return dict(origin=origin, heading=math.pi / 2 + sweep, pitch=-0.35, sun_angle=0.7)
