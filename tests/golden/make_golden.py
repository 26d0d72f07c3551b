#!/usr/bin/env python
"""Generates tests/golden/*.npz by running THE REFERENCE'S OWN SHADER TEXT (oracle/_ref/libref.so, built by
oracle/make_ref.py from /root/reference/shaders/glsl/*.comp) on seeded inputs.  Run in the build container,
where /root/reference exists; the fixtures travel to the GPU box, the reference does not.

    python oracle/make_ref.py && python tests/golden/make_golden.py

Inputs are reproducible from seeds alone (world: orc.generate_world + pack_into; noise: orc.blue_noise();
frame block: stored in the fixture), so only the reference OUTPUTS are stored.
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))

from oracle import orc  # noqa: E402
from raytrace_b200 import scenes  # noqa: E402

# name: (width, height, camera index of an 8-frame path, seed, window centre lr, pong_swap)
CASES = {
    "ref_mid_a": (96, 64, 0, 5, (0, 0, 0), False),
    "ref_mid_b": (80, 136, 5, 777, (16, 0, -16), True),      # lr != 0, faithful pong-set bug, rows past 128
}


def main():
    dims, kind, seed = scenes.WORLDS["mid"]
    m, f = orc.build_minefield(orc.generate_world(*dims, kind, seed))
    noise = orc.blue_noise()
    for name, (W, H, k, fseed, lr, swap) in CASES.items():
        cam = scenes.camera(dims, k, 8)
        blk = orc.fill_frame(cam["origin"], cam["heading"], cam["pitch"], cam["sun_angle"], fseed, lr)
        fr = orc.ref_render(m, f, noise, blk, W, H, pong_swap=swap)
        np.savez_compressed(os.path.join(HERE, name + ".npz"), frame_block=blk, pong_swap=np.array(swap),
                            lighting_raw=fr.lighting_raw, lighting=fr.lighting, depth=fr.depth, normal=fr.normal,
                            albedo=fr.albedo, emission=fr.emission, fog=fr.fog, output=fr.output)
        print(name, os.path.getsize(os.path.join(HERE, name + ".npz")), "bytes")


if __name__ == "__main__":
    main()
