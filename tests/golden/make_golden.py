"""Generate the golden vectors under tests/golden/ with the cv2-driven model (pyref.py).

Run here (cv2 4.13.0 present): python tests/golden/make_golden.py
The pixel primitives are real OpenCV calls (resize, FAST, GaussianBlur, fastAtan2, norm), so these
vectors pin the oracle (and through it the CUDA path) to OpenCV's arithmetic. The reference binary
itself cannot be built in this image (no OpenCV C++ headers), see DESIGN.md.
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(HERE, "..", ".."))
from orb_slam2_2021_b200 import synth  # noqa: E402
from tests.golden import pyref  # noqa: E402

CASES = [("tum_mono", 0, False), ("arducam", 1, True), ("kitti", 0, True), ("euroc", 4, True)]


def main():
    for name, seed, stereo in CASES:
        W, H, nf, nl, sf, ini, mn, fx, bf = synth.CONFIGS[name]
        left, right = synth.stereo_pair(W, H, seed)
        exl = pyref.PyRefExtractor(nf, sf, nl, ini, mn)
        kl, dl = exl(left)
        out = dict(config=name, seed=seed, kps_left=kl, desc_left=dl)
        if stereo:
            exr = pyref.PyRefExtractor(nf, sf, nl, ini, mn)
            kr, dr = exr(right)
            ur, dp = pyref.stereo_match(exl, exr, kl, dl, kr, dr, bf, bf / fx)
            ur0, dp0 = pyref.stereo_match(exl, exr, kl, dl, kr, dr, bf, 0.0)
            out.update(kps_right=kr, desc_right=dr, u_right=ur, depth=dp, u_right_mb0=ur0, depth_mb0=dp0)
        path = os.path.join(HERE, f"{name}_seed{seed}.npz")
        np.savez_compressed(path, **out)
        print(path, len(kl), os.path.getsize(path))


if __name__ == "__main__":
    main()
