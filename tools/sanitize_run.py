#!/usr/bin/env python
"""Small workload for compute-sanitizer (memcheck / racecheck / initcheck): every kernel of the path once, incl. the FAST
minTh attempt (nearly flat image), the WIDE resize (level ratio above 2) and the wide-cell FAST variant.
  compute-sanitizer --tool racecheck python tools/sanitize_run.py"""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import orb_slam2_2021_b200 as pkg  # noqa: E402
from orb_slam2_2021_b200 import synth  # noqa: E402

W, H, nf, nl, sf, ini, mn, fx, bf = synth.CONFIGS["arducam"]
left, right = synth.stereo_pair(W, H, 5)
exl, exr = pkg.ORBextractor(nf, sf, nl, ini, mn), pkg.ORBextractor(nf, sf, nl, ini, mn)
kl, dl = exl(left)
kr, dr = exr(right)
ur, dp = pkg.compute_stereo_matches(exl, exr, kl, dl, kr, dr, bf, bf / fx)
print("arducam", len(kl), len(kr), int((ur >= 0).sum()))
rng = np.random.default_rng(1)
flat = np.clip(128 + rng.normal(0, 3.0, (H, W)), 0, 255).astype(np.uint8)
flat[100:140, 200:260] += 14
print("flat", len(exl(flat)[0]))
ex2 = pkg.ORBextractor(800, 2.0, 3, 20, 7)
img = synth.stereo_pair(701, 397, 23)[0]
print("scale 2.0", len(ex2(img)[0]))
ex3 = pkg.ORBextractor(500, 1.2, 4, 20, 7)  # 118-px-wide level: one cell column of 59 px -> the 112-byte FAST tile
print("wide cells", len(ex3(synth.stereo_pair(150, 140, 3)[0])[0]))
b = pkg.StereoFrameBatch(nf, sf, nl, ini, mn, H, W, 3, stereo=True)
L3, R3 = zip(*[synth.stereo_pair(W, H, 60 + k) for k in range(3)])
out = b.process(np.stack(L3), np.stack(R3), bf, bf / fx)
print("batch", out["n_left"].tolist())
