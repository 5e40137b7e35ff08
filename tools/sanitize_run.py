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
# round 2: the one-call stereo frame (direct, graph capture, replay), the resident match, a batch wide enough for the per-level
# TMA-staged pyramid (> 8 images) incl. scale factor 2, the one-launch pyramid forced on a batch, the static-scratch match,
# and the device-resident frame search / BoW
fl, fr = pkg.ORBextractor(nf, sf, nl, ini, mn), pkg.ORBextractor(nf, sf, nl, ini, mn)
for _ in range(3):
    o6 = pkg.stereo_frame(fl, fr, left, right, bf, bf / fx)
print("stereo_frame", len(o6[0]), int((o6[4] >= 0).sum()))
kl, dl = exl(left)
kr, dr = exr(right)
print("resident", int((pkg.compute_stereo_matches_resident(exl, exr, len(kl), len(kr), bf, bf / fx)[0] >= 0).sum()))
print("static", int((pkg.compute_stereo_matches_static(exl, exr, kl, dl, kr, dr, bf, bf / fx)[0] >= 0).sum()))
b6 = pkg.StereoFrameBatch(nf, sf, nl, ini, mn, H, W, 6, stereo=True)
L6, R6 = zip(*[synth.stereo_pair(W, H, 80 + k) for k in range(6)])
out6 = b6.process(np.stack(L6), np.stack(R6), bf, bf / fx)
print("batch of 12 images (TMA pyramid)", out6["n_left"].tolist())
b2 = pkg.StereoFrameBatch(600, 2.0, 3, 20, 7, 397, 701, 10, stereo=False)
print("scale 2.0 batch", b2.process(np.stack([synth.synth(701, 397, k) for k in range(10)]), None)["n_left"].tolist())
pkg.set_tuning("pyr_chain", 1)
print("forced chain", b6.process(np.stack(L6), np.stack(R6), bf, bf / fx)["n_left"].tolist())
pkg.set_tuning("pyr_chain", 0)
I = np.hstack([np.eye(3, dtype=np.float32), np.zeros((3, 1), np.float32)])
sfs = (np.float32(sf) ** np.arange(nl)).astype(np.float32)
n0, n1 = int(out6["n_left"][0]), int(out6["n_left"][1])
dummy = lambda n: pkg.FrameView(np.zeros(n, pkg.KP_DTYPE), np.zeros((n, 32), np.uint8), None, I, fx, fx, W / 2, H / 2, bf, bf / fx, (0.0, float(W), 0.0, float(H)), sfs)  # noqa: E731
z = out6["depth"][1, :n1]
k1 = out6["kps_left"][1, :n1]
zz = np.where(z > 0, z, 1).astype(np.float32)
mp = pkg.MapPointView(z > 0, np.stack([(k1["x"] - W / 2) * zz / fx, (k1["y"] - H / 2) * zz / fx, zz], 1), out6["desc_left"][1, :n1], np.ones(n1, np.uint8))
m = pkg.Matcher()
print("resident search", m.SearchByProjection(dummy(n0), dummy(n1), mp, 15, cur_src=(b6, 0, 0), last_src=(b6, 1, 0))[1])
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import bow_util as B  # noqa: E402
voc = pkg.ORBVocabulary(4, *B.synthetic_vocabulary(10, 4, seed=2))
print("resident bow", [len(t[0]) for t in voc.transform_resident(b6, out6["n_left"][:2])])
