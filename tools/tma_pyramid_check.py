"""The TMA-staged resize (tuning key pyr_tma) against the word-load kernel: every level byte for byte. Run on the GPU box."""
import sys, numpy as np
sys.path.insert(0, '/root/repo')
import orb_slam2_2021_b200 as pkg
from orb_slam2_2021_b200 import synth
for name in ("kitti", "euroc", "tum_mono"):
    W, H, nf, nl, sf, ini, mn, fx, bf = synth.CONFIGS[name]
    n = 12
    left, right = synth.stereo_batch(W, H, 4, seed0=5)
    idx = np.arange(n) % 4
    L, R = np.ascontiguousarray(left[idx]), np.ascontiguousarray(right[idx])
    lev = {}
    for mode in (0, 1):
        pkg.set_tuning("pyr_tma", 2 - mode); pkg.set_tuning("pyr_chain", 2)
        b = pkg.StereoFrameBatch(nf, sf, nl, ini, mn, H, W, n, stereo=True)
        b.process(L, R, bf, bf / fx)
        lev[mode] = [b.level(f, s, l) for f in (0, 5, 11) for s in (0, 1) for l in range(nl)]
    ok = all(np.array_equal(a, c) for a, c in zip(lev[0], lev[1]))
    print(name, "tma pyramid equals word-load pyramid:", ok)
for sfx, nlv in ((1.1, 4), (1.5, 4), (2.0, 3), (1.9, 3)):
    lev = {}
    for mode in (0, 1):
        pkg.set_tuning("pyr_tma", 2 - mode); pkg.set_tuning("pyr_chain", 2)
        b = pkg.StereoFrameBatch(500, sfx, nlv, 20, 7, 397, 701, 10, stereo=False)
        b.process(np.stack([synth.synth(701, 397, k) for k in range(10)]), None)
        lev[mode] = [b.level(f, 0, l) for f in (0, 9) for l in range(nlv)]
    print("scale", sfx, all(np.array_equal(a, c) for a, c in zip(lev[0], lev[1])))
