#!/usr/bin/env python
"""Random extractor parameters and image sizes: the CUDA path (single handle = one-launch pyramid + CUDA graph; batch engine =
TMA-staged per-level pyramid; orbx_stereo_frame) against the reference's own object code (oracle/_ref), every keypoint field and
descriptor byte and every stereo float. Run on the GPU box:  python tools/fuzz_vs_reference.py [seconds] [seed] [max cases]"""
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import orb_slam2_2021_b200 as pkg  # noqa: E402
from oracle import ref as R  # noqa: E402  (the checker)
from tools import fuzz_images  # noqa: E402

budget = float(sys.argv[1]) if len(sys.argv) > 1 else 60.0
rng = np.random.default_rng(int(sys.argv[2]) if len(sys.argv) > 2 else 1)
max_cases = int(sys.argv[3]) if len(sys.argv) > 3 else 1 << 30
t0, cases, bad = time.time(), 0, 0


def same(a, b):
    return len(a[0]) == len(b[0]) and a[0].tobytes() == b[0].tobytes() and np.array_equal(a[1], b[1])


while time.time() - t0 < budget and cases < max_cases:
    sf = float(rng.choice([1.05, 1.1, 1.2, 1.2, 1.3, 1.5, 1.8, 2.0]))
    nl = int(rng.integers(1, 13))
    W, H = int(rng.choice([rng.integers(160, 1400), rng.integers(1400, 2200)])), int(rng.integers(160, 1100))
    nf = int(rng.choice([rng.integers(20, 3000), rng.integers(3000, 9000)]))
    ini = int(rng.integers(1, 100))
    mn = int(rng.integers(1, 60))
    try:
        ex = pkg.ORBextractor(nf, sf, nl, ini, mn)
        if pkg.lib().orbx_max_keypoints(ex._h, H, W) < 0:
            continue  # geometry the reference cannot handle either (level below 62 px, portrait level)
    except pkg.OrbxError:
        continue
    seed = int(rng.integers(0, 1 << 30))
    style = str(rng.choice(fuzz_images.STYLES))
    left, right = fuzz_images.stereo_pair(style, W, H, seed)
    r_l, r_r = R.RefExtractor(nf, sf, nl, ini, mn), R.RefExtractor(nf, sf, nl, ini, mn)
    want_l, want_r = r_l(left), r_r(right)
    bf = float(rng.choice([40.0, 47.9, 386.1]))
    mb = float(rng.choice([bf / 500.0, bf / 718.9, bf / 300.0, 0.0]))  # 0: the reference reads mb before setting it (maxD = inf)
    want_u, want_d = R.stereo_match(r_l, r_r, want_l[0], want_l[1], want_r[0], want_r[1], bf, mb)
    ok = same(ex(left), want_l)
    fl, fr = pkg.ORBextractor(nf, sf, nl, ini, mn), pkg.ORBextractor(nf, sf, nl, ini, mn)
    for _ in range(2):
        kl, dl, kr, dr, ur, dp = pkg.stereo_frame(fl, fr, left, right, bf, mb)
        ok = ok and same((kl, dl), want_l) and same((kr, dr), want_r) and ur.tobytes() == want_u.tobytes() and dp.tobytes() == want_d.tobytes()
    n = 5
    b = pkg.StereoFrameBatch(nf, sf, nl, ini, mn, H, W, n, stereo=True)
    out = b.process(np.stack([left] * n), np.stack([right] * n), bf, mb)
    for k in (0, n - 1):
        a, c = int(out["n_left"][k]), int(out["n_right"][k])
        ok = ok and same((out["kps_left"][k, :a], out["desc_left"][k, :a]), want_l) and same((out["kps_right"][k, :c], out["desc_right"][k, :c]), want_r)
        ok = ok and out["u_right"][k, :a].tobytes() == want_u.tobytes() and out["depth"][k, :a].tobytes() == want_d.tobytes()
    cases += 1
    if not ok:
        bad += 1
        print("MISMATCH", dict(W=W, H=H, nf=nf, sf=sf, nl=nl, ini=ini, mn=mn, seed=seed, style=style, bf=bf, mb=mb), flush=True)
print(f"{cases} random configurations in {time.time() - t0:.0f} s, {bad} mismatches")
sys.exit(1 if bad else 0)
