#!/usr/bin/env python
"""CPU-only companion of tools/fuzz_vs_reference.py / fuzz_rows.py: the oracle (restatement) against the reference's own object
code (oracle/_ref) on random configurations -- no GPU needed, so it can run for as long as one likes.
  python tools/fuzz_oracle_cpu.py [seconds] [seed] [max cases]"""
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import oracle as O  # noqa: E402
from oracle import ref as R  # noqa: E402
from tools import fuzz_images  # noqa: E402
from tests import frames_util as U  # noqa: E402

budget = float(sys.argv[1]) if len(sys.argv) > 1 else 60.0
rng = np.random.default_rng(int(sys.argv[2]) if len(sys.argv) > 2 else 1)
max_cases = int(sys.argv[3]) if len(sys.argv) > 3 else 1 << 30
t0, cases, bad, skipped = time.time(), [0, 0, 0], 0, {}


def same(a, b):
    return len(a[0]) == len(b[0]) and a[0].tobytes() == b[0].tobytes() and np.array_equal(a[1], b[1])


while time.time() - t0 < budget and sum(cases) < max_cases:
    seed = int(rng.integers(0, 1 << 30))
    kind = int(rng.choice([0, 0, 0, 0, 1, 2]))
    if kind == 0:
        sf = float(rng.choice([1.05, 1.1, 1.2, 1.2, 1.3, 1.5, 1.8, 2.0]))
        nl, nf = int(rng.integers(1, 13)), int(rng.choice([rng.integers(20, 3000), rng.integers(3000, 9000)]))
        W, H = int(rng.choice([rng.integers(160, 1400), rng.integers(1400, 2200)])), int(rng.integers(160, 1100))
        ini, mn = int(rng.integers(1, 100)), int(rng.integers(1, 60))
        style = str(rng.choice(fuzz_images.STYLES))
        left, right = fuzz_images.stereo_pair(style, W, H, seed)
        try:
            o, o_r = O.OracleExtractor(nf, sf, nl, ini, mn), O.OracleExtractor(nf, sf, nl, ini, mn)  # each keeps its pyramid
            ol, orr = o(left), o_r(right)
        except (ValueError, RuntimeError) as e:
            skipped[str(e)] = skipped.get(str(e), 0) + 1  # geometry the reference cannot handle either (a level under 62 px)
            continue
        r_l, r_r = R.RefExtractor(nf, sf, nl, ini, mn), R.RefExtractor(nf, sf, nl, ini, mn)
        wl, wr = r_l(left), r_r(right)
        bf, fx = float(rng.choice([40.0, 47.9, 386.1])), float(rng.choice([300.0, 500.0, 718.0]))
        mb = float(rng.choice([bf / fx, bf / fx, 0.0]))  # 0: the reference reads mb before setting it (maxD = inf)
        wu, wd = R.stereo_match(r_l, r_r, wl[0], wl[1], wr[0], wr[1], bf, mb)
        ou, od = O.stereo_match(o, o_r, ol[0], ol[1], orr[0], orr[1], bf, mb)
        if not (same(ol, wl) and same(orr, wr) and ou.tobytes() == wu.tobytes() and od.tobytes() == wd.tobytes()):
            bad += 1
            print("MISMATCH extract/stereo", dict(W=W, H=H, nf=nf, sf=sf, nl=nl, ini=ini, mn=mn, seed=seed, style=style, bf=bf, mb=mb), flush=True)
    elif kind == 1:
        n_last = int(rng.integers(1, 4000))
        kw = dict(n_last=n_last, n_extra=int(rng.integers(0, 2000)), W=int(rng.integers(200, 1400)), H=int(rng.integers(150, 800)),
                  tz=float(rng.choice([0.0, 0.5, -0.5, 0.05])), yaw=float(rng.uniform(-0.03, 0.03)),
                  clones=int(rng.integers(0, max(1, n_last // 3))), with_uright=bool(rng.integers(0, 2)))
        if kw["clones"] >= n_last:
            kw["clones"] = 0
        cur, last, has, world, pdesc, obs = U.random_scenario(seed, **kw)
        mono, ori, th = bool(rng.integers(0, 2)), bool(rng.integers(0, 2)), float(rng.choice([7, 15, 3, 30]))
        a = R.search_by_projection(U.view(O.FrameView, cur), U.view(O.FrameView, last), has, world, pdesc, obs, th, mono, ori)
        b = O.search_by_projection(U.view(O.FrameView, cur), U.view(O.FrameView, last), has, world, pdesc, obs, th, mono, ori)
        if a[1] != b[1] or not np.array_equal(a[0], b[0]):
            bad += 1
            print("MISMATCH search_by_projection", dict(seed=seed, mono=mono, ori=ori, th=th, **kw), flush=True)
    else:
        n_points = int(rng.integers(1, 3000))
        kw = dict(n_points=n_points, n_extra=int(rng.integers(0, 1500)), W=int(rng.integers(200, 1400)), H=int(rng.integers(150, 800)),
                  clones=int(rng.integers(0, max(1, n_points // 3))), with_uright=bool(rng.integers(0, 2)))
        if kw["clones"] >= n_points:
            kw["clones"] = 0
        F, occupied, valid, proj, level, vc, pdesc, obs = U.local_map_scenario(seed, **kw)
        th, ratio = float(rng.choice([1, 3, 5])), float(rng.choice([0.8, 0.9, 0.6]))
        occ = occupied if rng.integers(0, 2) else None
        a = R.search_local_points(U.view(O.FrameView, F), occ, valid, proj, level, vc, pdesc, obs, th, ratio)
        b = O.search_local_points(U.view(O.FrameView, F), occ, valid, proj, level, vc, pdesc, obs, th, ratio)
        if a[1] != b[1] or not np.array_equal(a[0], b[0]):
            bad += 1
            print("MISMATCH search_local_points", dict(seed=seed, th=th, ratio=ratio, occ=occ is not None, **kw), flush=True)
    cases[kind] += 1
print(f"{cases[0]} extract+stereo, {cases[1]} frame-to-frame, {cases[2]} local-map cases in {time.time() - t0:.0f} s, {bad} mismatches")
print("skipped:", skipped)
sys.exit(1 if bad else 0)
