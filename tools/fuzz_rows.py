#!/usr/bin/env python
"""Randomised scenarios for the rows after the front-end (SearchByProjection frame to frame, SearchByProjection against the
local map, ComputeBoW): the CUDA path through the C-ABI against the reference's own object code (oracle/_ref: src/ORBmatcher.cc,
src/Frame.cc, vendored DBoW2). Run on the GPU box:  python tools/fuzz_rows.py [seconds] [seed] [max cases]"""
import os
import sys
import tempfile
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import orb_slam2_2021_b200 as pkg  # noqa: E402
from oracle import oracle as O  # noqa: E402  (the checker's view types)
from oracle import ref as R  # noqa: E402  (the checker)
from tests import bow_util as B  # noqa: E402
from tests import frames_util as U  # noqa: E402

budget = float(sys.argv[1]) if len(sys.argv) > 1 else 60.0
seed0 = int(sys.argv[2]) if len(sys.argv) > 2 else 1
max_cases = int(sys.argv[3]) if len(sys.argv) > 3 else 1 << 30
rng = np.random.default_rng(seed0)
t0, cases, bad = time.time(), [0, 0, 0], 0
tmp = tempfile.mkdtemp()


def report(kind, **kw):
    global bad
    bad += 1
    print("MISMATCH", kind, kw, flush=True)


while time.time() - t0 < budget and sum(cases) < max_cases:
    seed = int(rng.integers(0, 1 << 30))
    kind = int(rng.integers(0, 3))
    if kind == 0:
        n_last = int(rng.integers(1, 4000))
        kw = dict(n_last=n_last, n_extra=int(rng.integers(0, 2000)), W=int(rng.integers(200, 1400)), H=int(rng.integers(150, 800)),
                  tz=float(rng.choice([0.0, 0.5, -0.5, 0.05])), yaw=float(rng.uniform(-0.03, 0.03)),
                  clones=int(rng.integers(0, max(1, n_last // 3))), with_uright=bool(rng.integers(0, 2)))
        if kw["clones"] >= n_last:
            kw["clones"] = 0
        cur, last, has, world, pdesc, obs = U.random_scenario(seed, **kw)
        mono, ori, th = bool(rng.integers(0, 2)), bool(rng.integers(0, 2)), float(rng.choice([7, 15, 3, 30]))
        m = pkg.Matcher(checkOrientation=ori)
        ref, nref = R.search_by_projection(U.view(O.FrameView, cur), U.view(O.FrameView, last), has, world, pdesc, obs, th, mono, ori)
        got, n = m.SearchByProjection(U.view(pkg.FrameView, cur), U.view(pkg.FrameView, last), pkg.MapPointView(has, world, pdesc, obs), th, mono)
        if n != nref or not np.array_equal(got, ref):
            report("search_by_projection", seed=seed, mono=mono, ori=ori, th=th, **kw)
    elif kind == 1:
        n_points = int(rng.integers(1, 3000))
        kw = dict(n_points=n_points, n_extra=int(rng.integers(0, 1500)), W=int(rng.integers(200, 1400)), H=int(rng.integers(150, 800)),
                  clones=int(rng.integers(0, max(1, n_points // 3))), with_uright=bool(rng.integers(0, 2)))
        if kw["clones"] >= n_points:
            kw["clones"] = 0
        F, occupied, valid, proj, level, vc, pdesc, obs = U.local_map_scenario(seed, **kw)
        th, ratio = float(rng.choice([1, 3, 5])), float(rng.choice([0.8, 0.9, 0.6]))
        occ = occupied if rng.integers(0, 2) else None
        m = pkg.Matcher()
        ref, nref = R.search_local_points(U.view(O.FrameView, F), occ, valid, proj, level, vc, pdesc, obs, th, ratio)
        got, n = m.SearchLocalPoints(U.view(pkg.FrameView, F), pkg.TrackPointView(valid, proj, level, vc, pdesc, obs), th, occ, ratio)
        if n != nref or not np.array_equal(got, ref):
            report("search_local_points", seed=seed, th=th, ratio=ratio, occ=occ is not None, **kw)
    else:
        k, L = int(rng.integers(2, 12)), int(rng.integers(2, 6))
        if k ** L > 200000:
            continue
        ragged = bool(rng.integers(0, 2))
        arrays = B.synthetic_vocabulary(k, L, seed=seed % 100000, ragged=ragged)
        path = os.path.join(tmp, "voc.txt")
        R.write_vocabulary_text(path, k, L, *arrays)
        rv, gv = R.RefVocabulary(path), pkg.ORBVocabulary(L, *arrays)
        for n in (int(rng.integers(0, 40)), int(rng.integers(40, 3000))):
            d = B.features_near_words(arrays[2], arrays[1], n, seed=seed % 9973 + n) if n else np.zeros((0, 32), np.uint8)
            a, b = gv.transform(d, 4), rv.transform(d)
            if not all(x.dtype == y.dtype and np.array_equal(x, y) for x, y in zip(a, b)):
                report("compute_bow", seed=seed, k=k, L=L, ragged=ragged, n=n)
        del rv, gv
    cases[kind] += 1
print(f"{cases[0]} frame-to-frame searches, {cases[1]} local-map searches, {cases[2]} vocabularies in {time.time() - t0:.0f} s, {bad} mismatches")
sys.exit(1 if bad else 0)
