#!/usr/bin/env python
"""Measurement for the N1 row (SearchByProjection, frame to frame): frame pairs/s through the C-ABI with HOST
buffers (packed H2D + 3 kernels + D2H inside the timed region), the kernel-only time from CUDA events is not
separable here (the matcher owns its stream), so the per-kernel split comes from the ncu launch list.
Beside it: the oracle (C++ port of the reference loop) on all host threads, and single-pair latency of both.

  python tools/bench_frame.py [--pairs 256] [--points 2000] > gpurun_out/frame_bench.json
"""
import argparse
import json
import os
import sys
import time
from concurrent.futures import ThreadPoolExecutor

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import orb_slam2_2021_b200 as pkg  # noqa: E402
from oracle import oracle as O  # noqa: E402  (CPU baseline only)
from tests import frames_util as U  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--pairs", type=int, default=256)
    ap.add_argument("--points", type=int, default=2000)
    ap.add_argument("--reps", type=int, default=10)
    a = ap.parse_args()
    distinct = 8
    scen = [U.random_scenario(100 + k, n_last=a.points, n_extra=a.points // 10, clones=a.points // 20, W=1241, H=376,
                              tz=[0.0, 0.5, -0.5][k % 3]) for k in range(distinct)]
    gv = [(U.view(pkg.FrameView, s[0]), U.view(pkg.FrameView, s[1]), pkg.MapPointView(s[2], s[3], s[4], s[5])) for s in scen]
    ov = [(U.view(O.FrameView, s[0]), U.view(O.FrameView, s[1])) + tuple(s[2:]) for s in scen]
    idx = [k % distinct for k in range(a.pairs)]
    m = pkg.Matcher()
    cur, last, pts = [gv[k][0] for k in idx], [gv[k][1] for k in idx], [gv[k][2] for k in idx]
    res = m.SearchByProjection(cur, last, pts, 7)
    for k in range(distinct):  # parity of what is being timed
        ref, nref = O.search_by_projection(*ov[k], 7, False, True)
        assert res[k][1] == nref and np.array_equal(res[k][0], ref)
    for _ in range(3):
        m.SearchByProjection(cur, last, pts, 7)
    t0 = time.perf_counter()
    for _ in range(a.reps):
        m.SearchByProjection(cur, last, pts, 7)
    gpu_batch = a.pairs * a.reps / (time.perf_counter() - t0)
    ts = []
    for r in range(40):
        t0 = time.perf_counter()
        m.SearchByProjection(*gv[r % distinct], 7)
        ts.append(time.perf_counter() - t0)
    gpu_single_ms = float(np.median(ts[5:]) * 1e3)
    threads = len(os.sched_getaffinity(0))

    def one(k):
        return O.search_by_projection(*ov[k], 7, False, True)[1]

    t0 = time.perf_counter()
    with ThreadPoolExecutor(threads) as ex:
        list(ex.map(one, idx * 2))
    cpu_batch = 2 * a.pairs / (time.perf_counter() - t0)
    ts = []
    for r in range(20):
        t0 = time.perf_counter()
        one(r % distinct)
        ts.append(time.perf_counter() - t0)
    cpu_single_ms = float(np.median(ts[3:]) * 1e3)
    nm = float(np.mean([r[1] for r in res]))
    print(json.dumps({"metric": "SearchByProjection frame pairs/s", "pairs": a.pairs, "points_per_frame": a.points,
                      "matches_per_pair": nm, "gpu_pairs_per_s_e2e": gpu_batch, "gpu_single_pair_ms": gpu_single_ms,
                      "cpu_pairs_per_s": cpu_batch, "cpu_threads": threads, "cpu_single_pair_ms": cpu_single_ms,
                      "cpu_kind": "port (oracle/frame.cpp)", "h2d_bytes_per_pair": int(a.points * 1.1 * 64 + a.points * 74),
                      "d2h_bytes_per_pair": int(a.points * 1.1 * 4 + 4)}))


if __name__ == "__main__":
    main()
