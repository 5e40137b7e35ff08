#!/usr/bin/env python
"""Measurement for the N2 row (Frame::ComputeBoW): frames/s through the C-ABI with host buffers against the oracle
port on all host threads. Vocabulary: synthetic k=10, L=6 tree (1,111,110 nodes, the shape of ORBvoc.txt, which is
not in the reference checkout).   python tools/bench_bow.py > gpurun_out/bow_bench.json"""
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


def big_vocabulary(k, L, seed):
    rng = np.random.default_rng(seed)
    parent, leaf, desc, weight = [], [], [], []
    ids = np.zeros(1, np.int64)
    d = rng.integers(0, 256, (1, 32), dtype=np.uint8)
    nxt = 1
    for depth in range(1, L + 1):
        n = len(ids) * k
        pd = np.repeat(d, k, axis=0)
        flip = rng.integers(0, 256, (n, 32), dtype=np.uint8) & rng.integers(0, 256, (n, 32), dtype=np.uint8) & rng.integers(0, 256, (n, 32), dtype=np.uint8)
        cd = pd ^ flip
        parent.append(np.repeat(ids, k))
        leaf.append(np.full(n, depth == L, np.uint8))
        desc.append(cd)
        weight.append(rng.uniform(0.1, 12.0, n) if depth == L else np.zeros(n))
        ids = np.arange(nxt, nxt + n)
        nxt += n
        d = cd
    return np.concatenate(parent).astype(np.int32), np.concatenate(leaf), np.concatenate(desc), np.concatenate(weight)


def main():
    frames, npts, reps = 256, 2000, 5
    arrays = big_vocabulary(10, 6, 0)
    gv, ov = pkg.ORBVocabulary(6, *arrays), O.Vocabulary(6, *arrays)
    rng = np.random.default_rng(1)
    leaves = np.flatnonzero(arrays[1])
    sets = []
    for f in range(8):
        d = arrays[2][rng.choice(leaves, npts)].copy()
        d ^= (rng.integers(0, 256, d.shape, dtype=np.uint8) & rng.integers(0, 256, d.shape, dtype=np.uint8) & rng.integers(0, 256, d.shape, dtype=np.uint8)
              & rng.integers(0, 256, d.shape, dtype=np.uint8))
        sets.append(d)
    batch = [sets[f % 8] for f in range(frames)]
    res = gv.transform(batch)
    for f in range(8):
        ref = ov.transform(sets[f])
        assert all(np.array_equal(a, b) for a, b in zip(res[f], ref))
    gv.transform(batch)
    t0 = time.perf_counter()
    for _ in range(reps):
        gv.transform(batch)
    gpu = frames * reps / (time.perf_counter() - t0)
    ts = []
    for r in range(30):
        t0 = time.perf_counter()
        gv.transform(sets[r % 8])
        ts.append(time.perf_counter() - t0)
    threads = len(os.sched_getaffinity(0))
    t0 = time.perf_counter()
    with ThreadPoolExecutor(threads) as ex:
        list(ex.map(lambda f: ov.transform(sets[f % 8]), range(frames)))
    cpu = frames / (time.perf_counter() - t0)
    t1 = []
    for r in range(10):
        t0 = time.perf_counter()
        ov.transform(sets[r % 8])
        t1.append(time.perf_counter() - t0)
    print(json.dumps({"metric": "ComputeBoW frames/s", "vocabulary": "synthetic k=10 L=6, 1,111,110 nodes", "features_per_frame": npts,
                      "frames_per_call": frames, "words_per_frame": int(len(res[0][0])), "gpu_frames_per_s_e2e": gpu,
                      "gpu_single_frame_ms": float(np.median(ts[5:]) * 1e3), "cpu_frames_per_s": cpu, "cpu_threads": threads,
                      "cpu_single_frame_ms": float(np.median(t1[2:]) * 1e3), "cpu_kind": "port (oracle/bow.cpp)"}))


if __name__ == "__main__":
    main()
