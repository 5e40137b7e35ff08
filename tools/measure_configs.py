#!/usr/bin/env python
"""Per-config CPU and GPU numbers for BASELINE.md §3 (CPU-1, CPU-2, CPU-3) next to the CUDA path.

Run on the GPU box:  python tools/measure_configs.py > gpurun_out/configs.json
Each leg is bounded (a few seconds to ~20 s) so the whole script ends within a few minutes.
The oracle is executed here only as the reported CPU baseline (the same role as bench.py's
`cpu_baseline` leg); nothing on the product path imports it.
"""
import json
import os
import statistics
import sys
import time
from concurrent.futures import ThreadPoolExecutor

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from orb_slam2_2021_b200 import synth  # noqa: E402

# (name, stereo, distinct frames, frames for CPU-1, frames for CPU-2, GPU batch)
PLAN = [
    ("tum_mono", False, 8, 16, 96, 256),
    ("kitti", True, 8, 16, 96, 256),
    ("euroc", True, 8, 16, 96, 256),
    ("uhd", True, 2, 4, 24, 16),
    ("arducam", True, 8, 16, 96, 256),
]


def host_threads():
    try:
        return len(os.sched_getaffinity(0))
    except AttributeError:
        return os.cpu_count() or 1


def frames_for(name, stereo, distinct):
    W, H = synth.CONFIGS[name][:2]
    L, R = synth.stereo_batch(W, H, distinct, seed0=1000)
    return L, (R if stereo else None)


def cpu1(O, cfg, L, R, n):
    """One frame at a time: extract L || extract R on two threads, then the stereo match."""
    W, H, nf, nl, sf, ini, mn, fx, bf = cfg
    el, er = O.OracleExtractor(nf, sf, nl, ini, mn), O.OracleExtractor(nf, sf, nl, ini, mn)
    ts = []
    with ThreadPoolExecutor(2) as ex:
        for i in range(n + 2):
            k = i % len(L)
            t0 = time.perf_counter()
            if R is None:
                el(L[k])
            else:
                fl, fr = ex.submit(el, L[k]), ex.submit(er, R[k])
                (kl, dl), (kr, dr) = fl.result(), fr.result()
                O.stereo_match(el, er, kl, dl, kr, dr, bf, bf / fx)
            if i >= 2:
                ts.append(time.perf_counter() - t0)
    return statistics.median(ts) * 1e3


def cpu2(O, cfg, L, R, n, threads):
    """Frame-parallel over `threads` host threads, one frame (pair) per task."""
    W, H, nf, nl, sf, ini, mn, fx, bf = cfg

    def one(i):
        k = i % len(L)
        el = O.OracleExtractor(nf, sf, nl, ini, mn)
        kl, dl = el(L[k])
        if R is not None:
            er = O.OracleExtractor(nf, sf, nl, ini, mn)
            kr, dr = er(R[k])
            O.stereo_match(el, er, kl, dl, kr, dr, bf, bf / fx)
        return len(kl)

    t0 = time.perf_counter()
    with ThreadPoolExecutor(threads) as ex:
        list(ex.map(one, range(n)))
    return n / (time.perf_counter() - t0)


def cpu3(cfg, img, nthreads):
    """OpenCV's own kernels through cv2: resize chain, clone+GaussianBlur x levels, per-cell FAST with fallback."""
    import cv2

    cv2.setNumThreads(nthreads)
    W, H, nf, nl, sf, ini, mn = cfg[:7]
    from bench import level_sizes  # same float32 rounding as the extractor

    lv = level_sizes(W, H, nl, sf)

    def best(fn, reps=5):
        b = 1e9
        for _ in range(reps):
            t0 = time.perf_counter()
            fn()
            b = min(b, time.perf_counter() - t0)
        return b * 1e3

    pyr = [img]

    def resize_chain():
        del pyr[1:]
        for w, h in lv[1:]:
            pyr.append(cv2.resize(pyr[-1], (w, h), interpolation=cv2.INTER_LINEAR))

    t_resize = best(resize_chain)

    def blur_all():
        for p in pyr:
            cv2.GaussianBlur(p.copy(), (7, 7), 2, 2, borderType=cv2.BORDER_REFLECT_101)

    t_blur = best(blur_all)
    f_ini = cv2.FastFeatureDetector_create(ini, True)
    f_min = cv2.FastFeatureDetector_create(mn, True)

    def fast_cells():
        # the reference's 30x30 cell loop (src/ORBextractor.cc:778-846): ini threshold, min threshold on empty cells
        for p in pyr:
            h, w = p.shape
            x0, y0, x1, y1 = 13, 13, w - 13, h - 13  # EDGE_THRESHOLD - 3 borders
            ncols, nrows = max((x1 - x0) // 30, 1), max((y1 - y0) // 30, 1)
            wc, hc = -(-(x1 - x0) // ncols), -(-(y1 - y0) // nrows)
            for i in range(nrows):
                iy = y0 + i * hc
                if iy >= y1 - 3:
                    continue
                ey = min(iy + hc + 6, y1)
                for j in range(ncols):
                    ix = x0 + j * wc
                    if ix >= x1 - 6:
                        continue
                    roi = p[iy:ey, ix:min(ix + wc + 6, x1)]
                    if not f_ini.detect(roi):
                        f_min.detect(roi)

    t_fast = best(fast_cells, reps=2)
    return {"threads": nthreads, "resize_ms": round(t_resize, 3), "blur_ms": round(t_blur, 3), "fast_cells_ms": round(t_fast, 2)}


def gpu_legs(cfg, stereo, L, R, batch):
    import torch

    from orb_slam2_2021_b200 import ORBextractor, StereoFrameBatch, compute_stereo_matches
    from orb_slam2_2021_b200.extractor import PinnedArray

    W, H, nf, nl, sf, ini, mn, fx, bf = cfg
    out = {}
    # single frame through the reference-shaped API (two handles, host buffers in and out)
    el, er = ORBextractor(nf, sf, nl, ini, mn), ORBextractor(nf, sf, nl, ini, mn)
    ts = []
    for i in range(24):
        k = i % len(L)
        t0 = time.perf_counter()
        kl, dl = el(L[k])
        if stereo:
            kr, dr = er(R[k])
            compute_stereo_matches(el, er, kl, dl, kr, dr, bf, bf / fx)
        if i >= 4:
            ts.append(time.perf_counter() - t0)
    out["single_frame_ms"] = round(statistics.median(ts) * 1e3, 3)
    if stereo:  # the same frame through ONE call (orbx_stereo_frame: both images per launch, CUDA graph, one round trip)
        from orb_slam2_2021_b200 import stereo_frame

        ts = []
        for i in range(64):
            k = i % len(L)
            t0 = time.perf_counter()
            stereo_frame(el, er, L[k], R[k], bf, bf / fx)
            if i >= 4:
                ts.append(time.perf_counter() - t0)
        out["stereo_frame_call_ms"] = round(statistics.median(ts) * 1e3, 3)
    del el, er
    # the same frame through the batch engine with a batch of one: both images share every launch, one round trip
    one = StereoFrameBatch(nf, sf, nl, ini, mn, H, W, 1, stereo=stereo)
    p1l, p1r = PinnedArray((1, H, W), np.uint8), PinnedArray((1, H, W), np.uint8)
    o1 = one.alloc_outputs(1, pinned=True)
    ts = []
    for i in range(44):
        k = i % len(L)
        p1l.array[0] = L[k]
        if stereo:
            p1r.array[0] = R[k]
        t0 = time.perf_counter()
        one.process(p1l.array, p1r.array if stereo else None, bf if stereo else 0.0, bf / fx if stereo else 1.0, out=o1)
        if i >= 4:
            ts.append(time.perf_counter() - t0)
    out["single_frame_batch_engine_ms"] = round(statistics.median(ts) * 1e3, 3)
    del one
    # batch engine: device-resident and end to end
    idx = np.arange(batch) % len(L)
    bl = np.ascontiguousarray(L[idx])
    br = np.ascontiguousarray(R[idx]) if stereo else None
    eng = StereoFrameBatch(nf, sf, nl, ini, mn, H, W, batch, stereo=stereo)
    eng.upload(bl, br)
    mbf, mb = (bf, bf / fx) if stereo else (0.0, 1.0)
    for _ in range(3):
        eng.run(batch, mbf, mb)
    reps = 5
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(reps):
        eng.run(batch, mbf, mb, sync=False)
    torch.cuda.synchronize()  # device-wide: covers the engine's own streams
    dt = time.perf_counter() - t0
    lib_sync = eng.download(batch)
    out["batch"] = batch
    out["device_resident_fps"] = round(batch * reps / dt, 1)
    outs = eng.alloc_outputs(batch, pinned=True)
    pl = PinnedArray(bl.shape, np.uint8)
    pl.array[...] = bl
    pr = None
    if stereo:
        pr = PinnedArray(br.shape, np.uint8)
        pr.array[...] = br
    for _ in range(2):
        eng.process(pl.array, pr.array if stereo else None, mbf, mb, out=outs)
    t0 = time.perf_counter()
    for _ in range(reps):
        eng.process(pl.array, pr.array if stereo else None, mbf, mb, out=outs)
    dt = time.perf_counter() - t0
    out["e2e_fps"] = round(batch * reps / dt, 1)
    out["keypoints_per_image"] = round(float(lib_sync["n_left"].mean()), 1)
    if stereo:
        out["matches_per_frame"] = round(float((lib_sync["u_right"] >= 0).sum()) / batch, 1)
    return out


def main():
    res = {"host_threads": host_threads(), "configs": {}}
    try:
        for line in open("/proc/cpuinfo"):
            if line.startswith("model name"):
                res["cpu_model"] = line.split(":", 1)[1].strip()
                break
    except OSError:
        pass
    from oracle import oracle as O  # CPU baseline only
    from oracle import ref as R

    if R.available():  # the reference's own ORBextractor.cc / Frame.cc object code (oracle/_ref) instead of the port
        class O:  # noqa: N801 - same two names the CPU legs use
            OracleExtractor = R.RefExtractor
            stereo_match = staticmethod(R.stereo_match)
        res["cpu_kind"] = "reference (oracle/_ref)"
    else:
        res["cpu_kind"] = "port (oracle)"

    import torch

    have_gpu = torch.cuda.is_available()
    only = sys.argv[1:] or None
    for name, stereo, distinct, n1, n2, batch in PLAN:
        if only and name not in only:
            continue
        cfg = synth.CONFIGS[name]
        L, R = frames_for(name, stereo, distinct)
        r = {"W": cfg[0], "H": cfg[1], "nfeatures": cfg[2], "stereo": stereo}
        r["cpu1_ms_per_frame_2thr"] = round(cpu1(O, cfg, L, R, n1), 2)
        r["cpu2_fps_all_cores"] = round(cpu2(O, cfg, L, R, n2, res["host_threads"]), 1)
        r["cpu3_cv2"] = [cpu3(cfg, L[0], 1), cpu3(cfg, L[0], res["host_threads"])]
        if have_gpu:
            r["gpu"] = gpu_legs(cfg, stereo, L, R, batch)
        res["configs"][name] = r
        print(name, json.dumps(r), file=sys.stderr, flush=True)
    print(json.dumps(res, indent=1))


if __name__ == "__main__":
    main()
