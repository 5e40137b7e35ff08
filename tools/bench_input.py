#!/usr/bin/env python
"""Measurement for the N3 row: raw frames -> level 0 (remap + crop + gray) on the GPU against OpenCV's own
kernels through cv2 on the host (the reference's code path for this step IS those two OpenCV calls).

  python tools/bench_input.py > gpurun_out/input_bench.json      (run it again under ncu for kernel time / DRAM bytes)
"""
import json
import os
import sys
import time

import cv2
import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import orb_slam2_2021_b200 as pkg  # noqa: E402
from orb_slam2_2021_b200 import synth  # noqa: E402
from tests.test_input_oracle import rectification_maps  # noqa: E402


def main():
    import torch

    W, H, nf, nl, sf, ini, mn, fx, bf = synth.CONFIGS["kitti"]
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 256
    res = {"frames": n, "W": W, "H": H}
    base = np.stack([synth.synth(W, H, 900 + k) for k in range(4)])
    ml, mr = rectification_maps(W, H, 1), rectification_maps(W, H, 2)
    for ch in (1, 3):
        raw = base[np.arange(n) % 4]
        if ch == 3:
            raw = np.ascontiguousarray(np.repeat(raw[..., None], 3, axis=3))
        pin_l, pin_r = pkg.PinnedArray(raw.shape, np.uint8), pkg.PinnedArray(raw.shape, np.uint8)
        pin_l.array[...] = raw
        pin_r.array[...] = raw
        b = pkg.StereoFrameBatch(nf, sf, nl, ini, mn, H, W, n, stereo=True)
        b.set_rectification(0, *ml)
        b.set_rectification(1, *mr)
        stream = torch.cuda.ExternalStream(b.stream)
        for _ in range(3):
            b.upload_raw(pin_l.array, pin_r.array)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        reps = 10
        e0.record(stream)
        for _ in range(reps):
            b.upload_raw(pin_l.array, pin_r.array)
        e1.record(stream)
        e1.synchronize()
        ms = e0.elapsed_time(e1) / reps
        # parity of what is timed (frame 0, left) against cv2 itself
        ref = cv2.remap(raw[0], ml[0], ml[1], cv2.INTER_LINEAR)
        if ch == 3:
            ref = cv2.cvtColor(ref, cv2.COLOR_BGR2GRAY)
        assert np.array_equal(b.level(0, 0, 0), ref)
        # host: cv2.remap (+ cvtColor) per image, default threading and single thread
        cpu = {}
        for thr in (0, 1):
            cv2.setNumThreads(thr if thr else os.cpu_count())
            t0 = time.perf_counter()
            k = 0
            while time.perf_counter() - t0 < 2.0:
                r = cv2.remap(raw[k % n], ml[0], ml[1], cv2.INTER_LINEAR)
                if ch == 3:
                    cv2.cvtColor(r, cv2.COLOR_BGR2GRAY)
                k += 1
            cpu["all_threads" if thr == 0 else "1_thread"] = k / (time.perf_counter() - t0)
        out_bytes = 2 * n * W * H
        alg = 2 * n * W * H * (ch + 1) + 4 * W * H * 4  # raw read once + gray written once + the four map planes once
        res[f"channels_{ch}"] = {"gpu_ms_per_batch_incl_h2d": ms, "gpu_images_per_s_incl_h2d": 2 * n / (ms / 1e3),
                                 "h2d_bytes": int(2 * raw.nbytes), "alg_bytes_kernel": int(alg), "out_bytes": int(out_bytes),
                                 "cv2_images_per_s": cpu}
        del b
    print(json.dumps(res))


if __name__ == "__main__":
    main()
