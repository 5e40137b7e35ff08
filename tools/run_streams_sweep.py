"""Device-resident throughput of the batch engine against the chunk / stream layout of Engine::runBatch.
One process per setting (orbx_set_tuning keys run_chunks / run_streams): python tools/run_streams_sweep.py [chunks streams]."""
import os
import subprocess
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def one():
    import numpy as np
    import torch

    import orb_slam2_2021_b200 as pkg
    from orb_slam2_2021_b200 import synth

    W, H, nf, nl, sf, ini, mn, fx, bf = synth.CONFIGS["kitti"]
    chunks, streams = int(sys.argv[1]), int(sys.argv[2])
    pkg.set_tuning("run_chunks", chunks)
    pkg.set_tuning("run_streams", streams)
    F = 256
    bl, br = synth.stereo_batch(W, H, 8, seed0=1)
    idx = np.arange(F) % 8
    b = pkg.StereoFrameBatch(nf, sf, nl, ini, mn, H, W, F, stereo=True)
    b.upload(np.ascontiguousarray(bl[idx]), np.ascontiguousarray(br[idx]))
    for _ in range(3):
        b.run(F, bf, bf / fx)
    torch.cuda.synchronize()
    best = 1e9
    for rep in range(3):
        t0 = time.perf_counter()
        for i in range(10):
            b.run(F, bf, bf / fx, sync=(i == 9))
        best = min(best, (time.perf_counter() - t0) / 10)
    print(f"chunks={chunks} streams={streams} "
          f"ms_per_step={best * 1e3:.3f} frames_per_s={F / best:.0f}", flush=True)


if __name__ == "__main__":
    if len(sys.argv) == 3:
        one()
    else:
        for streams in (2, 3, 4):
            for chunks in (2, 3, 4, 6, 8, 12):
                if chunks < streams:
                    continue
                subprocess.run([sys.executable, __file__, str(chunks), str(streams)], timeout=300)
