#!/usr/bin/env python
"""bench.py — KITTI-stereo frames/s (ORB extract L+R + stereo match) on N B200, % of HBM roofline.

  python bench.py --gpus N --steps K --warmup W            # our CUDA path (torchrun for N > 1)
  python bench.py --impl reference --gpus N --steps K ...  # the CPU implementation on the host cores

One "step" = one pass of the hot path over a batch of `--frames` synthetic KITTI-shaped stereo frames
per GPU (weak scaling: per-GPU work fixed). `value` = frames/s with the batch resident in HBM;
`e2e` = the same through the public API with pinned HOST buffers (H2D + kernels + D2H of every step inside the
timed region): `e2e.value` through the streaming call (orbx_batch_process_async per step, one orbx_batch_wait at the
end), `e2e.blocking_call` through orbx_batch_process, which returns with each step's results in host memory.
Prints ONE JSON line on rank 0.
"""
import argparse
import json
import os
import statistics
import subprocess
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "KITTI-stereo frames/sec (ORB extract+stereo match)"
UNIT = "frames/s"
CONFIG_NAME = "kitti"
DISTINCT = 8  # distinct synthetic frames, tiled to the batch size


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--frames", type=int, default=256, help="stereo frames per GPU per step")
    ap.add_argument("--config", default=CONFIG_NAME)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--tune", action="append", default=[], metavar="KEY=VALUE",
                    help="orbx_set_tuning knob for experiments (tools/*_sweep.sh), e.g. --tune fast_cells=2")
    return ap.parse_args()


def peaks():
    """HBM roofline denominator: the driver-measured copy bandwidth, else the profiling guide's fallback."""
    try:
        m = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        return float(m["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md)"


def level_sizes(W, H, nlevels, sf):
    s, out = np.float32(1.0), []
    for _ in range(nlevels):
        inv = np.float32(1.0) / s
        out.append((int(np.rint(np.float32(W) * inv)), int(np.rint(np.float32(H) * inv))))
        s = np.float32(np.float64(s) * np.float64(np.float32(sf)))
    return out


def b_alg_per_frame(W, H, nlevels, sf, n_left, n_right):
    """compulsory bytes per stereo frame = inputs once + levels 1.. written once + outputs."""
    lv = level_sizes(W, H, nlevels, sf)
    return 2 * W * H + 2 * sum(w * h for w, h in lv[1:]) + 60 * (n_left + n_right) + 8 * n_left


def make_frames(cfg, n, seed0=1000):
    from orb_slam2_2021_b200 import synth

    W, H = cfg[0], cfg[1]
    d = min(DISTINCT, n)
    bl, br = synth.stereo_batch(W, H, d, seed0=seed0)
    idx = np.arange(n) % d
    return bl[idx], br[idx]


class ClockSampler:
    """nvidia-smi clocks/throttle reasons sampled every 200 ms while the timed region runs."""

    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index):
        self.idx, self.proc, self.path = gpu_index, None, f"/tmp/orbx_clocks_{os.getpid()}.csv"

    def start(self):
        try:
            self.fh = open(self.path, "w")
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-lms", "200", "-i", str(self.idx)], stdout=self.fh, stderr=subprocess.DEVNULL)
        except Exception:
            self.proc = None

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.25)
        self.proc.terminate()
        self.proc.wait()
        self.fh.close()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for line in open(self.path):
            p = [x.strip() for x in line.split(",")]
            if len(p) < 8:
                continue
            try:
                sm.append(float(p[1])); mx.append(float(p[2]))
            except ValueError:
                continue
            for nme, v in zip(names, p[4:8]):
                if v.lower().startswith("active"):
                    reasons.add(nme)
        os.unlink(self.path)
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "samples": len(sm), "reasons": sorted(reasons)}


def cpu_kind():
    """("reference", text) when oracle/_ref holds the reference's own ORBextractor.cc + Frame.cc (compiled where they lie
    against the OpenCV stand-in, oracle/ref_wrap.cpp), else ("port", text) for the oracle's restatement."""
    from oracle import ref as R

    if R.available():
        return "reference", ("the reference's own src/ORBextractor.cc + Frame::ComputeStereoMatches (oracle/_ref, g++ -O2) over the "
                             "oracle's scalar restatements of the OpenCV primitives (no OpenCV C++ in the image)")
    return "port", "C++ oracle port of the reference (oracle/_ref not built)"


def cpu_run(cfg, left, right, threads):
    """The CPU implementation of the path over the given frames, frame-parallel on `threads` host threads: the reference's
    own object code (oracle/_ref) where it was built, else the oracle port; returns (seconds, keypoints, matches)."""
    from concurrent.futures import ThreadPoolExecutor

    from oracle import oracle as O  # the checker, timed here only as the reported CPU baseline
    from oracle import ref as R

    W, H, nf, nl, sf, ini, mn, fx, bf = cfg
    use_ref = R.available()
    if use_ref:
        R.lib()

    def one(i):
        if use_ref:
            el, er = R.RefExtractor(nf, sf, nl, ini, mn), R.RefExtractor(nf, sf, nl, ini, mn)
        else:
            el, er = O.OracleExtractor(nf, sf, nl, ini, mn), O.OracleExtractor(nf, sf, nl, ini, mn)
        kl, dl = el(left[i])
        kr, dr = er(right[i])
        ur, _ = (R if use_ref else O).stereo_match(el, er, kl, dl, kr, dr, bf, bf / fx)
        return len(kl) + len(kr), int((ur >= 0).sum())

    t0 = time.perf_counter()
    with ThreadPoolExecutor(threads) as ex:
        res = list(ex.map(one, range(len(left))))  # ctypes releases the GIL inside the C++ calls
    dt = time.perf_counter() - t0
    return dt, sum(r[0] for r in res), sum(r[1] for r in res)


def host_threads():
    try:
        return max(1, len(os.sched_getaffinity(0)))
    except Exception:
        return os.cpu_count() or 1


def cpu_model():
    try:
        for line in open("/proc/cpuinfo"):
            if line.startswith("model name"):
                return line.split(":", 1)[1].strip()
    except Exception:
        pass
    return "unknown"


def reference_arm(args, rank):
    """--impl reference: the reference's CPU implementation of the path on the host cores: oracle/_ref (its own
    ORBextractor.cc and Frame.cc object code, cpu_baseline.kind "reference") where built, else the oracle port."""
    if rank != 0:
        return None
    from orb_slam2_2021_b200 import synth

    cfg = synth.CONFIGS[args.config]
    threads = host_threads()
    n = max(8, 2 * threads)
    left, right = make_frames(cfg, n)
    for _ in range(min(args.warmup, 1)):
        cpu_run(cfg, left[:threads], right[:threads], threads)
    total = 0.0
    for _ in range(args.steps):
        dt, _, _ = cpu_run(cfg, left, right, threads)
        total += dt
    fps = n * args.steps / total
    kind, what = cpu_kind()
    return {
        "impl": "reference", "metric": METRIC, "value": fps, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": 1e3 * total / args.steps, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "u8", "data": "synthetic",
        "config": {"workload": f"{args.config}: {cfg[0]}x{cfg[1]} stereo, {cfg[2]} features, {cfg[3]} levels; "
                               f"extract L + extract R + ComputeStereoMatches", "frames_per_step": n},
        "cpu_baseline": {"value": fps, "unit": UNIT, "cores": threads, "kind": kind,
                         "sample": f"{n} stereo frames per step, frame-parallel on {threads} threads ({cpu_model()}); {what}"},
        "e2e": {"value": fps, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }


def stage_alg_bytes(cfg, n_imgs, n_frames, kp_per_img, cand_per_img, cells_per_img):
    """Algorithmic (compulsory) HBM bytes of one launch of each stage over the whole batch."""
    W, H, nf, nl, sf = cfg[:5]
    lv = [w * h for w, h in level_sizes(W, H, nl, sf)]
    return {
        "pyramid": n_imgs * (sum(lv[:-1]) + sum(lv[1:])),           # read level l-1 once, write level l once
        "fast": n_imgs * (sum(lv) + 4 * cand_per_img + 4 * cells_per_img),  # read every level once, write candidates
        "octree": n_imgs * (4 * cells_per_img + 4 * cand_per_img + 8 * kp_per_img),
        "blur": n_imgs * 2 * sum(lv),                               # read + write every level
        "describe": n_imgs * kp_per_img * (709 + 512 + 8 + 60),      # disc + samples + kept record in, kp+desc out
        "stereo": n_frames * kp_per_img * (28 + 32 + 10 * 60 + 121 + 231 + 12),
    }


def ours_arm(args, rank, local_rank, world):
    import torch

    import orb_slam2_2021_b200 as pkg
    from orb_slam2_2021_b200 import synth

    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the product path has no CPU fallback")
    torch.cuda.set_device(local_rank)
    dist = None
    if world > 1:
        import torch.distributed as dist

        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))

    def barrier():
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()

    def reduce(v, op):
        if dist is None:
            return v
        t = torch.tensor([v], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=op)
        return float(t.item())

    for kv in args.tune:
        k, v = kv.split("=", 1)
        pkg.set_tuning(k, int(v))
    cfg = synth.CONFIGS[args.config]
    W, H, nf, nl, sf, ini, mn, fx, bf = cfg
    F = args.frames
    left, right = make_frames(cfg, F, seed0=1000 + 16 * rank)
    b = pkg.StereoFrameBatch(nf, sf, nl, ini, mn, H, W, F, stereo=True, device=local_rank)
    stream = torch.cuda.ExternalStream(b.stream, device=local_rank)
    mbf, mb = bf, bf / fx

    # ---- device-resident throughput: inputs are in HBM before the timed region
    b.upload(left, right)
    for _ in range(args.warmup):
        b.run(F, mbf, mb, sync=False)
    sampler = ClockSampler(local_rank)
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    if rank == 0:
        sampler.start()
    ev0.record(stream)
    for _ in range(args.steps):
        b.run(F, mbf, mb, sync=False)
    ev1.record(stream)
    ev1.synchronize()
    barrier()
    ms = ev0.elapsed_time(ev1)
    launches_per_step = b.launches_per_run
    # per-stage times (roofline leg) come from separate, serial runs with CUDA events at the stage boundaries: the
    # timed runs above interleave chunks of the batch on two streams, where stage boundaries do not exist
    b.set_profiling(True)
    for _ in range(min(args.steps, 5)):
        b.run(F, mbf, mb, sync=False)
    stages, runs = b.stage_times()
    b.set_profiling(False)
    n_l, n_r, n_m = b.stats(F)
    ms_max = reduce(ms, dist.ReduceOp.MAX if dist else None)

    # ---- end to end through the public API: pinned host images in, host results out, every step
    pin_l, pin_r = pkg.PinnedArray(left.shape, np.uint8), pkg.PinnedArray(right.shape, np.uint8)
    pin_l.array[...] = left
    pin_r.array[...] = right
    out = b.alloc_outputs(F, pinned=True)
    for _ in range(args.warmup):
        b.process(pin_l.array, pin_r.array, mbf, mb, out=out)
    barrier()
    ev0.record(stream)
    t0 = time.perf_counter()
    for _ in range(args.steps):
        b.process(pin_l.array, pin_r.array, mbf, mb, out=out)
    ev1.record(stream)
    ev1.synchronize()
    wall_ms = 1e3 * (time.perf_counter() - t0)
    barrier()
    e2e_ms = reduce(max(ev0.elapsed_time(ev1), wall_ms), dist.ReduceOp.MAX if dist else None)
    # ---- the same end-to-end work as a stream of batches: process_async back to back (the uploads of step k+1 overlap the
    # kernels of step k), two output sets, ONE wait after the last step; wall clock around all K steps incl. the drain
    out2 = b.alloc_outputs(F, pinned=True)
    for i in range(max(args.warmup, 2)):
        b.process_async(pin_l.array, pin_r.array, mbf, mb, out=(out, out2)[i % 2])
    b.wait()
    barrier()
    t0 = time.perf_counter()
    for i in range(args.steps):
        b.process_async(pin_l.array, pin_r.array, mbf, mb, out=(out, out2)[i % 2])
    b.wait()
    stream_ms = reduce(1e3 * (time.perf_counter() - t0), dist.ReduceOp.MAX if dist else None)
    barrier()
    assert int(out2["n_left"].sum()) + int(out2["n_right"].sum()) == n_l + n_r, "streamed results differ from the device-resident run"
    clocks = sampler.stop() if rank == 0 else None
    h2d = int(left.nbytes + right.nbytes)
    d2h = int(sum(v.nbytes for v in out.values()))
    checksum = int(out["n_left"].sum()) + int(out["n_right"].sum())
    assert checksum == n_l + n_r, "e2e results differ from the device-resident run"

    tot_l = reduce(float(n_l), dist.ReduceOp.SUM if dist else None)
    tot_r = reduce(float(n_r), dist.ReduceOp.SUM if dist else None)
    tot_m = reduce(float(n_m), dist.ReduceOp.SUM if dist else None)
    if rank != 0:
        return None

    frames_total = F * world
    value = frames_total * args.steps / (ms_max / 1e3)
    e2e = frames_total * args.steps / (e2e_ms / 1e3)
    peak, peak_src = peaks()
    kp_img = (n_l + n_r) / (2.0 * F)
    # candidates / cells per image from the device (rank 0, frame 0, both sides averaged)
    cand_img = sum(len(b.candidates(0, s, l)) for s in (0, 1) for l in range(nl)) / 2.0
    cells_img = {"kitti": 1220, "euroc": 982, "tum_mono": 815}.get(args.config, 1000)
    alg = stage_alg_bytes(cfg, 2 * F, F, kp_img, cand_img, cells_img)
    top = max(stages, key=lambda k: stages[k])
    top_ms = stages[top] / max(runs, 1)
    achieved = alg[top] / (top_ms / 1e3) / 1e9
    balg = b_alg_per_frame(W, H, nl, sf, n_l / F, n_r / F)
    # DRAM traffic of the dominant stage per launch, from the committed ncu capture (profiles/), scaled to this batch
    traffic, traffic_src = None, None
    try:
        tj = json.load(open(os.path.join(ROOT, "profiles", "r01_traffic.json")))
        if args.config == "kitti":
            traffic = tj["stages"][top]["dram_bytes_per_frame"] * F
            traffic_src = "profiles/r01_traffic.json (ncu dram__bytes_read+write per stereo frame x frames per launch)"
    except Exception:
        pass
    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": ms_max / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "u8", "data": "synthetic",
        "config": {"workload": f"{args.config}: {W}x{H} stereo, {nf} features, {nl} levels, FAST {ini}/{mn}; "
                               f"extract L + extract R + ComputeStereoMatches",
                   "frames_per_gpu_per_step": F, "frames_per_step": frames_total,
                   "l2": f"no flush: per-step working set {b.device_bytes / 2**20:.0f} MiB per GPU exceeds the 126 MB L2",
                   "keypoints_per_frame": [tot_l / frames_total, tot_r / frames_total],
                   "matches_per_frame": tot_m / frames_total},
        # headline: the streaming call (a continuous feed of batches, which is what a throughput figure describes); the
        # blocking call, which pays the pipeline ramp and drain on every step, is reported beside it
        "e2e": {"value": frames_total * args.steps / (stream_ms / 1e3), "unit": UNIT, "h2d_bytes_per_step": h2d * world,
                "d2h_bytes_per_step": d2h * world, "ms_per_step": stream_ms / args.steps,
                "api": "orbx_batch_process_async back to back (pinned host images in, pinned host results out, two output sets), one "
                       "orbx_batch_wait after the last step; wall clock around all steps including the drain",
                "blocking_call": {"value": e2e, "ms_per_step": e2e_ms / args.steps,
                                  "api": "orbx_batch_process: returns with the step's results in host memory"},
                # rank 0's pinned buffers: NUMA node of its GPU, and whether orbx_host_alloc could place them there
                "host_numa": dict(zip(("node", "bound"), pkg.host_numa_node()))},
        "gpu_launches": launches_per_step * args.steps,
        "roofline": {"bound": "hbm", "kernel": top, "achieved": achieved, "peak": peak, "unit": "GB/s",
                     "frac": achieved / peak, "traffic": traffic, "traffic_source": traffic_src, "peak_source": peak_src,
                     "alg_bytes_per_launch": alg[top], "ms_per_launch": top_ms,
                     "stage_ms_per_step": {k: v / max(runs, 1) for k, v in stages.items()},
                     "path": {"alg_bytes_per_frame": balg, "achieved": value / world * balg / 1e9,
                              "frac": value / world * balg / 1e9 / peak}},
        "clocks": clocks,
    }
    if world == 1 and not args.no_cpu_baseline:
        threads = host_threads()
        n = max(16, 2 * threads)
        cl, cr = make_frames(cfg, n)
        dt, _, _ = cpu_run(cfg, cl, cr, threads)
        kind, what = cpu_kind()
        line["cpu_baseline"] = {"value": n / dt, "unit": UNIT, "cores": threads, "kind": kind,
                                "sample": f"{n} stereo frames, frame-parallel on {threads} threads ({cpu_model()}); {what}"}
    return line


def main():
    # Keep stdout clean for the ONE JSON line: libraries (NCCL prints its version banner to stdout) write to fd 1,
    # so fd 1 is pointed at stderr for the duration of the run and the line goes to the saved descriptor.
    sys.stdout.flush()
    saved_stdout = os.dup(1)
    os.dup2(2, 1)
    args = parse_args()
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if args.impl == "reference":
        line = reference_arm(args, rank)
    else:
        line = ours_arm(args, rank, local_rank, world)
    if rank == 0 and line is not None:
        sys.stdout.flush()
        os.write(saved_stdout, (json.dumps(line) + "\n").encode())
    if world > 1 and args.impl == "ours":
        import torch.distributed as dist

        dist.destroy_process_group()


if __name__ == "__main__":
    main()
