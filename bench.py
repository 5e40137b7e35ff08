#!/usr/bin/env python
"""bench.py — KITTI-stereo frames/s (ORB extract L+R + stereo match) on N B200, % of HBM roofline.

  python bench.py --gpus N --steps K --warmup W            # our CUDA path (torchrun for N > 1)
  python bench.py --impl reference --gpus N --steps K ...  # the CPU implementation on the host cores

One "step" = one pass of the hot path over a batch of `--frames` synthetic KITTI-shaped stereo frames
per GPU (weak scaling: per-GPU work fixed; `--scaling strong --total-frames 4096` is BASELINE.json's config 4 as written:
4096 frames split over the N GPUs, 4096/N per GPU and step). `value` = frames/s with the batch resident in HBM;
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
DISTINCT = 256  # distinct synthetic frames per GPU (= the default batch: no frame repeats inside a step); larger batches tile them


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--frames", type=int, default=256, help="stereo frames per GPU per step")
    ap.add_argument("--config", default=CONFIG_NAME)
    ap.add_argument("--scaling", default="weak", choices=["weak", "strong"],
                    help="weak: --frames per GPU and step; strong: --total-frames split over the GPUs (config 4 of BASELINE.json)")
    ap.add_argument("--total-frames", type=int, default=4096, help="frames per step over all GPUs with --scaling strong")
    ap.add_argument("--distinct", type=int, default=DISTINCT, help="distinct synthetic frames per GPU (tiled to the batch)")
    ap.add_argument("--no-latency", action="store_true", help="skip the single-frame latency block")
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


def make_frames(cfg, n, seed0=1000, distinct=DISTINCT, threads=None):
    """n stereo frames built from min(n, distinct) distinct seeded synthetic ones (generated on a few host threads)."""
    from concurrent.futures import ThreadPoolExecutor

    from orb_slam2_2021_b200 import synth

    W, H = cfg[0], cfg[1]
    d = max(1, min(distinct, n))
    bl, br = np.empty((d, H, W), np.uint8), np.empty((d, H, W), np.uint8)

    def one(k):
        bl[k], br[k] = synth.stereo_pair(W, H, seed0 + 2 * k)

    with ThreadPoolExecutor(threads or max(1, min(8, host_threads()))) as ex:
        list(ex.map(one, range(d)))
    if d == n:
        return bl, br
    idx = np.arange(n) % d
    return bl[idx], br[idx]


def workload(args, cfg):
    """The `config` both arms print (identical, so the two lines describe the same workload)."""
    W, H, nf, nl, sf, ini, mn = cfg[:7]
    c = {"workload": f"{args.config}: {W}x{H} stereo, {nf} features, {nl} levels, scale {sf}, FAST {ini}/{mn}; "
                     f"extract L + extract R + ComputeStereoMatches per frame",
         "scaling": args.scaling,
         "frames_per_gpu_per_step": args.frames if args.scaling == "weak" else None,
         "total_frames_per_step": args.total_frames if args.scaling == "strong" else None,
         "l2": "no flush: the per-step working set of a GPU (several GB) exceeds the 126 MB L2; the CPU arm streams frames "
               "from host memory"}
    return c


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
        return "reference", ("the reference's own src/ORBextractor.cc + Frame::ComputeStereoMatches (oracle/_ref, g++ -O3) over the "
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


def primitive_timing(cfg, fps, threads):
    """How much of the CPU arm is the scalar stand-ins for OpenCV's image primitives, and what a real OpenCV build would
    do instead: the same primitive calls of one image (7 resizes, FAST on every level at the initial threshold, 8 Gaussian
    blurs) timed once through the oracle's scalar restatements (what oracle/_ref runs) and once through cv2 (SIMD, one
    thread). `estimated_real_opencv` replaces the one by the other in the measured per-frame time; the GPU/CPU ratios of
    this bench are upper bounds by that factor."""
    try:
        import cv2

        from oracle import oracle as O
        from orb_slam2_2021_b200 import synth

        W, H, nf, nl, sf, ini, mn = cfg[:7]
        img = synth.synth(W, H, 12345)
        cv2.setNumThreads(1)

        def run(resize, fast, blur):
            t0 = time.perf_counter()
            lv = [img]
            for w, h in level_sizes(W, H, nl, sf)[1:]:
                lv.append(resize(lv[-1], w, h))
            for a in lv:
                fast(a)
                blur(a)
            return 1e3 * (time.perf_counter() - t0)

        fd = cv2.FastFeatureDetector_create(int(ini), True)
        t_cv = min(run(lambda a, w, h: cv2.resize(a, (w, h), interpolation=cv2.INTER_LINEAR), lambda a: fd.detect(a),
                       lambda a: cv2.GaussianBlur(a, (7, 7), 2, 2, borderType=cv2.BORDER_REFLECT_101)) for _ in range(3))
        t_sc = min(run(lambda a, w, h: O.resize_linear(a, w, h), lambda a: O.fast9(a, int(ini)), lambda a: O.gaussian7(a))
                   for _ in range(2))
        per_frame_ms = 1e3 * threads / fps  # one thread's time per stereo frame in the measured arm
        est = max(per_frame_ms - 2 * t_sc + 2 * t_cv, 2 * t_cv)
        return {"primitives": "scalar stand-ins for resize / FAST / GaussianBlur (no OpenCV C++ in the image): the ratio against "
                              "this arm is an upper bound",
                "scalar_primitives_ms_per_image": round(t_sc, 3), "opencv_primitives_ms_per_image": round(t_cv, 3),
                "measured_ms_per_frame_per_thread": round(per_frame_ms, 3),
                "estimated_real_opencv": {"value": threads * 1e3 / est, "unit": UNIT,
                                          "how": "measured per-frame time minus 2 x scalar primitives plus 2 x cv2 primitives, same threads"}}
    except Exception as e:  # the estimate is an annotation; the measured arm stands without it
        return {"primitives": f"scalar stand-ins (primitive timing unavailable: {e})"}


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
    left, right = make_frames(cfg, n, distinct=n)
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
        "warmup": args.warmup, "ms_per_step": 1e3 * total / args.steps, "higher_is_better": True, "scaling": args.scaling,
        "vs_baseline": None, "dtype": "u8", "data": "synthetic",
        "config": workload(args, cfg),
        "sample_frames_per_step": n,
        "cpu_baseline": dict({"value": fps, "unit": UNIT, "cores": threads, "kind": kind,
                              "sample": f"{n} distinct stereo frames per step, frame-parallel on {threads} threads ({cpu_model()}); {what}"},
                             **primitive_timing(cfg, fps, threads)),
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


def single_frame_latency(pkg, cfg, left, right, device, n=1000):
    """One stereo frame at a time through the C-ABI, wall clock per call from host images to host results, p50 / p99 over
    n frames (16 distinct): orbx_stereo_frame (one call per frame; pageable and page-locked input images), and the
    reference's own sequence orbx_extract (left), orbx_extract (right), orbx_stereo_match_resident — serial here, the
    reference runs the two extractions on two threads (src/Frame.cc:113-116)."""
    import ctypes as C

    from orb_slam2_2021_b200._lib import KP_DTYPE, lib, ptr

    W, H, nf, nl, sf, ini, mn, fx, bf = cfg
    d = min(16, len(left))
    fl, fr = pkg.ORBextractor(nf, sf, nl, ini, mn, device), pkg.ORBextractor(nf, sf, nl, ini, mn, device)
    el, er = pkg.ORBextractor(nf, sf, nl, ini, mn, device), pkg.ORBextractor(nf, sf, nl, ini, mn, device)
    L = lib()
    cap = L.orbx_max_keypoints(fl._h, H, W)
    kl, kr = np.zeros(cap, KP_DTYPE), np.zeros(cap, KP_DTYPE)
    dl, dr = np.zeros((cap, 32), np.uint8), np.zeros((cap, 32), np.uint8)
    ur, dp = np.zeros(cap, np.float32), np.zeros(cap, np.float32)
    nL, nR = C.c_int(0), C.c_int(0)
    pin_l, pin_r = pkg.PinnedArray((d, H, W), np.uint8), pkg.PinnedArray((d, H, W), np.uint8)
    pin_l.array[...], pin_r.array[...] = left[:d], right[:d]
    args = (ptr(kl), ptr(dl), C.byref(nL), ptr(kr), ptr(dr), C.byref(nR), cap, ptr(ur), ptr(dp))
    mbf, mb = float(bf), float(bf / fx)

    def pct(ts):
        a = np.sort(np.asarray(ts)) * 1e3
        return {"p50_ms": float(a[len(a) // 2]), "p99_ms": float(a[min(len(a) - 1, int(len(a) * 0.99))]), "mean_ms": float(a.mean())}

    def frame(src_l, src_r):
        def f(i):
            a, b = src_l[i % d], src_r[i % d]
            rc = L.orbx_stereo_frame(fl._h, fr._h, ptr(a), ptr(b), H, W, W, W, mbf, mb, *args)
            assert rc == 0 and nL.value > 0
        return f

    def seq(i):
        a, b = left[i % d], right[i % d]
        assert L.orbx_extract(el._h, ptr(a), H, W, W, ptr(kl), ptr(dl), cap, C.byref(nL)) == 0
        assert L.orbx_extract(er._h, ptr(b), H, W, W, ptr(kr), ptr(dr), cap, C.byref(nR)) == 0
        assert L.orbx_stereo_match_resident(el._h, er._h, nL.value, nR.value, mbf, mb, ptr(ur), ptr(dp)) == 0

    res = {}
    for name, fn, cnt in (("stereo_frame", frame(left, right), n), ("stereo_frame_pinned_input", frame(pin_l.array, pin_r.array), n),
                          ("extract_extract_match", seq, max(n // 4, 50))):
        for i in range(20):
            fn(i)
        ts = []
        for i in range(cnt):
            t0 = time.perf_counter()
            fn(i)
            ts.append(time.perf_counter() - t0)
        res[name] = dict(pct(ts), frames=cnt)
    res["single_frame_ms"] = res["stereo_frame"]["p50_ms"]
    res["api"] = ("orbx_stereo_frame through the C-ABI: host images in (pageable; page-locked in stereo_frame_pinned_input), host "
                  "keypoints / descriptors / uRight / depth of both sides out, one call per frame")
    return res


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
    if args.scaling == "strong":
        if args.total_frames % world:
            raise SystemExit("--total-frames must be a multiple of the number of GPUs")
        F = args.total_frames // world  # this rank's contiguous shard (orb_slam2_2021_b200/shard.py: frames are independent)
    else:
        F = args.frames
    left, right = make_frames(cfg, F, seed0=1000 + 2 * args.distinct * rank, distinct=args.distinct,
                              threads=max(1, host_threads() // max(world, 1)))
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
    # ---- the same stream for a caller that does not take the right image's keypoints and descriptors back (the ABI skips
    # NULL outputs; ORB-SLAM2 itself reads mvKeysRight / mDescriptorsRight only inside ComputeStereoMatches): less D2H
    slim = [{k: v for k, v in o.items() if k not in ("kps_right", "desc_right")} for o in (out, out2)]
    for i in range(2):
        b.process_async(pin_l.array, pin_r.array, mbf, mb, out=slim[i % 2])
    b.wait()
    barrier()
    t0 = time.perf_counter()
    for i in range(args.steps):
        b.process_async(pin_l.array, pin_r.array, mbf, mb, out=slim[i % 2])
    b.wait()
    slim_ms = reduce(1e3 * (time.perf_counter() - t0), dist.ReduceOp.MAX if dist else None)
    barrier()
    clocks = sampler.stop() if rank == 0 else None
    # ---- raw host-link ceiling at this N (tools/link_ceiling.py in short): the same bytes per step, pinned memory, one
    # cudaMemcpyAsync per direction on two streams, no kernels, all ranks at once; max over ranks
    link_ms = None
    try:
        nin, nout = int(left.nbytes + right.nbytes), int(sum(v.nbytes for v in out.values()))
        h_in = torch.empty(nin, dtype=torch.uint8, pin_memory=True)
        h_out = torch.empty(nout, dtype=torch.uint8, pin_memory=True)
        d_in = torch.empty(nin, dtype=torch.uint8, device="cuda")
        d_out = torch.zeros(nout, dtype=torch.uint8, device="cuda")
        s_up, s_dn = torch.cuda.Stream(), torch.cuda.Stream()
        ksteps = max(8, args.steps)
        best = None
        for rep in range(3):
            barrier()
            t0 = time.perf_counter()
            for _ in range(ksteps):
                with torch.cuda.stream(s_up):
                    d_in.copy_(h_in, non_blocking=True)
                with torch.cuda.stream(s_dn):
                    h_out.copy_(d_out, non_blocking=True)
            torch.cuda.synchronize()
            barrier()
            t = reduce(1e3 * (time.perf_counter() - t0) / ksteps, dist.ReduceOp.MAX if dist else None)
            if rep > 0:  # the first pass warms up
                best = t if best is None else min(best, t)
        link_ms = best
        del h_in, h_out, d_in, d_out
    except Exception as e:  # an annotation of the e2e number; never fails the bench
        print(f"link probe failed: {e}", file=sys.stderr)
    barrier()
    h2d = int(left.nbytes + right.nbytes)
    d2h = int(sum(v.nbytes for v in out.values()))
    checksum = int(out["n_left"].sum()) + int(out["n_right"].sum())
    assert checksum == n_l + n_r, "e2e results differ from the device-resident run"

    tot_l = reduce(float(n_l), dist.ReduceOp.SUM if dist else None)
    tot_r = reduce(float(n_r), dist.ReduceOp.SUM if dist else None)
    tot_m = reduce(float(n_m), dist.ReduceOp.SUM if dist else None)
    if rank != 0:
        return None

    latency = None
    if not args.no_latency:
        latency = single_frame_latency(pkg, cfg, left, right, local_rank)
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
        name = next(n for n in ("r02_traffic.json", "r01_traffic.json") if os.path.exists(os.path.join(ROOT, "profiles", n)))
        tj = json.load(open(os.path.join(ROOT, "profiles", name)))
        if args.config == "kitti":
            traffic = tj["stages"][top]["dram_bytes_per_frame"] * F
            traffic_src = f"profiles/{name} (ncu dram__bytes_read+write per stereo frame x frames per launch)"
    except Exception:
        pass
    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": ms_max / args.steps, "higher_is_better": True, "scaling": args.scaling, "vs_baseline": None,
        "dtype": "u8", "data": "synthetic",
        "config": workload(args, cfg),
        "run": {"frames_per_gpu_per_step": F, "frames_per_step": frames_total, "distinct_frames_per_gpu": min(args.distinct, F),
                "device_bytes_per_gpu": b.device_bytes,
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
                "without_right_outputs": {"value": frames_total * args.steps / (slim_ms / 1e3),
                                          "d2h_bytes_per_step": int(sum(v.nbytes for v in slim[0].values())) * world,
                                          "api": "the streaming call with kps_right / desc_right = NULL (not returned to the host)"},
                # rank 0's pinned buffers: NUMA node of its GPU, and whether orbx_host_alloc could place them there
                "host_numa": dict(zip(("node", "bound"), pkg.host_numa_node())),
                # what the host link alone allows at this N: the same bytes per step as plain pinned copies without kernels
                "link_ceiling_fps": (frames_total / (link_ms / 1e3)) if link_ms else None,
                "frac_of_link": (link_ms / (stream_ms / args.steps)) if link_ms else None,
                "link_probe": "pinned cudaMemcpyAsync H2D + D2H of the step's bytes on two streams, no kernels, all ranks at once "
                              "(tools/link_ceiling.py; profiles/r02_link_ceiling.json)"},
        "latency": latency,
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
        cl, cr = left[:n], right[:n]
        dt, _, _ = cpu_run(cfg, cl, cr, threads)
        kind, what = cpu_kind()
        line["cpu_baseline"] = dict({"value": n / dt, "unit": UNIT, "cores": threads, "kind": kind,
                                     "sample": f"the first {n} stereo frames of the batch, frame-parallel on {threads} threads "
                                               f"({cpu_model()}); {what}"}, **primitive_timing(cfg, n / dt, threads))
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
