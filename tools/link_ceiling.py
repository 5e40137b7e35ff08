#!/usr/bin/env python
"""Raw host-link ceiling for bench.py's end-to-end numbers: pinned cudaMemcpyAsync H2D + D2H of one step's bytes, NO kernels.

Launch with one rank per GPU (like bench.py):
  python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29531 tools/link_ceiling.py
Every measurement is the wall time of K steps (each: H2D of `--h2d-mb` from pinned memory on one stream, D2H of `--d2h-mb`
into pinned memory on another) between two barriers, max over the ACTIVE ranks. Active sets: the first 1, 2, 4, 8 ranks
(what bench.py --gpus N uses), every GPU alone, GPUs {0..3} and {4..7}. Also H2D alone and D2H alone on the first N ranks,
and the engine's chunking (`--chunks` copies per direction and step instead of one). Rank 0 prints one JSON document.
"""
import argparse
import json
import os
import time

import torch
import torch.distributed as dist


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--h2d-mb", type=float, default=238.9)  # 256 KITTI stereo frames: 2 x 256 x 1241 x 376 bytes
    ap.add_argument("--d2h-mb", type=float, default=67.4)   # their keypoints, descriptors, uRight, depth, counts (full pitch)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--chunks", type=int, default=4)
    ap.add_argument("--frame-bytes", type=int, default=2 * 1241 * 376)
    ap.add_argument("--frames-per-step", type=int, default=256)
    args = ap.parse_args()
    rank, world, local = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    nin, nout = int(args.h2d_mb * 1e6), int(args.d2h_mb * 1e6)
    h_in = torch.empty(nin, dtype=torch.uint8, pin_memory=True); h_in.fill_(1)
    h_out = torch.empty(nout, dtype=torch.uint8, pin_memory=True); h_out.fill_(0)
    d_in = torch.empty(nin, dtype=torch.uint8, device="cuda")
    d_out = torch.ones(nout, dtype=torch.uint8, device="cuda")
    s_up, s_dn = torch.cuda.Stream(), torch.cuda.Stream()

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()

    def run(active, up=True, down=True, chunks=1):
        """ms per step, max over the active ranks"""
        mine = rank in active
        barrier()
        t = 0.0
        if mine:
            for rep in range(2):  # first pass = warm-up
                torch.cuda.synchronize()
                t0 = time.perf_counter()
                for _ in range(args.steps):
                    for c in range(chunks):
                        if up:
                            a, b = nin * c // chunks, nin * (c + 1) // chunks
                            with torch.cuda.stream(s_up):
                                d_in[a:b].copy_(h_in[a:b], non_blocking=True)
                        if down:
                            a, b = nout * c // chunks, nout * (c + 1) // chunks
                            with torch.cuda.stream(s_dn):
                                h_out[a:b].copy_(d_out[a:b], non_blocking=True)
                torch.cuda.synchronize()
                t = 1e3 * (time.perf_counter() - t0) / args.steps
        barrier()
        if world > 1:
            v = torch.tensor([t], dtype=torch.float64, device="cuda")
            dist.all_reduce(v, op=dist.ReduceOp.MAX)
            t = float(v.item())
        return t

    def entry(ms, n, up=True, down=True):
        gb = ((nin if up else 0) + (nout if down else 0)) * n / 1e9
        return {"ms_per_step": round(ms, 4), "agg_GBps": round(gb / (ms / 1e3), 2),
                "h2d_GBps_per_gpu": round(nin / 1e9 / (ms / 1e3), 2) if up else None,
                "d2h_GBps_per_gpu": round(nout / 1e9 / (ms / 1e3), 2) if down else None,
                "frames_per_s_ceiling": round(args.frames_per_step * n / (ms / 1e3)) if up else None}

    res = {"h2d_bytes_per_step_per_gpu": nin, "d2h_bytes_per_step_per_gpu": nout, "steps": args.steps, "world": world,
           "first_n": {}, "first_n_chunked": {}, "first_n_h2d_only": {}, "first_n_d2h_only": {}, "each_alone": {}, "halves": {}}
    for n in (1, 2, 4, 8):
        if n > world:
            break
        act = set(range(n))
        res["first_n"][str(n)] = entry(run(act), n)
        res["first_n_chunked"][str(n)] = entry(run(act, chunks=args.chunks), n)
        res["first_n_h2d_only"][str(n)] = entry(run(act, down=False), n, down=False)
        res["first_n_d2h_only"][str(n)] = entry(run(act, up=False), n, up=False)
    for g in range(world):
        res["each_alone"][str(g)] = entry(run({g}), 1)
    if world >= 8:
        res["halves"]["0-3"] = entry(run({0, 1, 2, 3}), 4)
        res["halves"]["4-7"] = entry(run({4, 5, 6, 7}), 4)
        res["halves"]["even"] = entry(run({0, 2, 4, 6}), 4)
    if rank == 0:
        print(json.dumps(res, indent=1))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
