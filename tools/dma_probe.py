"""Does concurrent PCIe DMA slow the device-resident step? (measurement tool; run on the GPU box)

Times StereoFrameBatch.run() alone, with a concurrent pinned H2D stream, with a concurrent D2H stream and with both.
"""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import orb_slam2_2021_b200 as pkg
from orb_slam2_2021_b200 import synth
from bench import make_frames

cfg = synth.CONFIGS["kitti"]
W, H, nf, nl, sf, ini, mn, fx, bf = cfg
F = 256
left, right = make_frames(cfg, F, seed0=1000)
b = pkg.StereoFrameBatch(nf, sf, nl, ini, mn, H, W, F, stereo=True, device=0)
b.upload(left, right)
stream = torch.cuda.ExternalStream(b.stream, device=0)
hbuf = torch.empty(240 << 20, dtype=torch.uint8).pin_memory()
dbuf = torch.empty(240 << 20, dtype=torch.uint8, device="cuda")
hout = torch.empty(64 << 20, dtype=torch.uint8).pin_memory()
dout = torch.empty(64 << 20, dtype=torch.uint8, device="cuda")
s_up, s_dn = torch.cuda.Stream(), torch.cuda.Stream()

def timed(up, dn, steps=10):
    for _ in range(3):
        b.run(F, bf, bf / fx, sync=False)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(stream)
    for _ in range(steps):
        if up:
            with torch.cuda.stream(s_up):
                dbuf.copy_(hbuf, non_blocking=True)
        if dn:
            with torch.cuda.stream(s_dn):
                hout.copy_(dout, non_blocking=True)
        b.run(F, bf, bf / fx, sync=False)
    e1.record(stream)
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / steps

for up, dn in ((0, 0), (1, 0), (0, 1), (1, 1), (0, 0), (1, 0), (1, 1), (1, 0), (0, 0)):
    print(f"h2d={up} d2h={dn}: {timed(up, dn):.3f} ms per 256-frame step", flush=True)
