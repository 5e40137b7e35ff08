"""Seeded synthetic KITTI/EuRoC/TUM/4K-shaped images and stereo pairs.

There is no network and the reference ships no image data, so every test and benchmark runs on
these. The generator is deterministic given (W, H, seed) and identical on both machines (same
numpy / cv2 wheels); frame k of a batch uses seed0 + 2k.
"""
import cv2
import numpy as np

# name -> (W, H, nfeatures, nlevels, scaleFactor, iniThFAST, minThFAST, fx, bf)
CONFIGS = {
    "tum_mono": (640, 480, 1000, 8, 1.2, 20, 7, 517.306408, 0.0),
    "kitti": (1241, 376, 2000, 8, 1.2, 20, 7, 718.856, 386.1448),
    "euroc": (752, 480, 1200, 8, 1.2, 20, 7, 435.2047, 47.9064),
    "uhd": (3840, 2160, 8000, 8, 1.2, 20, 7, 2200.0, 440.0),
    # the one config the reference ships: Examples/Stereo/arducam.yaml:12,29,114-127, crop 640x400
    "arducam": (640, 400, 2000, 8, 1.2, 12, 7, 590.079638, 47.206371),
}


def synth(W, H, seed):
    """Left image: random rotated rectangles on grey, blurred, plus N(0,2) noise -> uint8 [H,W]."""
    rng = np.random.default_rng(seed)
    canvas = np.full((H, W), 128.0, np.float32)
    n = W * H // 900
    cx = rng.uniform(0, W, n)
    cy = rng.uniform(0, H, n)
    sw = rng.integers(4, 60, n)
    sh = rng.integers(4, 60, n)
    ang = rng.uniform(0, 180, n)
    val = rng.uniform(20, 235, n)
    for i in range(n):
        box = cv2.boxPoints(((float(cx[i]), float(cy[i])), (float(sw[i]), float(sh[i])), float(ang[i])))
        cv2.fillConvexPoly(canvas, np.round(box).astype(np.int32), float(val[i]))
    canvas = cv2.GaussianBlur(canvas, (0, 0), 0.8)
    canvas += rng.normal(0, 2, canvas.shape).astype(np.float32)
    return np.clip(np.rint(canvas), 0, 255).astype(np.uint8)


def synth_right(left, seed):
    """Right image: left warped by disparity d(x,y)=2+58(y/H)^2+4 sin(x/97) cos(y/53), + N(0,1) noise."""
    H, W = left.shape
    ys, xs = np.mgrid[0:H, 0:W].astype(np.float32)
    d = 2 + 58 * (ys / H) ** 2 + 4 * np.sin(xs / 97) * np.cos(ys / 53)
    right = cv2.remap(left, xs + d.astype(np.float32), ys, cv2.INTER_LINEAR, borderMode=cv2.BORDER_REFLECT_101)
    rng = np.random.default_rng(seed + 1)
    right = right.astype(np.float32) + rng.normal(0, 1, right.shape).astype(np.float32)
    return np.clip(np.rint(right), 0, 255).astype(np.uint8)


def stereo_pair(W, H, seed):
    left = synth(W, H, seed)
    return left, synth_right(left, seed)


def stereo_batch(W, H, n_frames, seed0=0):
    """(left[n,H,W], right[n,H,W]) uint8; frame k uses seed0 + 2k."""
    L = np.empty((n_frames, H, W), np.uint8)
    R = np.empty((n_frames, H, W), np.uint8)
    for k in range(n_frames):
        L[k], R[k] = stereo_pair(W, H, seed0 + 2 * k)
    return L, R
