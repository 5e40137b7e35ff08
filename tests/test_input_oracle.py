"""CPU tests: the oracle's gray conversion (C-8) and bilinear remap (C-9) against the real OpenCV kernels (cv2)."""
import cv2
import numpy as np
import pytest

from oracle import oracle as O


def rectification_maps(W, H, seed=0):
    """CV_32F maps from cv2.initUndistortRectifyMap for a plausible distorted, slightly rotated camera
    (what Examples/Stereo/arducam_images.cpp:257-272 builds from the settings file)."""
    rng = np.random.default_rng(seed)
    K = np.array([[0.92 * W, 0, W / 2 + 3.1], [0, 0.92 * W, H / 2 - 2.2], [0, 0, 1]])
    D = np.array([-0.28, 0.07, 0.0006, -0.0002, 0.0]) * (1 + 0.1 * rng.normal())
    rv = rng.normal(0, 0.01, 3)
    R, _ = cv2.Rodrigues(rv)
    P = K.copy()
    P[0, 2] += 4
    return cv2.initUndistortRectifyMap(K, D, R, P, (W, H), cv2.CV_32F)


def test_gray_all_colours():
    r, g, b = np.meshgrid(np.arange(256, dtype=np.uint8), np.arange(256, dtype=np.uint8), np.arange(256, dtype=np.uint8), indexing="ij")
    img = np.stack([b, g, r], -1).reshape(4096, 4096, 3)
    assert np.array_equal(O.cvt_gray(img, rgb_order=False), cv2.cvtColor(img, cv2.COLOR_BGR2GRAY))
    sub = np.ascontiguousarray(img[::7, ::5])
    assert np.array_equal(O.cvt_gray(sub, rgb_order=True), cv2.cvtColor(sub, cv2.COLOR_RGB2GRAY))
    sub4 = np.dstack([sub, np.full(sub.shape[:2], 91, np.uint8)])
    assert np.array_equal(O.cvt_gray(sub4, rgb_order=False), cv2.cvtColor(sub4, cv2.COLOR_BGRA2GRAY))
    assert np.array_equal(O.cvt_gray(sub4, rgb_order=True), cv2.cvtColor(sub4, cv2.COLOR_RGBA2GRAY))


@pytest.mark.parametrize("channels", [1, 3, 4])
def test_remap_against_cv2(channels):
    rng = np.random.default_rng(channels)
    for it in range(12):
        H, W = int(rng.integers(20, 260)), int(rng.integers(20, 330))
        src = rng.integers(0, 256, (H, W, channels), dtype=np.uint8)
        oh, ow = int(rng.integers(10, 260)), int(rng.integers(10, 330))
        ys, xs = np.mgrid[0:oh, 0:ow].astype(np.float32)
        mx = (xs * W / ow + rng.normal(0, 3, (oh, ow))).astype(np.float32)
        my = (ys * H / oh + rng.normal(0, 3, (oh, ow))).astype(np.float32)
        if it % 3 == 0:  # exact fractions incl. integers and .5/32 ties
            mx, my = np.round(mx * 4) / 4, np.round(my * 64) / 64
        if it % 5 == 0:  # far outside the image
            mx, my = mx * 1.5 - 20, my * 1.5 - 20
        mx, my = mx.astype(np.float32), my.astype(np.float32)
        s = src[:, :, 0] if channels == 1 else src
        assert np.array_equal(O.remap_linear(s, mx, my), cv2.remap(s, mx, my, cv2.INTER_LINEAR))


def test_rectification_maps_full_size():
    from orb_slam2_2021_b200 import synth

    W, H = 640, 480
    img = synth.synth(W, H, 5)
    m1, m2 = rectification_maps(W, H)
    assert np.array_equal(O.remap_linear(img, m1, m2), cv2.remap(img, m1, m2, cv2.INTER_LINEAR))
