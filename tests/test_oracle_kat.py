"""Known-answer tests: the oracle's OpenCV primitives vs the real OpenCV kernels (cv2 4.13).

The reference has no tests of its own for this path; these KATs are the pin.
"""
import cv2
import numpy as np
import pytest

from oracle import oracle as O
from orb_slam2_2021_b200 import synth

SIZES = {"tum_mono": (640, 480), "kitti": (1241, 376), "euroc": (752, 480), "arducam": (640, 400)}


def level_sizes(W, H, n=8, sf=1.2):
    s = np.float32(1.0)
    out = []
    for _ in range(n):
        inv = np.float32(1.0) / s
        out.append((int(np.rint(np.float32(W) * inv)), int(np.rint(np.float32(H) * inv))))
        s = np.float32(np.float64(s) * np.float64(np.float32(sf)))
    return out


@pytest.mark.parametrize("name", list(SIZES))
@pytest.mark.parametrize("kind", ["random", "synthetic"])
def test_resize_chain_matches_cv2(name, kind):
    W, H = SIZES[name]
    img = (np.random.default_rng(5).integers(0, 256, (H, W), dtype=np.uint8) if kind == "random" else synth.synth(W, H, 11))
    cur = img
    for (w, h) in level_sizes(W, H)[1:]:
        ref = cv2.resize(cur, (w, h), interpolation=cv2.INTER_LINEAR)
        got = O.resize_linear(cur, w, h)
        assert np.array_equal(ref, got), (name, w, h)
        cur = ref


def test_resize_odd_ratios():
    rng = np.random.default_rng(6)
    for (sw, sh, dw, dh) in [(100, 80, 83, 67), (333, 211, 278, 176), (64, 64, 53, 53), (200, 100, 167, 83)]:
        img = rng.integers(0, 256, (sh, sw), dtype=np.uint8)
        assert np.array_equal(cv2.resize(img, (dw, dh), interpolation=cv2.INTER_LINEAR), O.resize_linear(img, dw, dh))


@pytest.mark.parametrize("shape", [(480, 640), (105, 346), (376, 1241), (62, 62), (134, 179), (7, 9)])
def test_gaussian_matches_cv2(shape):
    rng = np.random.default_rng(shape[0])
    for img in (rng.integers(0, 256, shape, dtype=np.uint8), np.full(shape, 255, np.uint8),
                (rng.integers(0, 2, shape) * 255).astype(np.uint8)):
        ref = cv2.GaussianBlur(img.copy(), (7, 7), 2, sigmaY=2, borderType=cv2.BORDER_REFLECT_101)
        assert np.array_equal(ref, O.gaussian7(img))
    img = rng.integers(0, 256, shape, dtype=np.uint8)
    inplace = img.copy()
    cv2.GaussianBlur(inplace, (7, 7), 2, dst=inplace, sigmaY=2, borderType=cv2.BORDER_REFLECT_101)  # the reference blurs in place
    assert np.array_equal(inplace, O.gaussian7(img))


def test_reflect101_matches_copymakeborder():
    row = np.arange(10, dtype=np.uint8).reshape(1, 10)
    ref = cv2.copyMakeBorder(row, 0, 0, 7, 7, cv2.BORDER_REFLECT_101)[0]
    got = [row[0, O.lib().orc_reflect101(p, 10)] for p in range(-7, 17)]
    assert list(ref) == got


def _cv_fast(roi, t):
    det = cv2.FastFeatureDetector_create(t, True, cv2.FAST_FEATURE_DETECTOR_TYPE_9_16)
    return [(k.pt[0], k.pt[1], k.response) for k in det.detect(roi)]


@pytest.mark.parametrize("t", [20, 12, 7])
def test_fast_matches_cv2(t):
    rng = np.random.default_rng(t)
    img = synth.synth(640, 480, 3)
    noise = rng.integers(0, 256, (200, 200), dtype=np.uint8)
    total = 0
    for k in range(40):
        h, w = int(rng.integers(5, 67)), int(rng.integers(5, 67))
        y, x = int(rng.integers(0, 400)), int(rng.integers(0, 560))
        roi = img[y:y + h, x:x + w]  # a view, like the reference's rowRange/colRange
        ref = _cv_fast(roi, t)
        got = O.fast9(roi, t)
        assert [(float(g["x"]), float(g["y"]), float(g["response"])) for g in got] == ref
        total += len(ref)
        roi = noise[y % 130:y % 130 + h, x % 130:x % 130 + w]
        ref = _cv_fast(roi, t)
        got = O.fast9(roi, t)
        assert [(float(g["x"]), float(g["y"]), float(g["response"])) for g in got] == ref
        total += len(ref)
    assert total > 500
