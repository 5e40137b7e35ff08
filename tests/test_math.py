"""Known-answer tests of the shared deterministic math (orbx_math.h) against cv2 / float64."""
import cv2
import numpy as np

from oracle import oracle as O


def test_fast_atan2_matches_cv2():
    rng = np.random.default_rng(0)
    ys = rng.integers(-200000, 200000, 100000).astype(np.float32)
    xs = rng.integers(-200000, 200000, 100000).astype(np.float32)
    ref = cv2.phase(xs.reshape(-1, 1), ys.reshape(-1, 1), angleInDegrees=True).ravel()  # vector fastAtan2
    for i in range(0, 100000, 37):
        assert O.fast_atan2(ys[i], xs[i]) == cv2.fastAtan2(float(ys[i]), float(xs[i]))
    got = np.array([O.fast_atan2(y, x) for y, x in zip(ys[:20000], xs[:20000])], np.float32)
    scal = np.array([cv2.fastAtan2(float(y), float(x)) for y, x in zip(ys[:20000], xs[:20000])], np.float32)
    assert np.array_equal(got, scal)
    assert np.abs(got - ref[:20000]).max() < 1e-3  # the vector path may differ in the last ulp; scalar is the pin


def test_fast_atan2_axes():
    for y, x in [(0, 0), (0, 1), (1, 0), (0, -1), (-1, 0), (1, 1), (-1, -1), (5, -5), (-7, 7)]:
        assert O.fast_atan2(y, x) == cv2.fastAtan2(float(y), float(x))


def test_sincos_correctly_rounded():
    rng = np.random.default_rng(1)
    deg = np.concatenate([rng.uniform(0, 360, 200000), np.arange(0, 360.5, 0.5)]).astype(np.float32)
    rad = (deg * np.float32(np.pi / np.float64(np.float32(180.0)))).astype(np.float32)
    bad = 0
    for r in rad:
        s, c = O.sincos(r)
        bad += s != np.float32(np.sin(np.float64(r)))
        bad += c != np.float32(np.cos(np.float64(r)))
    assert bad == 0  # float64 kernel + one rounding == correctly rounded on this sample


def test_hamming():
    rng = np.random.default_rng(2)
    a = rng.integers(0, 256, (64, 32), dtype=np.uint8)
    b = rng.integers(0, 256, (64, 32), dtype=np.uint8)
    for i in range(64):
        assert O.hamming256(a[i], b[i]) == int(np.unpackbits(a[i] ^ b[i]).sum())
        assert O.hamming256(a[i], b[i]) == int(cv2.norm(a[i], b[i], cv2.NORM_HAMMING))
