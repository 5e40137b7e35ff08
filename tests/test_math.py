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


def test_sincos_is_glibc_sincosf_on_every_angle():
    """src/ORBextractor.cc:109-110: the reference's object code calls glibc sincosf. The shared routine restates glibc's algorithm;
    this sweeps EVERY float from 0 to 2*pi*1.0001 (the angle is fastAtan2 * pi/180) against the host's libm: 1.09e9 inputs."""
    import os
    from concurrent.futures import ThreadPoolExecutor

    hi = int(np.array([2 * np.pi * 1.0001], np.float32).view(np.uint32)[0])
    threads = max(1, min(16, len(os.sched_getaffinity(0))))
    with ThreadPoolExecutor(threads) as ex:  # ctypes drops the GIL; thread t takes bit patterns t, t+threads, ...
        bad = list(ex.map(lambda t: O.sincos_vs_libm(t, hi, threads), range(threads)))
    assert sum(bad) == 0
    # negative angles and the rest of the fast path (|x| < 120) on a stride
    assert O.sincos_vs_libm(0x80000000, 0x80000000 + hi, 997) == 0
    assert O.sincos_vs_libm(hi, int(np.array([119.0], np.float32).view(np.uint32)[0]), 101) == 0


def test_sincos_known_values():
    for deg in (0.0, 90.0, 180.0, 270.0, 359.99997):
        r = np.float32(deg) * np.float32(np.pi / np.float64(np.float32(180.0)))
        s, c = O.sincos(r)
        assert abs(s - np.sin(np.float64(r))) < 1.2e-7 and abs(c - np.cos(np.float64(r))) < 1.2e-7
    assert O.sincos(0.0) == (0.0, 1.0)


def test_hamming():
    rng = np.random.default_rng(2)
    a = rng.integers(0, 256, (64, 32), dtype=np.uint8)
    b = rng.integers(0, 256, (64, 32), dtype=np.uint8)
    for i in range(64):
        assert O.hamming256(a[i], b[i]) == int(np.unpackbits(a[i] ^ b[i]).sum())
        assert O.hamming256(a[i], b[i]) == int(cv2.norm(a[i], b[i], cv2.NORM_HAMMING))
