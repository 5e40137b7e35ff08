"""orbx_stereo_match_static (include/orbx.h): the stereo match with its device scratch in a static array of the library
instead of allocated memory. Collected last on purpose: it must pass alone in a fresh
process and after all the other tests. Bar: identical to orbx_stereo_match (same kernels, same inputs), which the parity
tests pin against the oracle and the reference."""
import ctypes as C

import numpy as np
import pytest

import orb_slam2_2021_b200 as pkg
from orb_slam2_2021_b200 import synth
from orb_slam2_2021_b200._lib import ERR_ARG, KP_DTYPE, lib, ptr


def _static(exl, exr, kl, dl, kr, dr, mbf, mb):
    ur = np.full(len(kl), -7, np.float32)
    dp = np.full(len(kl), -7, np.float32)
    rc = lib().orbx_stereo_match_static(exl._h, exr._h, ptr(kl), ptr(dl), len(kl), ptr(kr), ptr(dr), len(kr), C.c_float(mbf),
                                        C.c_float(mb), ptr(ur), ptr(dp))
    return rc, ur, dp


@pytest.mark.gpu
@pytest.mark.parametrize("name", ["kitti", "euroc"])
def test_static_scratch_stereo_match_equals_the_allocating_entry_point(name):
    W, H, nf, nl, sf, ini, mn, fx, bf = synth.CONFIGS[name]
    exl, exr = pkg.ORBextractor(nf, sf, nl, ini, mn), pkg.ORBextractor(nf, sf, nl, ini, mn)
    for call, seed in enumerate((3, 11)):  # two calls in a row on the shared array, different frames
        left, right = synth.stereo_pair(W, H, seed)
        kl, dl = exl(left)
        kr, dr = exr(right)
        if call == 1:  # fewer keypoints than the call before: stale scratch contents must not leak into the result
            kl, dl, kr, dr = kl[: len(kl) // 2].copy(), dl[: len(dl) // 2].copy(), kr[: len(kr) // 3].copy(), dr[: len(dr) // 3].copy()
        want_u, want_d = pkg.compute_stereo_matches(exl, exr, kl, dl, kr, dr, bf, bf / fx)
        rc, got_u, got_d = _static(exl, exr, kl, dl, kr, dr, bf, bf / fx)
        assert rc == 0, lib().orbx_last_error()
        assert (want_u >= 0).sum() > 50
        assert np.array_equal(got_u, want_u) and np.array_equal(got_d, want_d)
        # and once more on the same inputs: the call is repeatable on the array it just used
        rc, again_u, again_d = _static(exl, exr, kl, dl, kr, dr, bf, bf / fx)
        assert rc == 0 and np.array_equal(again_u, want_u) and np.array_equal(again_d, want_d)


@pytest.mark.gpu
def test_static_scratch_limits_are_refused_with_a_clear_error():
    ex_l, ex_r = pkg.ORBextractor(500, 1.2, 2, 20, 7), pkg.ORBextractor(500, 1.2, 2, 20, 7)
    tall = synth.stereo_pair(2112, 4100, 5)  # 4100 rows > ORBX_STEREO_STATIC_MAX_ROWS (width >= rows / 2: the reference's nIni >= 1)
    kl, dl = ex_l(tall[0])
    kr, dr = ex_r(tall[1])
    rc, _, _ = _static(ex_l, ex_r, kl, dl, kr, dr, 40.0, 0.1)
    assert rc == ERR_ARG and b"static scratch" in lib().orbx_last_error()
    ur, dp = pkg.compute_stereo_matches(ex_l, ex_r, kl, dl, kr, dr, 40.0, 0.1)  # the allocating entry point takes it
    assert len(ur) == len(kl)
    # inside the limits, no keypoints on the right: every left keypoint unmatched
    W, H, nf, nl, sf, ini, mn, fx, bf = synth.CONFIGS["arducam"]
    sl, sr = pkg.ORBextractor(nf, sf, nl, ini, mn), pkg.ORBextractor(nf, sf, nl, ini, mn)
    left, right = synth.stereo_pair(W, H, 2)
    kl, dl = sl(left)
    sr(right)
    rc, u, d = _static(sl, sr, kl, dl, np.zeros(0, KP_DTYPE), np.zeros((0, 32), np.uint8), bf, bf / fx)
    assert rc == 0 and np.all(u == -1) and np.all(d == -1)
    # empty left side: defined, nothing launched
    rc, u, d = _static(sl, sr, kl[:0], dl[:0], kl, dl, bf, bf / fx)
    assert rc == 0 and len(u) == 0
