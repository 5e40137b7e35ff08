"""The drop-in, exercised by the reference's own caller: /root/reference/src/Frame.cc's stereo constructor (:80-152) run
(a) unchanged on the reference's CPU ORBextractor (oracle/_ref/libref_orbextractor.so) and
(b) with INTEGRATION.md's patch and nothing else changed on ORBextractorGPU / liborbx.so (oracle/_ref/libref_frame_gpu.so:
    `ORBextractor` = the shim class built with ORBX_WITH_OPENCV, i.e. the overload with the reference's exact
    operator()(cv::InputArray, cv::InputArray, std::vector<cv::KeyPoint>&, cv::OutputArray); body of ComputeStereoMatches
    = ComputeStereoMatchesGPU).
Both go through the constructor's own two std::threads, UndistortKeyPoints, ComputeStereoMatches, ComputeImageBounds and
AssignFeaturesToGrid; everything the constructor leaves in the Frame must be identical."""
import numpy as np
import pytest

from oracle import ref as R
from orb_slam2_2021_b200 import synth

pytestmark = pytest.mark.skipif(not R.available(), reason="oracle/_ref not built and /root/reference absent")


def _cam(name):
    W, H, nf, nl, sf, ini, mn, fx, bf = synth.CONFIGS[name]
    return np.array([fx, fx, W / 2.0, H / 2.0], np.float32)


@pytest.mark.parametrize("name", ["arducam", "euroc"])
def test_reference_constructor_equals_its_parts(name):
    """CPU only: the reference's constructor (extraction on its own two threads, node addresses from the shared arena)
    gives what the single calls of tests/test_reference_pin.py give — ORBextractor::operator() left and right,
    ComputeStereoMatches with the same mb, AssignFeaturesToGrid."""
    W, H, nf, nl, sf, ini, mn, fx, bf = synth.CONFIGS[name]
    left, right = synth.stereo_pair(W, H, 21)
    el, er = R.RefExtractor(nf, sf, nl, ini, mn), R.RefExtractor(nf, sf, nl, ini, mn)
    for mb in (bf / fx, 0.0):
        F = R.frame_stereo(el, er, left, right, _cam(name), bf, 35.0 * bf / fx, mb)
        kl, dl = el(left)
        kr, dr = er(right)
        ur, dp = R.stereo_match(el, er, kl, dl, kr, dr, bf, mb)
        assert F["N"] == len(kl) > 500
        assert F["mvKeys"].tobytes() == kl.tobytes() and F["mvKeysRight"].tobytes() == kr.tobytes()
        assert np.array_equal(F["mDescriptors"], dl) and np.array_equal(F["mDescriptorsRight"], dr)
        assert F["mvKeysUn"].tobytes() == kl.tobytes()  # zero distortion
        assert F["mvuRight"].tobytes() == ur.tobytes() and F["mvDepth"].tobytes() == dp.tobytes()
        start, items = R.assign_features_to_grid(kl, (0.0, float(W), 0.0, float(H)))
        assert np.array_equal(F["mGridStart"], start) and np.array_equal(F["mGridItems"], items)
        st = F["statics"]
        assert st["mb"] == np.float32(bf) / np.float32(fx) and st["fx"] == np.float32(fx) and st["mnMaxX"] == W and st["mnMaxY"] == H


def _same_frame(A, B):
    for k in ("mvKeys", "mvKeysRight", "mvKeysUn"):
        assert len(A[k]) == len(B[k]), k
        for f in A[k].dtype.names:  # every cv::KeyPoint field, the angle included, bit for bit
            assert A[k][f].tobytes() == B[k][f].tobytes(), (k, f)
    for k in ("mDescriptors", "mDescriptorsRight", "mGridStart", "mGridItems", "counts", "scaleTables"):
        assert np.array_equal(A[k], B[k]), k
    assert A["mvuRight"].tobytes() == B["mvuRight"].tobytes()
    assert A["mvDepth"].tobytes() == B["mvDepth"].tobytes()
    assert A["statics"] == B["statics"]


@pytest.mark.gpu
@pytest.mark.parametrize("name", ["kitti", "euroc", "arducam"])
def test_reference_frame_on_the_gpu_extractor_equals_reference_frame(name):
    W, H, nf, nl, sf, ini, mn, fx, bf = synth.CONFIGS[name]
    cl, cr = R.RefExtractor(nf, sf, nl, ini, mn), R.RefExtractor(nf, sf, nl, ini, mn)
    gl, gr = R.GpuShimExtractor(nf, sf, nl, ini, mn), R.GpuShimExtractor(nf, sf, nl, ini, mn)
    for seed, mb in ((31, bf / fx), (32, bf / fx), (33, 0.0)):  # consecutive frames on the same extractor objects
        left, right = synth.stereo_pair(W, H, seed)
        want = R.frame_stereo(cl, cr, left, right, _cam(name), bf, 35.0 * bf / fx, mb)
        got = R.frame_stereo_gpu(gl, gr, left, right, _cam(name), bf, 35.0 * bf / fx, mb)
        assert want["N"] > 500 and (want["mvuRight"] >= 0).sum() > 50
        _same_frame(got, want)


@pytest.mark.gpu
def test_input_array_overload_alone_and_degenerate_frames():
    """ORBextractorGPU::operator()(cv::InputArray, ...) — the overload with the reference's signature — against the
    reference's operator(); a frame without keypoints takes the constructor's early return (:120-121) on both sides."""
    W, H, nf, nl, sf, ini, mn, fx, bf = synth.CONFIGS["tum_mono"]
    c, g = R.RefExtractor(nf, sf, nl, ini, mn), R.GpuShimExtractor(nf, sf, nl, ini, mn)
    img = synth.synth(W, H, 77)
    (kc, dc), (kg, dg) = c(img), g(img)
    assert kc.tobytes() == kg.tobytes() and np.array_equal(dc, dg) and len(kc) > 500
    view = np.zeros((H, W + 24), np.uint8)  # a cv::Mat header with step > cols
    view[:, :W] = img
    kv, dv = g(view[:, :W])
    assert kv.tobytes() == kc.tobytes() and np.array_equal(dv, dc)
    flat = np.full((H, W), 128, np.uint8)
    cl, cr = R.RefExtractor(nf, sf, nl, ini, mn), R.RefExtractor(nf, sf, nl, ini, mn)
    gl, gr = R.GpuShimExtractor(nf, sf, nl, ini, mn), R.GpuShimExtractor(nf, sf, nl, ini, mn)
    K4 = np.array([fx, fx, W / 2, H / 2], np.float32)
    want = R.frame_stereo(cl, cr, flat, flat, K4, bf, 40.0, 0.0)
    got = R.frame_stereo_gpu(gl, gr, flat, flat, K4, bf, 40.0, 0.0)
    assert want["N"] == got["N"] == 0
