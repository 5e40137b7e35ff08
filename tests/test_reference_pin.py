"""The parity pin: the REFERENCE'S OWN ORB_SLAM2::ORBextractor (src/ORBextractor.cc compiled where it lies into
oracle/_ref/libref_orbextractor.so, see oracle/ref_wrap.cpp) against the oracle's restatement (CPU tests) and against the
CUDA path through the C-ABI (-m gpu tests).

What runs as reference object code: the constructor tables, ComputePyramid, the cell loop with the iniTh/minTh fallback,
DistributeOctTree / DivideNode (std::list, std::sort on (size, node pointer)), IC_Angle, computeOrbDescriptor with its
pattern, the octave/size/scale assembly. What is substituted underneath: the five OpenCV functions it calls, by the oracle's
primitives that tests/test_oracle_kat.py pins bit for bit against cv2 4.13. The pointer tie-break of :687 is made
deterministic by a bump allocator (addresses grow in allocation order), see ref_wrap.cpp.

Bar: every keypoint field including the angle, and every descriptor byte, identical.
"""
import numpy as np
import pytest

from oracle import oracle as O
from oracle import ref as R
from orb_slam2_2021_b200 import synth

needs_ref = pytest.mark.skipif(not R.available(), reason="oracle/_ref not built and /root/reference absent")

FIELDS = ("x", "y", "size", "angle", "response", "octave", "class_id")


def assert_same(a, b, what):
    (ka, da), (kb, db) = a, b
    assert len(ka) == len(kb), (what, len(ka), len(kb))
    for f in FIELDS:
        assert np.array_equal(ka[f], kb[f]), (what, f, int((ka[f] != kb[f]).sum()))
    assert np.array_equal(da, db), (what, "descriptor bytes", int((da != db).sum()))


def images(W, H, seed):
    """Synthetic scene, its stereo partner, noise only, and a nearly flat image (cells fall back to minTh or stay empty)."""
    left, right = synth.stereo_pair(W, H, seed)
    rng = np.random.default_rng(seed)
    noise = rng.integers(0, 256, (H, W), dtype=np.uint8)
    flat = np.clip(128 + rng.normal(0, 3.0, (H, W)), 0, 255).astype(np.uint8)
    flat[H // 3:H // 3 + 40, W // 4:W // 4 + 60] += 14  # one faint rectangle: corners only below iniTh
    return {"left": left, "right": right, "noise": noise, "flat": flat}


@needs_ref
@pytest.mark.parametrize("name", ["tum_mono", "arducam", "kitti", "euroc"])
def test_oracle_equals_reference_configs(name):
    W, H, nf, nl, sf, ini, mn, fx, bf = synth.CONFIGS[name]
    o, r = O.OracleExtractor(nf, sf, nl, ini, mn), R.RefExtractor(nf, sf, nl, ini, mn)
    to, tr = o.tables(), r.tables()
    for k in tr:
        assert np.array_equal(to[k], tr[k]), k
    for seed in (5, 6):
        for what, img in images(W, H, seed).items():
            assert_same(o(img), r(img), f"{name}/{seed}/{what}")
            for l in range(nl):
                assert np.array_equal(o.level(l), r.level(l)), (name, seed, what, "pyramid level", l)


@needs_ref
@pytest.mark.parametrize("nf,sf,nl,ini,mn", [(500, 1.1, 4, 20, 7), (3000, 1.2, 8, 12, 7), (1000, 1.5, 5, 30, 10),
                                             (800, 2.0, 3, 20, 7), (1500, 1.2, 1, 20, 7), (2000, 1.2, 12, 20, 7),
                                             (1000, 1.3, 6, 7, 7), (200, 1.2, 8, 60, 40), (1000, 1.2, 8, 5, 20),
                                             (600, 1.2, 8, 130, 7), (600, 1.2, 8, 200, 128)])
def test_oracle_equals_reference_parameters(nf, sf, nl, ini, mn):
    W, H = (1280, 720) if nl == 12 else (701, 397)
    o, r = O.OracleExtractor(nf, sf, nl, ini, mn), R.RefExtractor(nf, sf, nl, ini, mn)
    for what, img in images(W, H, 17).items():
        assert_same(o(img), r(img), f"{(nf, sf, nl, ini, mn)}/{what}")


@needs_ref
def test_reference_reuse_and_empty():
    """One extractor over many images (the arena is dropped per call), strided input, a blank image (no keypoints)."""
    o, r = O.OracleExtractor(1000, 1.2, 8, 20, 7), R.RefExtractor(1000, 1.2, 8, 20, 7)
    for seed in range(4):
        left, _ = synth.stereo_pair(640, 480, 40 + seed)
        assert_same(o(left), r(left), f"reuse/{seed}")
    wide = np.zeros((480, 700), np.uint8)
    wide[:, :640] = synth.stereo_pair(640, 480, 3)[0]
    view = wide[:, :640]
    ro = r(view) if view.flags["C_CONTIGUOUS"] else r(np.ascontiguousarray(view))
    assert_same(o(np.ascontiguousarray(view)), ro, "strided")
    blank = np.full((480, 640), 77, np.uint8)
    ko, do = o(blank)
    kr, dr = r(blank)
    assert len(ko) == len(kr) == 0


@needs_ref
@pytest.mark.gpu
@pytest.mark.parametrize("name,seeds", [("tum_mono", range(3)), ("arducam", range(3)), ("kitti", range(4)), ("euroc", range(3))])
def test_cuda_equals_reference(name, seeds):
    """The CUDA path through the C-ABI against the reference's own extractor, no oracle in between."""
    import orb_slam2_2021_b200 as pkg

    W, H, nf, nl, sf, ini, mn, fx, bf = synth.CONFIGS[name]
    ex, r = pkg.ORBextractor(nf, sf, nl, ini, mn), R.RefExtractor(nf, sf, nl, ini, mn)
    for seed in seeds:
        for what, img in images(W, H, 200 + seed).items():
            assert_same(ex(img), r(img), f"{name}/{seed}/{what}")
        for l, lev in enumerate(ex.mvImagePyramid):
            assert np.array_equal(lev, r.level(l)), f"pyramid level {l}"


PARAMS = [(500, 1.1, 4, 20, 7), (3000, 1.2, 8, 12, 7), (1000, 1.5, 5, 30, 10), (800, 2.0, 3, 20, 7), (1500, 1.2, 1, 20, 7),
          (2000, 1.2, 12, 20, 7), (1000, 1.3, 6, 7, 7), (200, 1.2, 8, 60, 40), (1000, 1.2, 8, 5, 20), (600, 1.2, 8, 130, 7),
          (600, 1.2, 8, 200, 128)]


@needs_ref
@pytest.mark.gpu
@pytest.mark.parametrize("nf,sf,nl,ini,mn", PARAMS)
def test_cuda_equals_reference_parameters(nf, sf, nl, ini, mn):
    """Parameter space incl. iniTh < minTh and thresholds above 127 (the SWAR pre-test saturates there)."""
    import orb_slam2_2021_b200 as pkg

    W, H = (1280, 720) if nl == 12 else (701, 397)
    ex, r = pkg.ORBextractor(nf, sf, nl, ini, mn), R.RefExtractor(nf, sf, nl, ini, mn)
    for what, img in images(W, H, 23).items():
        assert_same(ex(img), r(img), f"{(nf, sf, nl, ini, mn)}/{what}")


@needs_ref
@pytest.mark.gpu
def test_cuda_batch_equals_reference():
    """The batch engine (the benchmarked call) on KITTI-shaped stereo frames against the reference's extractor."""
    import orb_slam2_2021_b200 as pkg

    W, H, nf, nl, sf, ini, mn, fx, bf = synth.CONFIGS["kitti"]
    n = 6
    lefts, rights = zip(*[synth.stereo_pair(W, H, 300 + 2 * k) for k in range(n)])
    batch = pkg.StereoFrameBatch(nf, sf, nl, ini, mn, H, W, n, stereo=True)
    out = batch.process(np.stack(lefts), np.stack(rights), bf, bf / fx)
    r = R.RefExtractor(nf, sf, nl, ini, mn)
    for k in range(n):
        for side, img in (("left", lefts[k]), ("right", rights[k])):
            m = int(out["n_" + side][k])
            assert_same((out["kps_" + side][k, :m], out["desc_" + side][k, :m]), r(img), f"batch/{k}/{side}")
