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


# ---------------------------------------------------------------- Frame.cc: ComputeStereoMatches and the Frame helpers

def ref_pair(cfg, left, right):
    W, H, nf, nl, sf, ini, mn, fx, bf = cfg
    rl, rr = R.RefExtractor(nf, sf, nl, ini, mn), R.RefExtractor(nf, sf, nl, ini, mn)
    kl, dl = rl(left)
    kr, dr = rr(right)
    return rl, rr, kl, dl, kr, dr


@needs_ref
@pytest.mark.parametrize("name", ["arducam", "kitti", "euroc"])
def test_stereo_oracle_equals_reference(name):
    """Frame::ComputeStereoMatches as reference object code (src/Frame.cc:522-700 on the reference's own Frame, extractors
    and pyramids) against the oracle: mvuRight and mvDepth bit-identical, for mb = bf/fx and mb = 0 (first frame)."""
    cfg = synth.CONFIGS[name]
    W, H, nf, nl, sf, ini, mn, fx, bf = cfg
    for seed in (7, 8, 9):
        left, right = synth.stereo_pair(W, H, seed)
        rl, rr, kl, dl, kr, dr = ref_pair(cfg, left, right)
        ol, orr = O.OracleExtractor(nf, sf, nl, ini, mn), O.OracleExtractor(nf, sf, nl, ini, mn)
        assert_same(ol(left), (kl, dl), "L")
        assert_same(orr(right), (kr, dr), "R")
        for mb in (bf / fx, 0.0):
            our, odp = O.stereo_match(ol, orr, kl, dl, kr, dr, bf, mb)
            rur, rdp = R.stereo_match(rl, rr, kl, dl, kr, dr, bf, mb)
            assert np.array_equal(our, rur) and np.array_equal(odp, rdp), (name, seed, mb)
            assert (rur >= 0).sum() > 100
    # right image unrelated to the left one: few or no matches survive, same answer
    left, _ = synth.stereo_pair(W, H, 31)
    _, right = synth.stereo_pair(W, H, 32)
    rl, rr, kl, dl, kr, dr = ref_pair(cfg, left, right)
    ol, orr = O.OracleExtractor(nf, sf, nl, ini, mn), O.OracleExtractor(nf, sf, nl, ini, mn)
    ol(left), orr(right)
    our, odp = O.stereo_match(ol, orr, kl, dl, kr, dr, bf, bf / fx)
    rur, rdp = R.stereo_match(rl, rr, kl, dl, kr, dr, bf, bf / fx)
    assert np.array_equal(our, rur) and np.array_equal(odp, rdp)


@needs_ref
def test_frame_helpers_oracle_equal_reference():
    """AssignFeaturesToGrid / PosInGrid / GetFeaturesInArea / ComputeStereoFromRGBD of the reference's Frame (src/Frame.cc
    :279-294, :435-445, :376-433, :702-723) against the oracle's restatements (oracle/frame.cpp)."""
    rng = np.random.default_rng(12)
    for case in range(6):
        n = [0, 1, 50, 400, 1500, 3000][case]
        W, H = [(640, 400), (1241, 376)][case % 2]
        keys = np.zeros(n, O.KP_DTYPE)
        keys["x"], keys["y"] = rng.uniform(-8, W + 8, n), rng.uniform(-8, H + 8, n)  # some fall outside the grid
        if n > 10:  # exact cell-boundary and half-way coordinates
            keys["x"][:5] = [0, W, W / 64 * 0.5, W / 64 * 1.5, W - 1e-3]
            keys["y"][:5] = [0, H, H / 48 * 0.5, H / 48 * 2.5, H - 1e-3]
        keys["octave"] = rng.integers(0, 8, n)
        bounds = (0.0, float(W), 0.0, float(H))
        fr = O.FrameView(keys, None, None, np.eye(4), 500, 500, W / 2, H / 2, 40, 0.08, bounds, 1.2 ** np.arange(8))
        so, io = O.assign_features_to_grid(fr)
        sr, ir = R.assign_features_to_grid(keys, bounds)
        assert np.array_equal(so, sr) and np.array_equal(io, ir), case
        for q in range(40):
            x, y, r = rng.uniform(-30, W + 30), rng.uniform(-30, H + 30), rng.choice([0.5, 7, 25, 60, 400])
            lv = [(-1, -1), (0, 3), (2, -1), (3, 3), (1, 0)][q % 5]
            assert np.array_equal(O.features_in_area(fr, x, y, r, *lv), R.features_in_area(keys, bounds, x, y, r, *lv)), (case, q)
    # RGB-D: keypoints inside the depth map only (the reference reads imDepth.at<float>(v, u) unchecked)
    Hd, Wd, n = 48, 64, 500
    depth = rng.uniform(-0.5, 6, (Hd, Wd)).astype(np.float32)
    depth[rng.random((Hd, Wd)) < 0.1] = np.nan
    keys = np.zeros(n, O.KP_DTYPE)
    keys["x"], keys["y"] = rng.uniform(0, Wd - 0.01, n), rng.uniform(0, Hd - 0.01, n)
    un = keys.copy()
    un["x"] = (keys["x"] + rng.normal(0, 0.3, n)).astype(np.float32)
    our, odp = O.stereo_from_rgbd(keys, depth, 40.0, un["x"])
    rur, rdp = R.stereo_from_rgbd(keys, un, depth, 40.0)
    assert np.array_equal(our, rur) and np.array_equal(odp, rdp)


@needs_ref
@pytest.mark.gpu
@pytest.mark.parametrize("name", ["arducam", "kitti", "euroc"])
def test_cuda_stereo_equals_reference(name):
    """extract L + extract R + ComputeStereoMatches through the C-ABI against the same three steps of the reference."""
    import orb_slam2_2021_b200 as pkg

    cfg = synth.CONFIGS[name]
    W, H, nf, nl, sf, ini, mn, fx, bf = cfg
    exl, exr = pkg.ORBextractor(nf, sf, nl, ini, mn), pkg.ORBextractor(nf, sf, nl, ini, mn)
    for seed in (41, 42, 43):
        left, right = synth.stereo_pair(W, H, seed)
        rl, rr, rkl, rdl, rkr, rdr = ref_pair(cfg, left, right)
        kl, dl = exl(left)
        kr, dr = exr(right)
        assert_same((kl, dl), (rkl, rdl), "L")
        assert_same((kr, dr), (rkr, rdr), "R")
        for mb in (bf / fx, 0.0):
            ur, dp = pkg.compute_stereo_matches(exl, exr, kl, dl, kr, dr, bf, mb)
            rur, rdp = R.stereo_match(rl, rr, rkl, rdl, rkr, rdr, bf, mb)
            assert np.array_equal(ur, rur) and np.array_equal(dp, rdp), (name, seed, mb)


@needs_ref
@pytest.mark.gpu
def test_cuda_batch_stereo_equals_reference():
    """The benchmarked batch call: every output of every frame against the reference's extractor + ComputeStereoMatches."""
    import orb_slam2_2021_b200 as pkg

    cfg = synth.CONFIGS["kitti"]
    W, H, nf, nl, sf, ini, mn, fx, bf = cfg
    n = 5
    lefts, rights = zip(*[synth.stereo_pair(W, H, 500 + 2 * k) for k in range(n)])
    batch = pkg.StereoFrameBatch(nf, sf, nl, ini, mn, H, W, n, stereo=True)
    out = batch.process(np.stack(lefts), np.stack(rights), bf, bf / fx)
    for k in range(n):
        rl, rr, rkl, rdl, rkr, rdr = ref_pair(cfg, lefts[k], rights[k])
        m = int(out["n_left"][k])
        assert_same((out["kps_left"][k, :m], out["desc_left"][k, :m]), (rkl, rdl), f"{k}/L")
        rur, rdp = R.stereo_match(rl, rr, rkl, rdl, rkr, rdr, bf, bf / fx)
        assert np.array_equal(out["u_right"][k, :m], rur) and np.array_equal(out["depth"][k, :m], rdp), k


# ---------------------------------------------------------------- Frame::ComputeBoW over the vendored DBoW2

def _bow_cases(tmp_path):
    from . import bow_util as B

    for k, L, ragged in [(10, 3, False), (4, 5, False), (7, 4, True), (2, 6, False), (10, 5, False), (20, 3, True)]:
        arrays = B.synthetic_vocabulary(k, L, seed=k * 10 + L, ragged=ragged)
        path = str(tmp_path / f"voc_{k}_{L}.txt")
        R.write_vocabulary_text(path, k, L, *arrays)
        sets = [B.features_near_words(arrays[2], arrays[1], n, seed=n) if n else np.zeros((0, 32), np.uint8)
                for n in (0, 1, 33, 500, 2000)]
        yield L, arrays, R.RefVocabulary(path), sets


@needs_ref
def test_bow_oracle_equals_dbow2(tmp_path):
    """Frame::ComputeBoW (src/Frame.cc:447-454) through the reference's vendored DBoW2 — ORBVocabulary::loadFromTextFile,
    TemplatedVocabulary::transform with levelsup 4, FORB::distance, L1 scoring, all as compiled reference code — against the
    oracle (oracle/bow.cpp): word ids, the float64 weights bit for bit, node ids and feature lists."""
    for L, arrays, rv, sets in _bow_cases(tmp_path):
        ov = O.Vocabulary(L, *arrays)
        assert rv.n_words == ov.n_words
        for d in sets:
            for x, y in zip(ov.transform(d, 4), rv.transform(d)):
                assert x.dtype == y.dtype and np.array_equal(x, y), (L, len(d))


@needs_ref
@pytest.mark.gpu
def test_cuda_bow_equals_dbow2(tmp_path):
    """orbx_compute_bow (include/orbx_bow.h) against the same DBoW2 object code, one frame per call and all in one call."""
    import orb_slam2_2021_b200 as pkg

    for L, arrays, rv, sets in _bow_cases(tmp_path):
        gv = pkg.ORBVocabulary(L, *arrays)
        for got, d in zip(gv.transform(sets, 4), sets):
            for x, y in zip(got, rv.transform(d)):
                assert x.dtype == y.dtype and np.array_equal(x, y), (L, len(d))
    # descriptors of a real extraction
    W, H, nf, nl, sf, ini, mn, fx, bf = synth.CONFIGS["kitti"]
    _, desc = pkg.ORBextractor(nf, sf, nl, ini, mn)(synth.synth(W, H, 3))
    for L, arrays, rv, sets in _bow_cases(tmp_path):
        if L == 5:
            for x, y in zip(pkg.ORBVocabulary(L, *arrays).transform(desc, 4), rv.transform(desc)):
                assert np.array_equal(x, y)


# ---------------------------------------------------------------- ORBmatcher.cc: SearchByProjection, both overloads

SEARCH_CASES = [(0, 0.0, False, True, True, 300, 30), (1, 0.5, False, True, True, 300, 30), (2, -0.5, False, True, False, 300, 30),
                (3, 0.5, True, False, True, 300, 30), (4, 0.0, False, False, False, 300, 30), (5, 0.0, False, True, True, 2000, 400),
                (7, 0.0, False, True, True, 300, 200)]
LOCAL_CASES = [(0, 1.0, True, True, 400, 40), (1, 3.0, True, False, 400, 40), (2, 5.0, False, True, 400, 40),
               (3, 1.0, False, False, 400, 200), (4, 1.0, True, True, 3000, 500)]


def _scenario(seed, tz, ur, n_last, clones):
    from . import frames_util as U

    big = n_last > 1000
    return U.random_scenario(seed, n_last=n_last, n_extra=n_last // 2, tz=tz, clones=clones, with_uright=ur,
                             W=1241 if big else 640, H=376 if big else 400)


def _local(seed, ur, n_points, clones):
    from . import frames_util as U

    big = n_points > 1000
    return U.local_map_scenario(seed, n_points=n_points, n_extra=n_points // 2, clones=clones, with_uright=ur,
                                W=1241 if big else 640, H=376 if big else 400)


@needs_ref
@pytest.mark.parametrize("seed,tz,mono,ori,ur,n_last,clones", SEARCH_CASES)
def test_search_by_projection_oracle_equals_reference(seed, tz, mono, ori, ur, n_last, clones):
    """ORBmatcher::SearchByProjection(CurrentFrame, LastFrame, th, bMono) as the reference's own src/ORBmatcher.cc:1348-1491
    (with the real Frame::GetFeaturesInArea, DescriptorDistance, ComputeThreeMaxima) against the oracle: which map point ends
    up in which current keypoint, incl. blocked keypoints and the rotation-histogram removal."""
    from . import frames_util as U

    cur, last, has, world, pdesc, obs = _scenario(seed, tz, ur, n_last, clones)
    fc, fl = U.view(O.FrameView, cur), U.view(O.FrameView, last)
    for th in (7, 15):
        a, na = O.search_by_projection(fc, fl, has, world, pdesc, obs, th, mono, ori)
        b, nb = R.search_by_projection(fc, fl, has, world, pdesc, obs, th, mono, ori)
        assert na == nb and np.array_equal(a, b), (seed, th, na, nb)
        assert nb > 20


@needs_ref
@pytest.mark.parametrize("seed,th,ur,occ,n_points,clones", LOCAL_CASES)
def test_search_local_points_oracle_equals_reference(seed, th, ur, occ, n_points, clones):
    """ORBmatcher::SearchByProjection(F, vpMapPoints, th), src/ORBmatcher.cc:45-127, same comparison."""
    from . import frames_util as U

    F, occupied, valid, proj, level, vc, pdesc, obs = _local(seed, ur, n_points, clones)
    occupied = occupied if occ else None
    fv = U.view(O.FrameView, F)
    for ratio in (0.8, 0.6, 10.0):
        a, na = O.search_local_points(fv, occupied, valid, proj, level, vc, pdesc, obs, th, ratio)
        b, nb = R.search_local_points(fv, occupied, valid, proj, level, vc, pdesc, obs, th, ratio)
        assert na == nb and np.array_equal(a, b), (seed, ratio, na, nb)
    assert nb > 30


@needs_ref
@pytest.mark.gpu
@pytest.mark.parametrize("seed,tz,mono,ori,ur,n_last,clones", SEARCH_CASES)
def test_cuda_search_by_projection_equals_reference(seed, tz, mono, ori, ur, n_last, clones):
    """orbx_search_by_projection (include/orbx_frame.h) against the reference's own ORBmatcher.cc."""
    import orb_slam2_2021_b200 as pkg
    from . import frames_util as U

    cur, last, has, world, pdesc, obs = _scenario(seed, tz, ur, n_last, clones)
    m = pkg.Matcher(checkOrientation=ori)
    for th in (7, 15):
        b, nb = R.search_by_projection(U.view(O.FrameView, cur), U.view(O.FrameView, last), has, world, pdesc, obs, th, mono, ori)
        g, ng = m.SearchByProjection(U.view(pkg.FrameView, cur), U.view(pkg.FrameView, last), pkg.MapPointView(has, world, pdesc, obs), th, mono)
        assert ng == nb and np.array_equal(g, b), (seed, th, ng, nb)


@needs_ref
@pytest.mark.gpu
@pytest.mark.parametrize("seed,th,ur,occ,n_points,clones", LOCAL_CASES)
def test_cuda_search_local_points_equals_reference(seed, th, ur, occ, n_points, clones):
    import orb_slam2_2021_b200 as pkg
    from . import frames_util as U

    F, occupied, valid, proj, level, vc, pdesc, obs = _local(seed, ur, n_points, clones)
    occupied = occupied if occ else None
    m = pkg.Matcher()
    for ratio in (0.8, 0.6, 10.0):
        b, nb = R.search_local_points(U.view(O.FrameView, F), occupied, valid, proj, level, vc, pdesc, obs, th, ratio)
        g, ng = m.SearchLocalPoints(U.view(pkg.FrameView, F), pkg.TrackPointView(valid, proj, level, vc, pdesc, obs), th, occupied, ratio)
        assert ng == nb and np.array_equal(g, b), (seed, ratio, ng, nb)


# ---------------------------------------------------------------- the 4K configuration and the benchmarked batch sizes
@needs_ref
def test_oracle_equals_reference_uhd():
    """Config 5 of BASELINE.json (3840x2160, 8000 features): 31k-candidate levels through the reference's std::list quadtree."""
    W, H, nf, nl, sf, ini, mn, fx, bf = synth.CONFIGS["uhd"]
    o, r = O.OracleExtractor(nf, sf, nl, ini, mn), R.RefExtractor(nf, sf, nl, ini, mn)
    ims = images(W, H, 7)
    for what in ("left", "flat"):
        a, b = o(ims[what]), r(ims[what])
        assert_same(a, b, f"uhd/{what}")
        if what == "left":
            assert len(a[0]) >= nf


@needs_ref
@pytest.mark.gpu
def test_cuda_equals_reference_uhd():
    import orb_slam2_2021_b200 as pkg

    W, H, nf, nl, sf, ini, mn, fx, bf = synth.CONFIGS["uhd"]
    ex, r = pkg.ORBextractor(nf, sf, nl, ini, mn), R.RefExtractor(nf, sf, nl, ini, mn)
    ims = images(W, H, 8)
    for what in ("left", "right"):
        assert_same(ex(ims[what]), r(ims[what]), f"uhd/{what}")
    for l, lev in enumerate(ex.mvImagePyramid):
        assert np.array_equal(lev, r.level(l)), f"pyramid level {l}"


@needs_ref
@pytest.mark.gpu
@pytest.mark.parametrize("name", ["kitti", "euroc"])
def test_cuda_256_frame_stream_equals_reference_sampled(name):
    """The benchmarked call at its real size: 256 distinct stereo frames per call through orbx_batch_process_async (four
    chunks on two compute streams, two calls back to back on two output sets), every 16th frame of the SECOND call checked
    against the reference's extractor and its Frame::ComputeStereoMatches; configs 2/4 (KITTI) and 3 (EuRoC batch of 256)."""
    from concurrent.futures import ThreadPoolExecutor

    import orb_slam2_2021_b200 as pkg

    W, H, nf, nl, sf, ini, mn, fx, bf = synth.CONFIGS[name]
    n = 256
    pl, pr = pkg.PinnedArray((n, H, W), np.uint8), pkg.PinnedArray((n, H, W), np.uint8)

    def gen(k):
        pl.array[k], pr.array[k] = synth.stereo_pair(W, H, 5000 + 2 * k)

    with ThreadPoolExecutor(8) as ex:
        list(ex.map(gen, range(n)))
    batch = pkg.StereoFrameBatch(nf, sf, nl, ini, mn, H, W, n, stereo=True)
    out_a, out_b = batch.alloc_outputs(n, pinned=True), batch.alloc_outputs(n, pinned=True)
    mbf, mb = bf, bf / fx
    batch.process_async(pl.array, pr.array, mbf, mb, out_a)
    batch.process_async(pl.array, pr.array, mbf, mb, out_b)
    batch.wait()
    for key in out_a:  # the two calls saw the same frames
        assert np.array_equal(out_a[key].view(np.uint8), out_b[key].view(np.uint8)), key
    rl, rr = R.RefExtractor(nf, sf, nl, ini, mn), R.RefExtractor(nf, sf, nl, ini, mn)
    for k in range(3, n, 16):
        a, b = int(out_b["n_left"][k]), int(out_b["n_right"][k])
        kl, dl = rl(pl.array[k])
        kr, dr = rr(pr.array[k])
        assert_same((out_b["kps_left"][k, :a], out_b["desc_left"][k, :a]), (kl, dl), f"{name}/stream/{k}/left")
        assert_same((out_b["kps_right"][k, :b], out_b["desc_right"][k, :b]), (kr, dr), f"{name}/stream/{k}/right")
        ur, dp = R.stereo_match(rl, rr, kl, dl, kr, dr, mbf, mb)
        assert out_b["u_right"][k, :a].tobytes() == ur.tobytes() and out_b["depth"][k, :a].tobytes() == dp.tobytes(), (name, k)
        assert (ur >= 0).sum() > 50


@needs_ref
@pytest.mark.gpu
@pytest.mark.parametrize("nf,sf,nl,ini,mn", PARAMS)
def test_cuda_batch_equals_reference_parameters(nf, sf, nl, ini, mn):
    """The batch engine's per-level path (more than 8 images: TMA-staged resize, one launch per level) over the same
    parameter space as the single handle above (which takes the one-launch pyramid): scale factors 1.1 .. 2.0, 1 .. 12 levels."""
    import orb_slam2_2021_b200 as pkg

    W, H = (1280, 720) if nl == 12 else (701, 397)
    n = 10
    imgs = np.stack([synth.synth(W, H, 400 + k) for k in range(n)])
    b = pkg.StereoFrameBatch(nf, sf, nl, ini, mn, H, W, n, stereo=False)
    out = b.process(imgs, None)
    r = R.RefExtractor(nf, sf, nl, ini, mn)
    for k in (0, 4, 9):
        m = int(out["n_left"][k])
        assert_same((out["kps_left"][k, :m], out["desc_left"][k, :m]), r(imgs[k]), f"batch/{(nf, sf, nl, ini, mn)}/{k}")
        for l in range(nl):
            assert np.array_equal(b.level(k, 0, l), r.level(l)), (k, l)
