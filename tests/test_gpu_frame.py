"""GPU parity for the frame-level rows through the C-ABI of include/orbx_frame.h:
every integer output bit-exact against the oracle, floats bit-exact (single IEEE operations)."""
import os

import numpy as np
import pytest

import orb_slam2_2021_b200 as pkg
from oracle import oracle as O
from orb_slam2_2021_b200 import synth

from . import frames_util as U

pytestmark = pytest.mark.gpu
GOLD = os.path.join(os.path.dirname(__file__), "golden", "frame_search.npz")


def oracle_search(cur, last, has, world, pdesc, obs, th, mono, ori):
    return O.search_by_projection(U.view(O.FrameView, cur), U.view(O.FrameView, last), has, world, pdesc, obs, th, mono, ori)


def gpu_search(m, cur, last, has, world, pdesc, obs, th, mono):
    return m.SearchByProjection(U.view(pkg.FrameView, cur), U.view(pkg.FrameView, last), pkg.MapPointView(has, world, pdesc, obs), th, mono)


@pytest.mark.parametrize("seed,tz,mono,ori,ur,n_last,clones", [
    (0, 0.0, False, True, True, 300, 30), (1, 0.5, False, True, True, 300, 30), (2, -0.5, False, True, False, 300, 30),
    (3, 0.5, True, False, True, 300, 30), (4, 0.0, False, False, False, 300, 30), (5, 0.0, False, True, True, 2000, 400),
    (6, 0.5, False, True, True, 8000, 1500), (7, 0.0, False, True, True, 300, 200)])
def test_search_by_projection_parity(seed, tz, mono, ori, ur, n_last, clones):
    cur, last, has, world, pdesc, obs = U.random_scenario(seed, n_last=n_last, n_extra=n_last // 2, tz=tz, clones=clones, with_uright=ur,
                                                          W=1241 if n_last > 1000 else 640, H=376 if n_last > 1000 else 400)
    m = pkg.Matcher(checkOrientation=ori)
    for th in (7, 15):
        ref, nref = oracle_search(cur, last, has, world, pdesc, obs, th, mono, ori)
        got, n = gpu_search(m, cur, last, has, world, pdesc, obs, th, mono)
        assert n == nref and np.array_equal(got, ref), f"th {th}: {n} vs {nref}, {(got != ref).sum()} keypoints differ"
        assert nref > 20


def test_search_golden_fixture():
    g = np.load(GOLD)
    m_on, m_off = pkg.Matcher(checkOrientation=True), pkg.Matcher(checkOrientation=False)
    for k in range(int(g["n_cases"])):
        seed, tz, th, mono, ori = [g[f"case{k}_{n}"].item() for n in ("seed", "tz", "th", "mono", "ori")]
        cur, last, has, world, pdesc, obs = U.random_scenario(int(seed), tz=float(tz))
        got, n = gpu_search(m_on if ori else m_off, cur, last, has, world, pdesc, obs, float(th), bool(mono))
        assert n == int(g[f"case{k}_nmatches"]) and np.array_equal(got, g[f"case{k}_match"])


def test_search_batch_of_pairs_and_ragged_sizes():
    """Many pairs in one call, different sizes, one with no map points and one with an empty current frame."""
    scen = [U.random_scenario(20 + k, n_last=[50, 700, 0, 333, 1200, 64][k], n_extra=[10, 300, 40, 0, 500, 7][k], clones=[5, 80, 0, 30, 100, 8][k],
                              tz=[0, 0.5, 0, -0.5, 0, 0.5][k]) for k in range(6)]
    # pair 5: empty current frame
    c5 = dict(scen[5][0])
    c5["keys_un"], c5["desc"], c5["u_right"] = c5["keys_un"][:0], c5["desc"][:0], c5["u_right"][:0]
    scen[5] = (c5,) + scen[5][1:]
    m = pkg.Matcher()
    res = m.SearchByProjection([U.view(pkg.FrameView, s[0]) for s in scen], [U.view(pkg.FrameView, s[1]) for s in scen],
                               [pkg.MapPointView(s[2], s[3], s[4], s[5]) for s in scen], 7)
    for k, s in enumerate(scen):
        ref, nref = oracle_search(*s, 7, False, True)
        assert res[k][1] == nref and np.array_equal(res[k][0], ref), k
    assert res[2][1] == 0 and res[5][1] == 0 and len(res[5][0]) == 0


def test_search_on_extracted_frames():
    """End to end with the extractor's own output: last frame = stereo frame k (map points from its stereo depth),
    current frame = the same scene shifted by a few pixels and re-extracted."""
    import cv2

    W, H, nf, nl, sf, ini, mn, fx, bf = synth.CONFIGS["arducam"]
    cx, cy = W / 2.0, H / 2.0
    L, R = synth.stereo_pair(W, H, 300)
    M = np.float32([[1, 0, 3.4], [0, 1, -1.7]])
    L2 = cv2.warpAffine(L, M, (W, H), flags=cv2.INTER_LINEAR, borderMode=cv2.BORDER_REFLECT_101)
    R2 = synth.synth_right(L2, 301)
    el, er = pkg.ORBextractor(nf, sf, nl, ini, mn), pkg.ORBextractor(nf, sf, nl, ini, mn)
    kl, dl = el(L)
    kr, dr = er(R)
    ur, dp = pkg.compute_stereo_matches(el, er, kl, dl, kr, dr, bf, bf / fx)
    kl2, dl2 = el(L2)
    kr2, dr2 = er(R2)
    ur2, _ = pkg.compute_stereo_matches(el, er, kl2, dl2, kr2, dr2, bf, bf / fx)
    z = np.where(dp > 0, dp, 1).astype(np.float32)
    world = np.stack([(kl["x"] - cx) * z / fx, (kl["y"] - cy) * z / fx, z], 1).astype(np.float32)
    has = (dp > 0).astype(np.uint8)
    obs = (np.arange(len(kl)) % 3 != 0).astype(np.uint8)
    sfs = el.GetScaleFactors()
    common = dict(fx=fx, fy=fx, cx=cx, cy=cy, mbf=bf, mb=bf / fx, bounds=(0.0, float(W), 0.0, float(H)), scale_factors=sfs)
    Tlw = np.hstack([np.eye(3, dtype=np.float32), np.zeros((3, 1), np.float32)])
    Tcw = Tlw.copy()
    Tcw[:, 3] = [0.02, -0.01, 0.0]
    cur = dict(keys_un=kl2, desc=dl2, u_right=ur2, Tcw=Tcw, **common)
    last = dict(keys_un=kl, desc=None, u_right=None, Tcw=Tlw, **common)
    m = pkg.Matcher()
    for th in (7, 15):
        ref, nref = oracle_search(cur, last, has, world, dl, obs, th, False, True)
        got, n = gpu_search(m, cur, last, has, world, dl, obs, th, False)
        assert n == nref and np.array_equal(got, ref)
    assert nref > 300  # most map points are found again


def test_grid_and_area_parity():
    rng = np.random.default_rng(5)
    m = pkg.Matcher()
    for n, (W, H) in [(0, (640, 400)), (1, (640, 400)), (700, (640, 400)), (2000, (1241, 376)), (8000, (3840, 2160))]:
        keys = np.zeros(n, pkg.KP_DTYPE)
        keys["x"], keys["y"] = rng.uniform(-8, W + 8, n), rng.uniform(-8, H + 8, n)
        if n > 10:
            keys["x"][:5] = [0, W, W / 64 * 0.5, W / 64 * 1.5, W - 1e-3]
            keys["y"][:5] = [0, H, H / 48 * 0.5, H / 48 * 2.5, H - 1e-3]
            keys["x"][5:40], keys["y"][5:40] = 100.25, 50.75  # one crowded cell
        keys["octave"] = rng.integers(0, 8, n)
        args = (np.eye(4), 500, 500, W / 2, H / 2, 40, 0.08, (0.0, float(W), 0.0, float(H)), 1.2 ** np.arange(8))
        fo, fg = O.FrameView(keys, None, None, *args), pkg.FrameView(keys, None, None, *args)
        s0, i0 = O.assign_features_to_grid(fo)
        s1, i1 = m.AssignFeaturesToGrid(fg)
        assert np.array_equal(s0, s1) and np.array_equal(i0, i1)
        for q in range(25):
            x, y, r = rng.uniform(-30, W + 30), rng.uniform(-30, H + 30), float(rng.choice([0.5, 7, 25, 60, 400]))
            if q == 0:
                x, y, r = 100.0, 50.0, 5.0
            lv = [(-1, -1), (0, 3), (2, -1), (3, 3), (1, 0)][q % 5]
            assert np.array_equal(m.GetFeaturesInArea(fg, x, y, r, *lv), O.features_in_area(fo, x, y, r, *lv))


def test_stereo_from_rgbd_parity():
    rng = np.random.default_rng(6)
    W, H, nf, nl, sf, ini, mn, fx, bf = synth.CONFIGS["tum_mono"]
    img = synth.synth(W, H, 77)
    keys, _ = pkg.ORBextractor(2 * nf, sf, nl, ini, mn)(img)  # mpIniORBextractor: 2x features (src/Tracking.cc:131)
    assert len(keys) >= 2 * nf
    depth = rng.uniform(-0.5, 8, (H, W + 5)).astype(np.float32)[:, :W]  # strided rows
    depth[rng.random((H, W)) < 0.1] = np.nan
    depth[rng.random((H, W)) < 0.05] = 0
    m = pkg.Matcher()
    for unx in (None, (keys["x"] + rng.normal(0, 0.4, len(keys))).astype(np.float32)):
        ur, dp = m.ComputeStereoFromRGBD(keys, depth, 40.0, unx)
        our, odp = O.stereo_from_rgbd(keys, depth, 40.0, unx)
        assert np.array_equal(ur, our) and np.array_equal(dp, odp)
        assert (dp > 0).sum() > 0.5 * len(keys)
    e, _ = m.ComputeStereoFromRGBD(keys[:0], depth, 40.0)
    assert len(e) == 0


def test_frame_abi_errors():
    m = pkg.Matcher()
    keys = np.zeros(4, pkg.KP_DTYPE)
    keys["octave"] = 9
    args = (np.eye(4), 500, 500, 320, 200, 40, 0.08, (0.0, 640.0, 0.0, 400.0), 1.2 ** np.arange(8))
    cur = pkg.FrameView(keys, np.zeros((4, 32), np.uint8), None, *args)
    last = pkg.FrameView(keys, None, None, *args)
    pts = pkg.MapPointView(np.ones(4), np.ones((4, 3)), np.zeros((4, 32)), np.ones(4))
    with pytest.raises(pkg.OrbxError) as ei:  # octave 9 does not index an 8-level scale table
        m.SearchByProjection(cur, last, pts, 7)
    assert ei.value.code == pkg._lib.ERR_ARG
    with pytest.raises(TypeError):
        m.ComputeStereoFromRGBD(keys, np.zeros((4, 4), np.float64), 40.0)


# ---- N3: input staging (include/orbx_input.h) ---------------------------------------------------------------
from .test_input_oracle import rectification_maps  # noqa: E402


def oracle_stage(img, m1, m2, rgb, top, rows):
    r = img if m1 is None else O.remap_linear(img, m1, m2)
    r = r[top:top + rows]
    return O.cvt_gray(r, rgb_order=rgb) if r.ndim == 3 else np.ascontiguousarray(r)


@pytest.mark.parametrize("channels,rgb", [(1, False), (3, False), (3, True), (4, True), (4, False)])
def test_rectify_gray_parity(channels, rgb):
    rng = np.random.default_rng(channels)
    m = pkg.Matcher()
    for (W, H), top, rows in [((640, 480), 0, 400), ((1241, 376), 0, 376), ((333, 97), 5, 80), ((64, 20), 0, 20)]:
        img = rng.integers(0, 256, (H, W, channels), dtype=np.uint8)
        img[:, :, 0] = synth.synth(W, H, 9)
        if channels == 1:
            img = img[:, :, 0]
        m1, m2 = rectification_maps(W, H, seed=W)
        if W == 333:  # extreme maps: far outside, exact integers, ties
            m1, m2 = (m1 * 1.4 - 30).astype(np.float32), (np.round(m2 * 64) / 64 - 3).astype(np.float32)
        assert np.array_equal(m.RectifyGray(img, m1, m2, rgb=rgb, crop_top=top, crop_rows=rows), oracle_stage(img, m1, m2, rgb, top, rows))
        assert np.array_equal(m.RectifyGray(img, rgb=rgb, crop_top=top, crop_rows=rows), oracle_stage(img, None, None, rgb, top, rows))


def test_batch_from_raw_frames():
    """Raw distorted BGR frames in -> the same keypoints, descriptors and stereo matches as extracting the images
    that cv::remap + crop + cvtColor produce (oracle path), i.e. the reference's arducam pipeline end to end."""
    W, H, nf, nl, sf, ini, mn, fx, bf = synth.CONFIGS["arducam"]
    RH = 480  # raw height; the reference crops rows 0..400 after rectification (arducam_images.cpp:105-133)
    n = 3
    rng = np.random.default_rng(8)
    raws_l, raws_r = np.zeros((n, RH, W, 3), np.uint8), np.zeros((n, RH, W, 3), np.uint8)
    for k in range(n):
        l, r = synth.stereo_pair(W, RH, 500 + 2 * k)
        for c in range(3):
            raws_l[k, :, :, c] = np.clip(l.astype(np.int32) + rng.integers(-6, 7, l.shape), 0, 255)
            raws_r[k, :, :, c] = np.clip(r.astype(np.int32) + rng.integers(-6, 7, r.shape), 0, 255)
    ml, mr = rectification_maps(W, RH, seed=1), rectification_maps(W, RH, seed=2)
    b = pkg.StereoFrameBatch(nf, sf, nl, ini, mn, H, W, n, stereo=True)
    b.set_rectification(0, ml[0], ml[1], crop_top=0)
    b.set_rectification(1, mr[0], mr[1], crop_top=0)
    b.upload_raw(raws_l, raws_r, rgb=False)
    b.run(n, bf, bf / fx)
    out = b.download(n)
    for k in range(n):
        gl, gr = oracle_stage(raws_l[k], ml[0], ml[1], False, 0, H), oracle_stage(raws_r[k], mr[0], mr[1], False, 0, H)
        assert np.array_equal(b.level(k, 0, 0), gl) and np.array_equal(b.level(k, 1, 0), gr)
        el, er = O.OracleExtractor(nf, sf, nl, ini, mn), O.OracleExtractor(nf, sf, nl, ini, mn)
        okl, odl = el(gl)
        okr, odr = er(gr)
        our, odp = O.stereo_match(el, er, okl, odl, okr, odr, bf, bf / fx)
        nl_, nr_ = out["n_left"][k], out["n_right"][k]
        assert nl_ == len(okl) and nr_ == len(okr)
        assert np.array_equal(out["kps_left"][k][:nl_]["x"], okl["x"]) and np.array_equal(out["desc_left"][k][:nl_], odl)
        assert np.array_equal(out["desc_right"][k][:nr_], odr)
        assert np.array_equal(out["u_right"][k][:nl_], our) and np.array_equal(out["depth"][k][:nl_], odp)
    # gray-only path (no maps): mono engine fed with RGBA frames
    mono = pkg.StereoFrameBatch(nf, sf, nl, ini, mn, H, W, n, stereo=False)
    mono.set_rectification(0, None, None, crop_top=40)
    rgba = np.concatenate([raws_l, np.full((n, RH, W, 1), 255, np.uint8)], axis=3)
    mono.upload_raw(rgba, rgb=True)
    mono.run(n)
    for k in range(n):
        assert np.array_equal(mono.level(k, 0, 0), oracle_stage(rgba[k], None, None, True, 40, H))


# ---- N1, second overload: local-map SearchByProjection (src/ORBmatcher.cc:45-127) ------------------------------
@pytest.mark.parametrize("seed,th,ur,occ,n_points,clones", [(0, 1.0, True, True, 400, 40), (1, 3.0, True, False, 400, 40), (2, 5.0, False, True, 400, 40),
                                                            (3, 1.0, False, False, 400, 200), (4, 1.0, True, True, 3000, 500), (5, 5.0, True, True, 6000, 1500)])
def test_search_local_points_parity(seed, th, ur, occ, n_points, clones):
    big = n_points > 1000
    F, occupied, valid, proj, level, vc, pdesc, obs = U.local_map_scenario(seed, n_points=n_points, n_extra=n_points // 2, clones=clones, with_uright=ur,
                                                                          W=1241 if big else 640, H=376 if big else 400)
    occupied = occupied if occ else None
    m = pkg.Matcher()
    for ratio in (0.8, 0.6, 10.0):
        ref, nref = O.search_local_points(U.view(O.FrameView, F), occupied, valid, proj, level, vc, pdesc, obs, th, ratio)
        got, n = m.SearchLocalPoints(U.view(pkg.FrameView, F), pkg.TrackPointView(valid, proj, level, vc, pdesc, obs), th, occupied, ratio)
        assert n == nref and np.array_equal(got, ref), f"ratio {ratio}: {n} vs {nref}, {(got != ref).sum()} keypoints differ"
    assert nref > 30


def test_search_local_points_batch():
    scen = [U.local_map_scenario(30 + k, n_points=[60, 900, 0, 350][k], n_extra=[10, 200, 30, 0][k], clones=[6, 100, 0, 35][k]) for k in range(4)]
    m = pkg.Matcher()
    res = m.SearchLocalPoints([U.view(pkg.FrameView, s[0]) for s in scen], [pkg.TrackPointView(*s[2:]) for s in scen], 3.0,
                              [s[1] if k % 2 == 0 else None for k, s in enumerate(scen)])
    for k, s in enumerate(scen):
        ref, nref = O.search_local_points(U.view(O.FrameView, s[0]), s[1] if k % 2 == 0 else None, *s[2:], 3.0, 0.8)
        assert res[k][1] == nref and np.array_equal(res[k][0], ref), k
    assert res[2][1] == 0


# ---- device-resident chaining: SearchByProjection on what a batch engine left on the device ---------------------------
def _resident_scene(n_pairs=3):
    """A batch of stereo frames (every image pair twice: frame 2p = current, 2p+1 = last of pair p, same image, identity
    poses), its downloaded results, and per pair the host views + map points made from the last frame's stereo depth."""
    W, H, nf, nl, sf, ini, mn, fx, bf = synth.CONFIGS["kitti"]
    imgs = [synth.stereo_pair(W, H, 900 + p) for p in range(n_pairs)]
    left = np.stack([imgs[p // 2][0] for p in range(2 * n_pairs)])
    right = np.stack([imgs[p // 2][1] for p in range(2 * n_pairs)])
    batch = pkg.StereoFrameBatch(nf, sf, nl, ini, mn, H, W, 2 * n_pairs, stereo=True)
    out = batch.process(left, right, bf, bf / fx)
    cx, cy = W / 2.0, H / 2.0
    sfs = (np.float32(sf) ** np.arange(nl)).astype(np.float32)
    I = np.hstack([np.eye(3, dtype=np.float32), np.zeros((3, 1), np.float32)])
    rng = np.random.default_rng(3)
    views, dummies, pts = [], [], []
    for f in range(2 * n_pairs):
        n = int(out["n_left"][f])
        k, d, u = out["kps_left"][f, :n], out["desc_left"][f, :n], out["u_right"][f, :n]
        views.append(pkg.FrameView(k, d, u, I, fx, fx, cx, cy, bf, bf / fx, (0.0, float(W), 0.0, float(H)), sfs))
        dummies.append(pkg.FrameView(np.zeros(n, pkg.KP_DTYPE), np.zeros((n, 32), np.uint8), None, I, fx, fx, cx, cy, bf, bf / fx,
                                     (0.0, float(W), 0.0, float(H)), sfs))
    total = sum(views[2 * p + 1].N for p in range(n_pairs))
    has, obs = np.zeros(total, np.uint8), np.zeros(total, np.uint8)
    pos, pdesc = np.zeros((total, 3), np.float32), np.zeros((total, 32), np.uint8)
    at = 0
    for p in range(n_pairs):
        f = 2 * p + 1
        n = views[f].N
        z = out["depth"][f, :n]
        k = out["kps_left"][f, :n]
        s = slice(at, at + n)
        has[s] = (z > 0) & (rng.random(n) < 0.8)
        obs[s] = rng.random(n) < 0.6
        zz = np.where(z > 0, z, 1).astype(np.float32)
        pos[s] = np.stack([(k["x"] - cx) * zz / fx, (k["y"] - cy) * zz / fx, zz], 1)
        pdesc[s] = out["desc_left"][f, :n]
        pts.append((at, n))
        at += n
    return batch, views, dummies, (has, pos, pdesc, obs), pts


def test_search_by_projection_on_device_resident_frames():
    """orbx_search_by_projection_resident: the frames' keypoints / descriptors / uRight read where the batch engine left
    them == the host-array form (which the tests above pin against the oracle and the reference), for map point arrays
    that are one concatenation (copied straight from the caller's memory) and for separate ones (repacked)."""
    n_pairs = 3
    batch, views, dummies, (has, pos, pdesc, obs), pts = _resident_scene(n_pairs)
    m = pkg.Matcher()
    cur = [views[2 * p] for p in range(n_pairs)]
    last = [views[2 * p + 1] for p in range(n_pairs)]
    concat = [pkg.MapPointView(has[a:a + n], pos[a:a + n], pdesc[a:a + n], obs[a:a + n]) for a, n in pts]
    separate = [pkg.MapPointView(has[a:a + n].copy(), pos[a:a + n].copy(), pdesc[a:a + n].copy(), obs[a:a + n].copy()) for a, n in pts]
    assert concat[1].has_point.ctypes.data == concat[0].has_point.ctypes.data + pts[0][1]  # really one concatenation
    want = m.SearchByProjection(cur, last, separate, 15)
    assert all(nm > 200 for _, nm in want)
    o_ref, o_n = O.search_by_projection(O.FrameView(cur[0].keys_un, cur[0].desc, cur[0].u_right, cur[0].Tcw, *[cur[0].c.fx, cur[0].c.fy, cur[0].c.cx, cur[0].c.cy, cur[0].c.mbf, cur[0].c.mb],
                                                     (0.0, cur[0].c.max_x, 0.0, cur[0].c.max_y), cur[0].scale_factors),
                                        O.FrameView(last[0].keys_un, last[0].desc, last[0].u_right, last[0].Tcw, *[last[0].c.fx, last[0].c.fy, last[0].c.cx, last[0].c.cy, last[0].c.mbf, last[0].c.mb],
                                                     (0.0, last[0].c.max_x, 0.0, last[0].c.max_y), last[0].scale_factors),
                                        separate[0].has_point, separate[0].world_pos, separate[0].point_desc, separate[0].observed, 15, False, True)
    assert o_n == want[0][1] and np.array_equal(o_ref, want[0][0])
    srcs_c = [(batch, 2 * p, 0) for p in range(n_pairs)]
    srcs_l = [(batch, 2 * p + 1, 0) for p in range(n_pairs)]
    for pts_views in (concat, separate):
        # the views handed over with the sources carry no keypoints or descriptors at all (dummies): nothing of them is read
        got = m.SearchByProjection([dummies[2 * p] for p in range(n_pairs)], [dummies[2 * p + 1] for p in range(n_pairs)], pts_views, 15,
                                   cur_src=srcs_c, last_src=srcs_l)
        for (a, na), (b, nb) in zip(got, want):
            assert na == nb and np.array_equal(a, b)
    # mixed: current frame resident, last frame from host arrays; and host views with concatenated map points
    got = m.SearchByProjection([dummies[2 * p] for p in range(n_pairs)], last, concat, 15, cur_src=srcs_c, last_src=[None] * n_pairs)
    assert all(na == nb and np.array_equal(a, b) for (a, na), (b, nb) in zip(got, want))
    got = m.SearchByProjection(cur, last, concat, 15)
    assert all(na == nb and np.array_equal(a, b) for (a, na), (b, nb) in zip(got, want))
    with pytest.raises(pkg.OrbxError):  # a frame index outside the batch
        m.SearchByProjection(dummies[0], dummies[1], concat[0], 15, cur_src=(batch, 99, 0), last_src=(batch, 1, 0))


def test_search_local_points_on_device_resident_frames():
    """orbx_search_local_points_resident == the host-array form: local map = the frames' own keypoints reprojected (same
    pixel, predicted level = octave), so that most points find their keypoint."""
    n_pairs = 3
    batch, views, dummies, _, _ = _resident_scene(n_pairs)
    rng = np.random.default_rng(11)
    frames, dums, srcs, pts, occ = [], [], [], [], []
    for p in range(n_pairs):
        f = 2 * p
        v = views[f]
        n = v.N
        k = v.keys_un
        sel = rng.permutation(n)[: n // 2]
        proj = np.stack([k["x"][sel] + rng.normal(0, 0.7, len(sel)), k["y"][sel] + rng.normal(0, 0.7, len(sel)),
                         np.where(v.u_right[sel] >= 0, v.u_right[sel], -1)], 1).astype(np.float32)
        pts.append(pkg.TrackPointView(rng.random(len(sel)) < 0.9, proj, k["octave"][sel], rng.uniform(0.99, 1.0, len(sel)).astype(np.float32),
                                      v.desc[sel], rng.random(len(sel)) < 0.7))
        occ.append((rng.random(n) < 0.1).astype(np.uint8))
        frames.append(v)
        dums.append(dummies[f])
        srcs.append((batch, f, 0))
    m = pkg.Matcher()
    want = m.SearchLocalPoints(frames, pts, 3, occupied=occ)
    got = m.SearchLocalPoints(dums, pts, 3, occupied=occ, frame_src=srcs)
    assert all(nm > 200 for _, nm in want)
    for (a, na), (b, nb) in zip(got, want):
        assert na == nb and np.array_equal(a, b)
