"""CPU tests of the oracle for the frame-level rows: OpenCV arithmetic pinned against
cv2, the C++ oracle cross-checked against a second pure-Python restatement, golden fixtures."""
import os

import cv2
import numpy as np
import pytest

from oracle import oracle as O

from . import frames_util as U

GOLD = os.path.join(os.path.dirname(__file__), "golden", "frame_search.npz")


def test_projection_arithmetic_matches_cv2_gemm():
    """C-6: Rcw * x3Dw + tcw == cv::gemm(R, x, 1, t, 1) on CV_32F (small-matrix path), bit for bit."""
    rng = np.random.default_rng(0)
    for _ in range(3000):
        T = np.zeros((3, 4), np.float32)
        T[:, :3] = rng.normal(0, 1, (3, 3))
        T[:, 3] = rng.normal(0, 2, 3)
        x = rng.normal(0, 10, 3).astype(np.float32)
        ref = cv2.gemm(np.ascontiguousarray(T[:, :3]), x.reshape(3, 1), 1.0, np.ascontiguousarray(T[:, 3:4]), 1.0)[:, 0]
        assert np.array_equal(ref, O.project(T, x))


def test_tlc_matches_cv2_gemm():
    """C-7: twc = -Rcw.t() * tcw (general gemm path, GEMM_1_T), tlc = Rlw * twc + tlw."""
    rng = np.random.default_rng(1)
    for _ in range(3000):
        Tc, Tl = np.zeros((3, 4), np.float32), np.zeros((3, 4), np.float32)
        Tc[:, :3], Tc[:, 3] = rng.normal(0, 1, (3, 3)), rng.normal(0, 2, 3)
        Tl[:, :3], Tl[:, 3] = rng.normal(0, 1, (3, 3)), rng.normal(0, 2, 3)
        twc = cv2.gemm(np.ascontiguousarray(Tc[:, :3]), np.ascontiguousarray(Tc[:, 3:4]), -1.0, None, 0.0, flags=cv2.GEMM_1_T)
        tlc = cv2.gemm(np.ascontiguousarray(Tl[:, :3]), twc, 1.0, np.ascontiguousarray(Tl[:, 3:4]), 1.0)
        assert np.float32(O.tlc_z(Tc, Tl)) == tlc[2, 0]


def test_grid_and_area_against_python():
    rng = np.random.default_rng(2)
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
        start, items = O.assign_features_to_grid(fr)
        grid, inv_w, inv_h = U.py_grid(keys, bounds)
        flat = [i for ix in range(64) for iy in range(48) for i in grid[ix][iy]]
        assert list(items) == flat
        for ix in (0, 13, 63):
            for iy in (0, 20, 47):
                c = ix * 48 + iy
                assert list(items[start[c]:start[c + 1]]) == grid[ix][iy]
        for q in range(40):
            x, y, r = rng.uniform(-30, W + 30), rng.uniform(-30, H + 30), rng.choice([0.5, 7, 25, 60, 400])
            lv = [(-1, -1), (0, 3), (2, -1), (3, 3), (1, 0)][q % 5]
            got = O.features_in_area(fr, x, y, r, *lv)
            assert list(got) == U.py_features_in_area(keys, grid, inv_w, inv_h, bounds, x, y, r, *lv)


@pytest.mark.parametrize("seed,tz,mono,ori,ur", [(0, 0.0, False, True, True), (1, 0.5, False, True, True), (2, -0.5, False, True, False),
                                                 (3, 0.5, True, False, True), (4, 0.0, False, False, False)])
def test_search_by_projection_against_python(seed, tz, mono, ori, ur):
    cur, last, has, world, pdesc, obs = U.random_scenario(seed, n_last=160, n_extra=120, tz=tz, with_uright=ur)
    th = 7 if seed % 2 == 0 else 15
    got, n = O.search_by_projection(U.view(O.FrameView, cur), U.view(O.FrameView, last), has, world, pdesc, obs, th, mono, ori)
    ref, nref = U.py_search_by_projection(cur, last, has, world, pdesc, obs, th, mono, ori)
    assert n == nref and np.array_equal(got, ref)
    assert n > 20  # the scenario does produce matches


def test_search_exercises_blocking_and_rotation_removal():
    """The scenarios reach the order-dependent branches: blocked keypoints and histogram removals."""
    cur, last, has, world, pdesc, obs = U.random_scenario(7, n_last=300, n_extra=100, clones=60)
    fc, fl = U.view(O.FrameView, cur), U.view(O.FrameView, last)
    a, na = O.search_by_projection(fc, fl, has, world, pdesc, obs, 7, False, True)
    b, nb = O.search_by_projection(fc, fl, has, world, pdesc, np.zeros_like(obs), 7, False, True)
    c, nc = O.search_by_projection(fc, fl, has, world, pdesc, obs, 7, False, False)
    assert not np.array_equal(a, b)  # Observations() > 0 changes who keeps a contested keypoint
    assert nc > na                   # the rotation check removes matches


def test_rgbd_against_numpy():
    rng = np.random.default_rng(3)
    H, W, n = 48, 64, 500
    depth = rng.uniform(-0.5, 6, (H, W)).astype(np.float32)
    depth[rng.random((H, W)) < 0.1] = np.nan
    keys = np.zeros(n, O.KP_DTYPE)
    keys["x"], keys["y"] = rng.uniform(0, W - 0.01, n), rng.uniform(0, H - 0.01, n)
    keys["x"][:3], keys["y"][:3] = [-0.5, W + 2, 3], [4, 4, H + 1]  # (-0.5 truncates to 0: inside)
    ux = (keys["x"] + rng.normal(0, 0.3, n)).astype(np.float32)
    for unx in (None, ux):
        ur, dp = O.stereo_from_rgbd(keys, depth, 40.0, unx)
        for i in range(n):
            v, u = int(keys["y"][i]), int(keys["x"][i])
            d = depth[v, u] if 0 <= v < H and 0 <= u < W else np.float32(-1)
            if d > 0:
                xx = keys["x"][i] if unx is None else unx[i]
                assert dp[i] == d and ur[i] == np.float32(xx - np.float32(40.0) / d)
            else:
                assert dp[i] == -1 and ur[i] == -1


def test_golden_search_fixture():
    """Committed outputs of the oracle (made by tests/golden/make_frame_golden.py); guards the oracle against drift."""
    g = np.load(GOLD)
    for k in range(int(g["n_cases"])):
        seed, tz, th, mono, ori = [g[f"case{k}_{n}"].item() for n in ("seed", "tz", "th", "mono", "ori")]
        cur, last, has, world, pdesc, obs = U.random_scenario(int(seed), tz=float(tz))
        got, n = O.search_by_projection(U.view(O.FrameView, cur), U.view(O.FrameView, last), has, world, pdesc, obs, float(th), bool(mono),
                                        bool(ori))
        assert n == int(g[f"case{k}_nmatches"]) and np.array_equal(got, g[f"case{k}_match"])


@pytest.mark.parametrize("seed,th,ur,occ", [(0, 1.0, True, True), (1, 3.0, True, False), (2, 5.0, False, True), (3, 1.0, False, False)])
def test_search_local_points_against_python(seed, th, ur, occ):
    F, occupied, valid, proj, level, vc, pdesc, obs = U.local_map_scenario(seed, n_points=200, n_extra=100, clones=30, with_uright=ur)
    occupied = occupied if occ else None
    got, n = O.search_local_points(U.view(O.FrameView, F), occupied, valid, proj, level, vc, pdesc, obs, th, 0.8)
    ref, nref = U.py_search_local_points(F, occupied, valid, proj, level, vc, pdesc, obs, th, 0.8)
    assert n == nref and np.array_equal(got, ref)
    assert n > 30
    # the ratio test and the blocking rule are both exercised
    loose, nl = O.search_local_points(U.view(O.FrameView, F), occupied, valid, proj, level, vc, pdesc, obs, th, 10.0)
    assert nl > n
    free, _ = O.search_local_points(U.view(O.FrameView, F), occupied, valid, proj, level, vc, pdesc, np.zeros_like(obs), th, 0.8)
    assert not np.array_equal(free, got)
