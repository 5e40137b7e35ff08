"""Parity of the CUDA path (through the C-ABI) against the CPU oracle and the golden vectors.

Bars (BASELINE.json north_star): keypoint x/y/octave/response/size/class_id, descriptor bytes and the
match validity mask bit-exact; angle <= 1e-4 deg; mvuRight <= 1e-4 px; mvDepth <= 1e-5 relative.
In practice every field is bit-identical, and the tests assert that stronger statement where the
shared math header guarantees it.
"""
import glob
import os
import threading

import numpy as np
import pytest

import orb_slam2_2021_b200 as pkg
from oracle import oracle as O
from orb_slam2_2021_b200 import synth

pytestmark = pytest.mark.gpu

ANGLE_TOL_DEG = 1e-4
URIGHT_TOL_PX = 1e-4
DEPTH_RTOL = 1e-5


def assert_kps_equal(got, ref, what=""):
    assert len(got) == len(ref), (what, len(got), len(ref))
    for f in ("x", "y", "size", "response", "octave", "class_id"):
        assert np.array_equal(got[f], ref[f]), (what, f, int((got[f] != ref[f]).sum()))
    assert np.abs(got["angle"] - ref["angle"]).max(initial=0) <= ANGLE_TOL_DEG, what
    assert np.array_equal(got["angle"], ref["angle"]), (what, "angle not bit-identical")


def assert_stereo_equal(ur, dp, our, odp, what=""):
    assert np.array_equal(ur >= 0, our >= 0), (what, "validity mask")
    assert np.abs(ur - our).max(initial=0) <= URIGHT_TOL_PX, what
    assert np.all(np.abs(dp - odp) <= DEPTH_RTOL * np.abs(odp)), what
    assert np.array_equal(ur, our) and np.array_equal(dp, odp), (what, "not bit-identical")


def oracle_pair(cfg, left, right):
    W, H, nf, nl, sf, ini, mn, fx, bf = cfg
    ol, orr = O.OracleExtractor(nf, sf, nl, ini, mn), O.OracleExtractor(nf, sf, nl, ini, mn)
    kl, dl = ol(left)
    kr, dr = orr(right)
    return ol, orr, kl, dl, kr, dr


@pytest.mark.parametrize("name,seeds", [("tum_mono", range(6)), ("arducam", range(4)), ("kitti", range(6)),
                                        ("euroc", range(4))])
def test_extract_and_stereo_handles(name, seeds):
    cfg = synth.CONFIGS[name]
    W, H, nf, nl, sf, ini, mn, fx, bf = cfg
    exl, exr = pkg.ORBextractor(nf, sf, nl, ini, mn), pkg.ORBextractor(nf, sf, nl, ini, mn)
    for seed in seeds:
        left, right = synth.stereo_pair(W, H, 100 + seed)
        ol, orr, okl, odl, okr, odr = oracle_pair(cfg, left, right)
        kl, dl = exl(left)
        kr, dr = exr(right)
        assert_kps_equal(kl, okl, f"{name}/{seed}/L")
        assert_kps_equal(kr, okr, f"{name}/{seed}/R")
        assert np.array_equal(dl, odl) and np.array_equal(dr, odr)
        if seed == seeds[0]:
            for l, lev in enumerate(exl.mvImagePyramid):
                assert np.array_equal(lev, ol.level(l)), f"pyramid level {l}"
        if bf > 0:
            for mb in (bf / fx, 0.0):  # steady state, and the reference's uninitialised-mb first frame
                ur, dp = pkg.compute_stereo_matches(exl, exr, kl, dl, kr, dr, bf, mb)
                our, odp = O.stereo_match(ol, orr, okl, odl, okr, odr, bf, mb)
                assert_stereo_equal(ur, dp, our, odp, f"{name}/{seed}/mb={mb}")
                assert (ur >= 0).sum() > 100


GOLD = sorted(glob.glob(os.path.join(os.path.dirname(__file__), "golden", "*_seed*.npz")))


@pytest.mark.parametrize("path", GOLD, ids=[os.path.basename(p) for p in GOLD])
def test_golden_vectors(path):
    """CUDA path vs vectors produced by the cv2-driven model (real OpenCV kernels)."""
    g = np.load(path)
    name, seed = str(g["config"]), int(g["seed"])
    W, H, nf, nl, sf, ini, mn, fx, bf = synth.CONFIGS[name]
    left, right = synth.stereo_pair(W, H, seed)
    exl = pkg.ORBextractor(nf, sf, nl, ini, mn)
    kl, dl = exl(left)
    assert kl.tobytes() == g["kps_left"].tobytes() and np.array_equal(dl, g["desc_left"])
    if "kps_right" in g:
        exr = pkg.ORBextractor(nf, sf, nl, ini, mn)
        kr, dr = exr(right)
        assert kr.tobytes() == g["kps_right"].tobytes() and np.array_equal(dr, g["desc_right"])
        ur, dp = pkg.compute_stereo_matches(exl, exr, kl, dl, kr, dr, bf, bf / fx)
        assert np.array_equal(ur, g["u_right"]) and np.array_equal(dp, g["depth"])
        ur, dp = pkg.compute_stereo_matches(exl, exr, kl, dl, kr, dr, bf, 0.0)
        assert np.array_equal(ur, g["u_right_mb0"]) and np.array_equal(dp, g["depth_mb0"])


def test_batch_matches_oracle_per_frame():
    cfg = synth.CONFIGS["kitti"]
    W, H, nf, nl, sf, ini, mn, fx, bf = cfg
    n = 6
    L, R = synth.stereo_batch(W, H, n, seed0=40)
    b = pkg.StereoFrameBatch(nf, sf, nl, ini, mn, H, W, n, stereo=True)
    out = b.process(L, R, mbf=bf, mb=bf / fx)
    tot_l = tot_r = tot_m = 0
    for f in range(n):
        ol, orr, okl, odl, okr, odr = oracle_pair(cfg, L[f], R[f])
        nl_, nr_ = out["n_left"][f], out["n_right"][f]
        assert_kps_equal(out["kps_left"][f][:nl_], okl, f"frame {f} L")
        assert_kps_equal(out["kps_right"][f][:nr_], okr, f"frame {f} R")
        assert np.array_equal(out["desc_left"][f][:nl_], odl) and np.array_equal(out["desc_right"][f][:nr_], odr)
        our, odp = O.stereo_match(ol, orr, okl, odl, okr, odr, bf, bf / fx)
        assert_stereo_equal(out["u_right"][f][:nl_], out["depth"][f][:nl_], our, odp, f"frame {f}")
        tot_l += nl_; tot_r += nr_; tot_m += int((our >= 0).sum())
        for l in (0, 3, 7):  # stage taps: pyramid, blur, FAST candidates in reference order
            assert np.array_equal(b.level(f, 0, l), ol.level(l))
            assert np.array_equal(b.level(f, 1, l, blurred=True), orr.level(l, blurred=True))
            gc, oc = b.candidates(f, 0, l), ol.candidates(l)
            assert len(gc) == len(oc) and all(np.array_equal(gc[k], oc[k]) for k in ("x", "y", "response"))
    assert b.stats(n) == (tot_l, tot_r, tot_m)


def test_mono_batch():
    W, H, nf, nl, sf, ini, mn, fx, bf = synth.CONFIGS["tum_mono"]
    imgs = np.stack([synth.synth(W, H, 60 + k) for k in range(3)])
    b = pkg.StereoFrameBatch(nf, sf, nl, ini, mn, H, W, 4, stereo=False)
    out = b.process(imgs)
    for f in range(3):
        ok, od = O.OracleExtractor(nf, sf, nl, ini, mn)(imgs[f])
        assert_kps_equal(out["kps_left"][f][:out["n_left"][f]], ok)
        assert np.array_equal(out["desc_left"][f][:out["n_left"][f]], od)


def adversarial_images(W, H):
    rng = np.random.default_rng(9)
    yy, xx = np.mgrid[0:H, 0:W]
    return {
        "flat": np.full((H, W), 90, np.uint8),
        "uniform_noise": rng.integers(0, 256, (H, W), dtype=np.uint8),
        "binary_noise": (rng.integers(0, 2, (H, W)) * 255).astype(np.uint8),
        "checker8": ((((xx // 8) + (yy // 8)) & 1) * 200 + 20).astype(np.uint8),
        "sparse_dots": np.where(rng.random((H, W)) < 0.002, 255, 40).astype(np.uint8),
        "gradient": ((xx * 255) // W).astype(np.uint8),
        "one_corner": np.pad(np.full((40, 40), 255, np.uint8), ((H // 2, H - H // 2 - 40), (W // 2, W - W // 2 - 40))),
    }


@pytest.mark.parametrize("kind", ["flat", "uniform_noise", "binary_noise", "checker8", "sparse_dots", "gradient", "one_corner"])
def test_adversarial_images(kind):
    """Ties everywhere (checker, binary noise), maximal candidate counts (noise), and no keypoints at all."""
    W, H, nf, nl, sf, ini, mn = 400, 300, 500, 6, 1.2, 20, 7
    img = adversarial_images(W, H)[kind]
    ok, od = O.OracleExtractor(nf, sf, nl, ini, mn)(img)
    ex = pkg.ORBextractor(nf, sf, nl, ini, mn)
    k, d = ex(img)
    assert_kps_equal(k, ok, kind)
    assert np.array_equal(d, od)
    if kind in ("flat", "gradient"):
        assert len(k) == 0
        exr = pkg.ORBextractor(nf, sf, nl, ini, mn)
        exr(img)
        ur, dp = pkg.compute_stereo_matches(ex, exr, k, d, k, d, 40.0, 0.1)  # empty match set: defined no-op
        assert len(ur) == 0


def test_stereo_identical_images_and_no_matches():
    """disparity 0 -> the 0.01 clamp (src/Frame.cc:674-678); unrelated right image -> few/no matches, median cut on tiny sets."""
    W, H, nf, nl, sf, ini, mn, fx, bf = synth.CONFIGS["arducam"]
    left = synth.synth(W, H, 5)
    for right in (left.copy(), synth.synth(W, H, 999)):
        exl, exr = pkg.ORBextractor(nf, sf, nl, ini, mn), pkg.ORBextractor(nf, sf, nl, ini, mn)
        kl, dl = exl(left)
        kr, dr = exr(right)
        ol, orr, okl, odl, okr, odr = oracle_pair(synth.CONFIGS["arducam"], left, right)
        ur, dp = pkg.compute_stereo_matches(exl, exr, kl, dl, kr, dr, bf, bf / fx)
        our, odp = O.stereo_match(ol, orr, okl, odl, okr, odr, bf, bf / fx)
        assert_stereo_equal(ur, dp, our, odp)


def test_uhd_single_frame():
    """3840x2160, 8000 features: ~31k candidates on level 0 stress the quadtree and the slot arrays."""
    cfg = synth.CONFIGS["uhd"]
    W, H, nf, nl, sf, ini, mn, fx, bf = cfg
    left, right = synth.stereo_pair(W, H, 2)
    ol, orr, okl, odl, okr, odr = oracle_pair(cfg, left, right)
    b = pkg.StereoFrameBatch(nf, sf, nl, ini, mn, H, W, 1, stereo=True)
    out = b.process(left[None], right[None], mbf=bf, mb=bf / fx)
    nl_, nr_ = out["n_left"][0], out["n_right"][0]
    assert_kps_equal(out["kps_left"][0][:nl_], okl, "uhd L")
    assert_kps_equal(out["kps_right"][0][:nr_], okr, "uhd R")
    assert np.array_equal(out["desc_left"][0][:nl_], odl)
    our, odp = O.stereo_match(ol, orr, okl, odl, okr, odr, bf, bf / fx)
    assert_stereo_equal(out["u_right"][0][:nl_], out["depth"][0][:nl_], our, odp, "uhd")


def test_abi_behaviour_on_device():
    W, H, nf, nl, sf, ini, mn, fx, bf = synth.CONFIGS["tum_mono"]
    img = synth.synth(W, H, 1)
    ex = pkg.ORBextractor(nf, sf, nl, ini, mn)
    k1, d1 = ex(img)
    # non-contiguous rows (stride > cols), like a cv::Mat ROI
    big = np.zeros((H, W + 37), np.uint8)
    big[:, :W] = img
    k2, d2 = ex(big[:, :W])
    assert k1.tobytes() == k2.tobytes() and np.array_equal(d1, d2)
    # the same handle on another image size rebuilds its device state
    small = synth.synth(400, 300, 2)
    k3, _ = ex(small)
    ok, _ = O.OracleExtractor(nf, sf, nl, ini, mn)(small)
    assert_kps_equal(k3, ok)
    # capacity error
    import ctypes as C
    kp = np.zeros(10, pkg.KP_DTYPE); ds = np.zeros((10, 32), np.uint8); n = C.c_int(0)
    rc = pkg.lib().orbx_extract(ex._h, img.ctypes.data_as(C.c_void_p), H, W, W, kp.ctypes.data_as(C.c_void_p),
                                ds.ctypes.data_as(C.c_void_p), 10, C.byref(n))
    assert rc == -2 and b"capacity" in pkg.lib().orbx_last_error()


def test_two_handles_two_threads():
    """Frame::Frame runs left and right extraction in two std::threads on two instances (src/Frame.cc:113-116)."""
    cfg = synth.CONFIGS["euroc"]
    W, H, nf, nl, sf, ini, mn, fx, bf = cfg
    left, right = synth.stereo_pair(W, H, 8)
    exs = [pkg.ORBextractor(nf, sf, nl, ini, mn) for _ in range(2)]
    res = [None, None]

    def work(i, im):
        for _ in range(5):
            res[i] = exs[i](im)

    ts = [threading.Thread(target=work, args=(i, im)) for i, im in enumerate((left, right))]
    [t.start() for t in ts]
    [t.join() for t in ts]
    ol, orr, okl, odl, okr, odr = oracle_pair(cfg, left, right)
    assert_kps_equal(res[0][0], okl)
    assert_kps_equal(res[1][0], okr)
    assert np.array_equal(res[0][1], odl) and np.array_equal(res[1][1], odr)


def test_full_size_batch_properties():
    """BASELINE-size properties where the oracle is too slow to check every frame: determinism,
    independence of a frame's result from its position in the batch, and count invariants."""
    W, H, nf, nl, sf, ini, mn, fx, bf = synth.CONFIGS["kitti"]
    base_l, base_r = synth.stereo_batch(W, H, 4, seed0=70)
    n = 64
    idx = np.random.default_rng(0).integers(0, 4, n)
    L, R = base_l[idx], base_r[idx]
    b = pkg.StereoFrameBatch(nf, sf, nl, ini, mn, H, W, n, stereo=True)
    out1 = {k: v.copy() for k, v in b.process(L, R, mbf=bf, mb=bf / fx).items()}
    out2 = b.process(L, R, mbf=bf, mb=bf / fx)
    first = {}
    for f in range(n):
        nl_ = out1["n_left"][f]
        assert nf <= nl_ <= nf + 2 * nl
        sig = (out1["kps_left"][f][:nl_].tobytes(), out1["desc_left"][f][:nl_].tobytes(),
               out1["u_right"][f][:nl_].tobytes(), out1["depth"][f][:nl_].tobytes())
        assert sig == (out2["kps_left"][f][:nl_].tobytes(), out2["desc_left"][f][:nl_].tobytes(),
                       out2["u_right"][f][:nl_].tobytes(), out2["depth"][f][:nl_].tobytes())
        assert first.setdefault(int(idx[f]), sig) == sig
        # descriptor rows are 4-byte aligned 32-byte records; every keypoint lies inside the 19-px edge at its level
        k = out1["kps_left"][f][:nl_]
        s = np.float32(1.2) ** k["octave"]
        assert np.all(k["x"] / s >= 18.99) and np.all(k["y"] / s >= 18.99)
        assert np.all(np.diff(k["octave"]) >= 0)  # levels concatenated 0..7


def test_device_resident_run_in_chunks():
    """upload + run + download (a batch of >= 32 frames runs as interleaved chunks on two streams) gives the oracle's
    results for every distinct frame, wherever the frame sits in the batch, including a ragged last chunk."""
    cfg = synth.CONFIGS["arducam"]
    W, H, nf, nl, sf, ini, mn, fx, bf = cfg
    base_l, base_r = synth.stereo_batch(W, H, 3, seed0=90)
    n = 37  # 2 chunks of 19 and 18
    idx = np.arange(n) % 3
    L, R = np.ascontiguousarray(base_l[idx]), np.ascontiguousarray(base_r[idx])
    b = pkg.StereoFrameBatch(nf, sf, nl, ini, mn, H, W, 40, stereo=True)
    b.upload(L, R)
    b.run(n, bf, bf / fx)
    out = b.download(n)
    assert b.launches_per_run > 15  # more than one chunk's worth of kernels
    ref = []
    for k in range(3):
        ol, orr, okl, odl, okr, odr = oracle_pair(cfg, base_l[k], base_r[k])
        ref.append((okl, odl, okr, odr) + O.stereo_match(ol, orr, okl, odl, okr, odr, bf, bf / fx))
    for f in range(n):
        okl, odl, okr, odr, our, odp = ref[idx[f]]
        nl_, nr_ = out["n_left"][f], out["n_right"][f]
        assert_kps_equal(out["kps_left"][f][:nl_], okl, f"frame {f} L")
        assert_kps_equal(out["kps_right"][f][:nr_], okr, f"frame {f} R")
        assert np.array_equal(out["desc_left"][f][:nl_], odl) and np.array_equal(out["desc_right"][f][:nr_], odr)
        assert_stereo_equal(out["u_right"][f][:nl_], out["depth"][f][:nl_], our, odp, f"frame {f}")


def test_streamed_process_matches_synchronous_calls():
    """process_async x4 back to back with changing inputs (the uploads of call k+1 overlap the kernels of call k), then
    wait(): every call's results equal those of the synchronous process() on the same inputs; a call with another frame
    count (different chunk layout) and other entry points in between drain the pipeline by themselves."""
    cfg = synth.CONFIGS["arducam"]
    W, H, nf, nl, sf, ini, mn, fx, bf = cfg
    base_l, base_r = synth.stereo_batch(W, H, 4, seed0=300)
    n = 70  # 8 chunks + split first chunk
    b = pkg.StereoFrameBatch(nf, sf, nl, ini, mn, H, W, n, stereo=True)
    ins = []
    for k in range(4):
        idx = (np.arange(n) + k) % 4
        pl, pr = pkg.PinnedArray((n, H, W), np.uint8), pkg.PinnedArray((n, H, W), np.uint8)
        pl.array[...] = base_l[idx]
        pr.array[...] = base_r[idx]
        ins.append((pl, pr))
    want = []
    for pl, pr in ins:
        o = b.process(pl.array, pr.array, bf, bf / fx)
        want.append({k: v.copy() for k, v in o.items()})
    outs = [b.alloc_outputs(n, pinned=True) for _ in range(4)]
    for k, (pl, pr) in enumerate(ins):  # four calls in flight back to back
        b.process_async(pl.array, pr.array, bf, bf / fx, outs[k])
    b.wait()
    res = outs
    for k in range(4):
        for key in ("n_left", "n_right"):
            assert np.array_equal(res[k][key], want[k][key]), (k, key)
        for f in range(n):
            a, c = res[k]["n_left"][f], res[k]["n_right"][f]
            assert np.array_equal(res[k]["kps_left"][f][:a], want[k]["kps_left"][f][:a]), (k, f)
            assert np.array_equal(res[k]["desc_left"][f][:a], want[k]["desc_left"][f][:a]), (k, f)
            assert np.array_equal(res[k]["kps_right"][f][:c], want[k]["kps_right"][f][:c]), (k, f)
            assert np.array_equal(res[k]["u_right"][f][:a], want[k]["u_right"][f][:a]), (k, f)
            assert np.array_equal(res[k]["depth"][f][:a], want[k]["depth"][f][:a]), (k, f)
    # in flight + another layout + another entry point
    b.process_async(ins[0][0].array, ins[0][1].array, bf, bf / fx, outs[0])
    small = b.process(ins[1][0].array[:20], ins[1][1].array[:20], bf, bf / fx)
    assert np.array_equal(small["n_left"], want[1]["n_left"][:20])
    assert np.array_equal(outs[0]["n_left"], want[0]["n_left"])  # the async call was drained before the layout changed
    b.process_async(ins[2][0].array, ins[2][1].array, bf, bf / fx, outs[1])
    nl_, nr_, nm_ = b.stats(n)  # drains
    assert nl_ == int(want[2]["n_left"].sum()) and np.array_equal(outs[1]["n_right"], want[2]["n_right"])


def test_cpp_shim_reference_signature(tmp_path):
    """include/ORBextractorGPU.h (the C++ class with the reference's constructor/operator()/getters) driven like
    Frame::Frame(stereo): two extractors on two std::threads, then ComputeStereoMatches — compiled here with g++."""
    import subprocess

    from orb_slam2_2021_b200._lib import KP_DTYPE

    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    exe = str(tmp_path / "shim_main")
    libdir = os.path.join(root, "orb_slam2_2021_b200")
    subprocess.run(["g++", "-std=c++17", "-O2", "-o", exe, os.path.join(root, "tests", "cpp", "shim_main.cpp"), f"-L{libdir}", "-lorbx",
                    f"-Wl,-rpath,{libdir}", "-lpthread"], check=True)
    cfg = synth.CONFIGS["kitti"]
    W, H, nf, nl, sf, ini, mn, fx, bf = cfg
    left, right = synth.stereo_pair(W, H, 321)
    left.tofile(tmp_path / "l.raw")
    right.tofile(tmp_path / "r.raw")
    out = tmp_path / "out.bin"
    subprocess.run([exe, str(tmp_path / "l.raw"), str(tmp_path / "r.raw"), str(H), str(W), str(nf), repr(float(sf)), str(nl), str(ini),
                    str(mn), repr(float(np.float32(bf))), repr(float(np.float32(bf / fx))), str(out)], check=True)
    raw = out.read_bytes()
    nL, nR = np.frombuffer(raw, np.int32, 2)
    off = 8
    kl = np.frombuffer(raw, KP_DTYPE, nL, off); off += 28 * nL
    dl = np.frombuffer(raw, np.uint8, 32 * nL, off).reshape(-1, 32); off += 32 * nL
    kr = np.frombuffer(raw, KP_DTYPE, nR, off); off += 28 * nR
    dr = np.frombuffer(raw, np.uint8, 32 * nR, off).reshape(-1, 32); off += 32 * nR
    ur = np.frombuffer(raw, np.float32, nL, off); off += 4 * nL
    dp = np.frombuffer(raw, np.float32, nL, off)
    ol, orr, okl, odl, okr, odr = oracle_pair(cfg, left, right)
    assert_kps_equal(kl, okl, "shim/L")
    assert_kps_equal(kr, okr, "shim/R")
    assert np.array_equal(dl, odl) and np.array_equal(dr, odr)
    our, odp = O.stereo_match(ol, orr, okl, odl, okr, odr, bf, bf / fx)
    assert_stereo_equal(ur, dp, our, odp, "shim/stereo")


@pytest.mark.parametrize("nf,sf,nl,ini,mn,W,H", [
    (500, 1.1, 8, 20, 7, 640, 480),      # fine pyramid
    (800, 1.5, 5, 20, 7, 752, 480),      # coarse pyramid: source windows of the resize span up to 7 bytes
    (600, 2.0, 4, 12, 7, 800, 600),      # the largest supported scale factor
    (300, 1.2, 1, 20, 7, 320, 240),      # a single level: no resize at all
    (3000, 1.2, 12, 30, 30, 1024, 768),  # 12 levels (the ABI maximum), minTh == iniTh
    (50, 1.3, 6, 254, 1, 640, 400),      # threshold extremes: the fallback to minTh fires in almost every cell
    (1500, 1.25, 8, 20, 7, 1023, 511),   # odd sizes: pitches with padding, ragged tiles and cells
    (1200, 1.2, 8, 20, 7, 1280, 720),    # width a multiple of 16: no row padding anywhere
    (9000, 1.2, 1, 12, 5, 800, 600),     # one level holding the whole quota: the quadtree's node list lives in global memory
    (20000, 1.2, 8, 10, 3, 1280, 720),   # every level above the shared-memory list; more features asked for than there are corners
])
def test_extractor_parameter_space(nf, sf, nl, ini, mn, W, H):
    """Other extractor parameters than the BASELINE configs, bit-exact against the oracle incl. every pyramid level."""
    cfg = (W, H, nf, nl, sf, ini, mn, 500.0, 40.0)
    left, right = synth.stereo_pair(W, H, 200 + nl)
    ex_l, ex_r = pkg.ORBextractor(nf, sf, nl, ini, mn), pkg.ORBextractor(nf, sf, nl, ini, mn)
    ol, orr, okl, odl, okr, odr = oracle_pair(cfg, left, right)
    kl, dl = ex_l(left)
    kr, dr = ex_r(right)
    assert_kps_equal(kl, okl, "L")
    assert_kps_equal(kr, okr, "R")
    assert np.array_equal(dl, odl) and np.array_equal(dr, odr)
    for l, lev in enumerate(ex_l.mvImagePyramid):
        assert np.array_equal(lev, ol.level(l)), f"pyramid level {l}"
    ur, dp = pkg.compute_stereo_matches(ex_l, ex_r, kl, dl, kr, dr, 40.0, 40.0 / 500.0)
    our, odp = O.stereo_match(ol, orr, okl, odl, okr, odr, 40.0, 40.0 / 500.0)
    assert_stereo_equal(ur, dp, our, odp, "stereo")
    # the same through the batch engine (other launch geometry: two images per launch)
    b = pkg.StereoFrameBatch(nf, sf, nl, ini, mn, H, W, 1, stereo=True)
    out = b.process(left[None], right[None], mbf=40.0, mb=40.0 / 500.0)
    assert_kps_equal(out["kps_left"][0][:out["n_left"][0]], okl, "batch L")
    assert np.array_equal(out["desc_right"][0][:out["n_right"][0]], odr)
    assert_stereo_equal(out["u_right"][0][:len(okl)], out["depth"][0][:len(okl)], our, odp, "batch stereo")


def test_scale_factor_above_two_is_rejected():
    with pytest.raises(pkg.OrbxError) as ei:
        pkg.ORBextractor(1000, 2.5, 4, 20, 7)
    assert ei.value.code == pkg._lib.ERR_ARG


@pytest.mark.gpu
def test_pinned_allocation_reports_its_numa_placement():
    """orbx_host_alloc places pinned buffers on the device's NUMA node where the host says which one that is; the
    buffer is usable either way, and orbx_set_tuning("numa_off", 1) switches the placement off."""
    a = pkg.PinnedArray((1 << 20,), np.uint8)
    a.array[...] = 7
    node, how = pkg.host_numa_node()
    assert node >= -1 and how in (0, 1, 2)
    assert (how == 0) == (node < 0) or how == 0  # no node -> plain allocation; a known node may still fall back to plain
    assert int(a.array.sum()) == 7 << 20
    pkg.set_tuning("numa_off", 1)
    try:
        b = pkg.PinnedArray((4096,), np.uint8)
        b.array[...] = 1
        assert pkg.host_numa_node() == (-1, 0)
    finally:
        pkg.set_tuning("numa_off", 0)
    with pytest.raises(pkg.OrbxError):
        pkg.set_tuning("no_such_knob", 1)


# ---- single-frame entry points: orbx_stereo_frame, orbx_stereo_match_resident (and the CUDA graph they replay) -------------
@pytest.mark.gpu
@pytest.mark.parametrize("name", ["kitti", "euroc", "arducam"])
def test_stereo_frame_equals_the_two_handle_path_and_the_oracle(name):
    """orbx_stereo_frame == orbx_extract x2 + orbx_stereo_match == oracle, over several frames (the first call runs the
    launches directly, the second captures the CUDA graph, later ones replay it; mbf/mb changes re-capture)."""
    W, H, nf, nl, sf, ini, mn, fx, bf = synth.CONFIGS[name]
    fl, fr = pkg.ORBextractor(nf, sf, nl, ini, mn), pkg.ORBextractor(nf, sf, nl, ini, mn)
    el, er = pkg.ORBextractor(nf, sf, nl, ini, mn), pkg.ORBextractor(nf, sf, nl, ini, mn)
    ol, orr = O.OracleExtractor(nf, sf, nl, ini, mn), O.OracleExtractor(nf, sf, nl, ini, mn)
    for i, (seed, mb) in enumerate(((41, bf / fx), (42, bf / fx), (43, bf / fx), (44, 0.0), (45, bf / fx))):
        left, right = synth.stereo_pair(W, H, seed)
        kl, dl, kr, dr, ur, dp = pkg.stereo_frame(fl, fr, left, right, bf, mb)
        k2l, d2l = el(left)
        k2r, d2r = er(right)
        u3, p3 = pkg.compute_stereo_matches_resident(el, er, len(k2l), len(k2r), bf, mb)  # nothing uploaded
        u2, p2 = pkg.compute_stereo_matches(el, er, k2l, d2l, k2r, d2r, bf, mb)
        assert kl.tobytes() == k2l.tobytes() and kr.tobytes() == k2r.tobytes()
        assert np.array_equal(dl, d2l) and np.array_equal(dr, d2r)
        assert ur.tobytes() == u2.tobytes() and dp.tobytes() == p2.tobytes()
        assert u3.tobytes() == u2.tobytes() and p3.tobytes() == p2.tobytes()
        if i in (0, 2, 3):
            okl, odl = ol(left)
            okr, odr = orr(right)
            our, odp = O.stereo_match(ol, orr, okl, odl, okr, odr, bf, mb)
            for f in ("x", "y", "size", "response", "octave", "class_id"):
                assert np.array_equal(kl[f], okl[f]), f
            assert np.abs(kl["angle"] - okl["angle"]).max() <= 1e-4
            assert np.array_equal(dl, odl) and np.array_equal(dr, odr)
            assert np.array_equal(ur >= 0, our >= 0) and np.abs(ur - our).max() <= 1e-4
        # the frame's pyramid is readable through both handles
        pl, pr_ = fl.mvImagePyramid, fr.mvImagePyramid
        ql, qr = el.mvImagePyramid, er.mvImagePyramid
        assert all(np.array_equal(a, b) for a, b in zip(pl, ql)) and all(np.array_equal(a, b) for a, b in zip(pr_, qr))


@pytest.mark.gpu
def test_single_frame_entry_points_state_and_strides():
    W, H, nf, nl, sf, ini, mn, fx, bf = synth.CONFIGS["arducam"]
    fl, fr = pkg.ORBextractor(nf, sf, nl, ini, mn), pkg.ORBextractor(nf, sf, nl, ini, mn)
    left, right = synth.stereo_pair(W, H, 9)
    ref_out = pkg.stereo_frame(fl, fr, left, right, bf, bf / fx)
    # strided views (a cv::Mat ROI) and page-locked tight buffers take different upload routes: same results
    wide = np.zeros((H, 2 * W + 16), np.uint8)
    wide[:, :W], wide[:, W + 16:] = left, right
    out2 = pkg.stereo_frame(fl, fr, wide[:, :W], wide[:, W + 16:], bf, bf / fx)
    pin_l, pin_r = pkg.PinnedArray((H, W), np.uint8), pkg.PinnedArray((H, W), np.uint8)
    pin_l.array[...], pin_r.array[...] = left, right
    out3 = pkg.stereo_frame(fl, fr, pin_l.array, pin_r.array, bf, bf / fx)
    for a, b, c in zip(ref_out, out2, out3):
        assert a.tobytes() == b.tobytes() == c.tobytes()
    # after orbx_stereo_frame the handles' own engines hold no frame: the two-handle match refuses
    with pytest.raises(pkg.OrbxError) as ei:
        pkg.compute_stereo_matches(fl, fr, ref_out[0], ref_out[1], ref_out[2], ref_out[3], bf, bf / fx)
    assert ei.value.code == pkg._lib.ERR_STATE
    # resident match: refuses counts that are not the extractors' own, and after an uploading match
    el, er = pkg.ORBextractor(nf, sf, nl, ini, mn), pkg.ORBextractor(nf, sf, nl, ini, mn)
    kl, dl = el(left)
    kr, dr = er(right)
    with pytest.raises(pkg.OrbxError):
        pkg.compute_stereo_matches_resident(el, er, len(kl) - 1, len(kr), bf, bf / fx)
    pkg.compute_stereo_matches(el, er, kl[:100], dl[:100], kr, dr, bf, bf / fx)
    with pytest.raises(pkg.OrbxError):
        pkg.compute_stereo_matches_resident(el, er, len(kl), len(kr), bf, bf / fx)
    # different extractor parameters on the two sides are refused
    with pytest.raises(pkg.OrbxError):
        pkg.stereo_frame(fl, pkg.ORBextractor(nf + 1, sf, nl, ini, mn), left, right, bf, bf / fx)
    # size change rebuilds the frame engine
    W2, H2 = synth.CONFIGS["tum_mono"][:2]
    l2, r2 = synth.stereo_pair(W2, H2, 3)
    a = pkg.stereo_frame(fl, fr, l2, r2, bf, bf / fx)
    b = pkg.stereo_frame(fl, fr, l2, r2, bf, bf / fx)
    assert all(x.tobytes() == y.tobytes() for x, y in zip(a, b)) and len(a[0]) > 300


@pytest.mark.gpu
def test_one_launch_pyramid_equals_the_per_level_launches():
    """orbx_pyramid_chain_kernel (every tile of every level in one launch, chained by readiness counters) against the
    per-level launches and the oracle: levels byte for byte, on a batch large enough that tiles of many images and
    levels are in flight at once (forced with the pyr_chain tuning key; small batches take it by default)."""
    W, H, nf, nl, sf, ini, mn, fx, bf = synth.CONFIGS["euroc"]
    n = 40
    left, right = synth.stereo_batch(W, H, 4, seed0=90)
    idx = np.arange(n) % 4
    L, R = np.ascontiguousarray(left[idx]), np.ascontiguousarray(right[idx])
    outs, levels = {}, {}
    for mode in (2, 1):
        pkg.set_tuning("pyr_chain", mode)
        try:
            b = pkg.StereoFrameBatch(nf, sf, nl, ini, mn, H, W, n, stereo=True)
            for rep in range(3):  # repeated runs: the counters are cleared per launch
                outs[mode] = b.process(L, R, bf, bf / fx)
            levels[mode] = [b.level(f, s, l) for f in (0, 17, 39) for s in (0, 1) for l in range(nl)]
        finally:
            pkg.set_tuning("pyr_chain", 0)
    for a, c in zip(levels[1], levels[2]):
        assert np.array_equal(a, c)
    for side in ("left", "right"):
        assert np.array_equal(outs[1]["n_" + side], outs[2]["n_" + side])
    for f in range(n):  # (rows beyond a frame's count are whatever the engine's fresh buffers held)
        a, c = int(outs[1]["n_left"][f]), int(outs[1]["n_right"][f])
        for k, m in (("kps_left", a), ("desc_left", a), ("u_right", a), ("depth", a), ("kps_right", c), ("desc_right", c)):
            assert outs[1][k][f, :m].tobytes() == outs[2][k][f, :m].tobytes(), (k, f)
    o = O.OracleExtractor(nf, sf, nl, ini, mn)
    o(left[1])
    for l in range(nl):
        assert np.array_equal(b.level(1, 0, l), o.level(l))
    # odd geometry: scale factor 2 (the WIDE resize variant) and a portrait-ish image, single handle (one-launch form)
    e = pkg.ORBextractor(800, 2.0, 3, 20, 7)
    oo = O.OracleExtractor(800, 2.0, 3, 20, 7)
    img = synth.synth(701, 397, 4)
    k1, d1 = e(img)
    k2, d2 = oo(img)
    assert np.array_equal(d1, d2) and k1["x"].tobytes() == k2["x"].tobytes()
    for l, lev in enumerate(e.mvImagePyramid):
        assert np.array_equal(lev, oo.level(l))


@pytest.mark.gpu
def test_stereo_frame_handle_lifetimes():
    """The frame engine lives in the left handle; right-hand partners only look into it. Destroying or re-pairing handles in
    any order leaves no dangling view: a partner whose engine is gone reports ORBX_ERR_STATE instead of reading freed memory."""
    W, H, nf, nl, sf, ini, mn, fx, bf = synth.CONFIGS["arducam"]
    left, right = synth.stereo_pair(W, H, 12)
    mk = lambda: pkg.ORBextractor(nf, sf, nl, ini, mn)  # noqa: E731
    l, r1, r2 = mk(), mk(), mk()
    a = pkg.stereo_frame(l, r1, left, right, bf, bf / fx)
    b = pkg.stereo_frame(l, r2, left, right, bf, bf / fx)  # re-paired: r1 no longer reads l's engine
    assert all(x.tobytes() == y.tobytes() for x, y in zip(a, b))
    with pytest.raises(pkg.OrbxError) as ei:
        r1.mvImagePyramid
    assert ei.value.code == pkg._lib.ERR_STATE
    assert np.array_equal(r2.mvImagePyramid[0], right)
    del l  # the owner goes first
    with pytest.raises(pkg.OrbxError):
        r2.mvImagePyramid
    k, _ = r2(right)  # the handle itself is still a working extractor
    assert len(k) == len(a[2]) and np.array_equal(r2.mvImagePyramid[0], right)
    l2, r3 = mk(), mk()
    pkg.stereo_frame(l2, r3, left, right, bf, bf / fx)
    del r3  # the partner goes first: the owner must not touch it afterwards
    c = pkg.stereo_frame(l2, r2, left, right, bf, bf / fx)
    assert all(x.tobytes() == y.tobytes() for x, y in zip(a, c))
    del l2, r2, r1


@pytest.mark.gpu
@pytest.mark.parametrize("case", ["kitti", "wide_cells", "flat"])
def test_fast_cta_layouts_agree(case):
    """orbx_fast_kernel runs one warp per cell; how many warps share a CTA (tuning key fast_cells: 1 = default, 2, 4, 8) is
    a layout choice that changes the generated code a lot (one warp per CTA keeps cell-only values on the uniform datapath),
    never the result. Every layout against the oracle, on the 80-byte tile (KITTI), the 112-byte tile (a level that is one
    59-px cell wide) and an image where most cells repeat at minThFAST."""
    if case == "kitti":
        W, H, nf, nl, sf, ini, mn = synth.CONFIGS["kitti"][:7]
        img = synth.stereo_pair(W, H, 41)[0]
    elif case == "wide_cells":
        nf, sf, nl, ini, mn = 500, 1.2, 4, 20, 7
        img = synth.stereo_pair(150, 140, 3)[0]
    else:
        W, H, nf, nl, sf, ini, mn = synth.CONFIGS["arducam"][:7]
        rng = np.random.default_rng(1)
        img = np.clip(128 + rng.normal(0, 3.0, (H, W)), 0, 255).astype(np.uint8)
        img[100:140, 200:260] += 14
    ok, od = O.OracleExtractor(nf, sf, nl, ini, mn)(img)
    try:
        for cells in (1, 2, 4, 8):
            pkg.set_tuning("fast_cells", cells)
            k, d = pkg.ORBextractor(nf, sf, nl, ini, mn)(img)
            assert_kps_equal(k, ok, f"{case} fast_cells={cells}")
            assert np.array_equal(d, od), (case, cells)
    finally:
        pkg.set_tuning("fast_cells", 0)
