"""ctypes binding of oracle/_ref/libref_orbextractor.so: the REFERENCE'S OWN ORB_SLAM2::ORBextractor
(/root/reference/src/ORBextractor.cc compiled where it lies, see oracle/Makefile and oracle/ref_wrap.cpp) over the
OpenCV stand-in of oracle/cvshim/. TEST INFRASTRUCTURE ONLY: tests/, smoke() and bench.py's CPU legs may import it.

/root/reference exists only in the build container; the built library is git-ignored but travels to the GPU box.
`available()` says whether the library can be used here; `build()` (re)builds it where the sources exist.
"""
import ctypes as C
import os
import subprocess

import numpy as np

from .oracle import KP_DTYPE, _ptr, _u8

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "_ref", "libref_orbextractor.so")
_REF_SRC = "/root/reference/src/ORBextractor.cc"


def build(force=False):
    """Compile the reference translation unit against the shim (only where /root/reference exists)."""
    if os.path.exists(_REF_SRC) and (force or not os.path.exists(_SO) or not os.path.exists(os.path.join(_HERE, "_ref", "libref_frame_gpu.so"))):
        subprocess.check_call(["make", "-C", _HERE, "-B" if force else "-s", "_ref"])
    return _SO if os.path.exists(_SO) else None


def available():
    return os.path.exists(_SO) or os.path.exists(_REF_SRC)


_lib = None


def lib():
    global _lib
    if _lib is None:
        if build() is None:
            raise RuntimeError("oracle/_ref/libref_orbextractor.so is missing and /root/reference is not here to build it")
        L = C.CDLL(_SO)
        vp, f32p, i32p = C.c_void_p, C.POINTER(C.c_float), C.POINTER(C.c_int)
        L.ref_create.restype = vp
        L.ref_create.argtypes = [C.c_int, C.c_float, C.c_int, C.c_int, C.c_int]
        L.ref_destroy.argtypes = [vp]
        L.ref_tables.argtypes = [vp, f32p, f32p, f32p, f32p]
        L.ref_extract.argtypes = [vp, vp, C.c_int, C.c_int, C.c_int, vp, vp, C.c_int]
        L.ref_level_size.argtypes = [vp, C.c_int, i32p, i32p]
        L.ref_level_copy.argtypes = [vp, C.c_int, vp, C.c_int]
        L.ref_stereo_match.argtypes = [vp, vp, vp, vp, C.c_int, vp, vp, C.c_int, C.c_float, C.c_float, vp, vp]
        L.ref_stereo_from_rgbd.argtypes = [vp, vp, C.c_int, vp, C.c_int, C.c_int, C.c_int, C.c_float, vp, vp]
        L.ref_assign_features_to_grid.argtypes = [vp, C.c_int, C.c_float, C.c_float, C.c_float, C.c_float, vp, vp]
        L.ref_features_in_area.argtypes = [vp, C.c_int, C.c_float, C.c_float, C.c_float, C.c_float, C.c_float, C.c_float,
                                           C.c_float, C.c_int, C.c_int, vp]
        L.ref_voc_load.restype = vp
        L.ref_voc_load.argtypes = [C.c_char_p]
        L.ref_voc_destroy.argtypes = [vp]
        L.ref_voc_words.argtypes = [vp]
        L.ref_compute_bow.argtypes = [vp, vp, C.c_int] + [vp] * 7
        _lib = L
    return _lib


class RefExtractor:
    """ORB_SLAM2::ORBextractor itself (include/ORBextractor.h:56-100): same constructor, operator() as __call__."""

    def __init__(self, nfeatures, scale_factor, nlevels, ini_th_fast, min_th_fast):
        self.nfeatures, self.nlevels = nfeatures, nlevels
        self._h = lib().ref_create(nfeatures, scale_factor, nlevels, ini_th_fast, min_th_fast)

    def __del__(self):
        if getattr(self, "_h", None):
            lib().ref_destroy(self._h)
            self._h = None

    def tables(self):
        f = [np.zeros(self.nlevels, np.float32) for _ in range(4)]
        lib().ref_tables(self._h, *[a.ctypes.data_as(C.POINTER(C.c_float)) for a in f])
        return dict(scale=f[0], inv_scale=f[1], sigma2=f[2], inv_sigma2=f[3])

    def __call__(self, image):
        img = _u8(image)
        cap = self.nfeatures + 64 * self.nlevels + 64
        kps = np.zeros(cap, KP_DTYPE)
        desc = np.zeros((cap, 32), np.uint8)
        n = lib().ref_extract(self._h, _ptr(img), img.shape[0], img.shape[1], img.strides[0], _ptr(kps), _ptr(desc), cap)
        if n < 0:
            raise RuntimeError("ref_extract: capacity too small")
        return kps[:n].copy(), desc[:n].copy()

    def level(self, l):
        r, c = C.c_int(), C.c_int()
        if lib().ref_level_size(self._h, l, C.byref(r), C.byref(c)) != 0:
            raise IndexError(l)
        out = np.zeros((r.value, c.value), np.uint8)
        lib().ref_level_copy(self._h, l, _ptr(out), out.strides[0])
        return out


def stereo_match(ex_l, ex_r, kps_l, desc_l, kps_r, desc_r, mbf, mb):
    """Frame::ComputeStereoMatches (src/Frame.cc:522-700) as the reference's own object code, on the reference's Frame with
    the members the stereo constructor sets (src/Frame.cc:98-125). ex_l / ex_r: RefExtractors that extracted this frame's
    images last (their pyramids are read). Returns (mvuRight, mvDepth)."""
    kps_l, kps_r = np.ascontiguousarray(kps_l, KP_DTYPE), np.ascontiguousarray(kps_r, KP_DTYPE)
    desc_l, desc_r = np.ascontiguousarray(desc_l, np.uint8), np.ascontiguousarray(desc_r, np.uint8)
    n = len(kps_l)
    ur, dp = np.zeros(n, np.float32), np.zeros(n, np.float32)
    rc = lib().ref_stereo_match(ex_l._h, ex_r._h, _ptr(kps_l), _ptr(desc_l), n, _ptr(kps_r), _ptr(desc_r), len(kps_r),
                                mbf, mb, _ptr(ur), _ptr(dp))
    if rc != 0:
        raise RuntimeError("ref_stereo_match failed")
    return ur, dp


def stereo_from_rgbd(keys, keys_un, depth_map, mbf):
    """Frame::ComputeStereoFromRGBD (src/Frame.cc:702-723) on the reference's Frame. Returns (mvuRight, mvDepth)."""
    keys, keys_un = np.ascontiguousarray(keys, KP_DTYPE), np.ascontiguousarray(keys_un, KP_DTYPE)
    d = np.ascontiguousarray(depth_map, np.float32)
    n = len(keys)
    ur, dp = np.zeros(n, np.float32), np.zeros(n, np.float32)
    lib().ref_stereo_from_rgbd(_ptr(keys), _ptr(keys_un), n, _ptr(d), d.shape[0], d.shape[1], d.strides[0], mbf, _ptr(ur), _ptr(dp))
    return ur, dp


def assign_features_to_grid(keys_un, bounds):
    """Frame::AssignFeaturesToGrid (src/Frame.cc:279-294) -> (cell_start[64*48+1], items); cell (ix, iy) at ix*48+iy."""
    keys_un = np.ascontiguousarray(keys_un, KP_DTYPE)
    start = np.zeros(64 * 48 + 1, np.int32)
    items = np.zeros(max(len(keys_un), 1), np.int32)
    k = lib().ref_assign_features_to_grid(_ptr(keys_un), len(keys_un), *[float(b) for b in bounds], _ptr(start), _ptr(items))
    return start, items[:k]


def features_in_area(keys_un, bounds, x, y, r, min_level=-1, max_level=-1):
    """Frame::GetFeaturesInArea (src/Frame.cc:376-433) on a frame whose grid the reference builds from keys_un."""
    keys_un = np.ascontiguousarray(keys_un, KP_DTYPE)
    out = np.zeros(max(len(keys_un), 1), np.int32)
    k = lib().ref_features_in_area(_ptr(keys_un), len(keys_un), *[float(b) for b in bounds], x, y, r, min_level, max_level,
                                   _ptr(out))
    return out[:k]


def write_vocabulary_text(path, k, L, parent, is_leaf, desc, weight):
    """The text format ORBVocabulary::loadFromTextFile reads (TemplatedVocabulary.h:1351-1439): "k L scoring weighting", then
    one line per node "parent isLeaf d0 .. d31 weight". No newline after the last node (the loader's eof loop would
    otherwise append an empty node to the root). Scoring L1_NORM = 0, weighting TF_IDF = 0, as in ORBvoc.txt."""
    lines = [f"{k} {L} 0 0"]
    for i in range(len(parent)):
        lines.append(f"{int(parent[i])} {int(is_leaf[i])} " + " ".join(str(int(b)) for b in desc[i]) + f" {float(weight[i])!r}")
    with open(path, "w") as f:
        f.write("\n".join(lines))


class RefVocabulary:
    """ORB_SLAM2::ORBVocabulary (the vendored DBoW2 TemplatedVocabulary<FORB>) loaded by its own loadFromTextFile."""

    def __init__(self, path):
        self._h = lib().ref_voc_load(os.fsencode(path))
        if not self._h:
            raise ValueError("loadFromTextFile failed")

    def __del__(self):
        h, self._h = getattr(self, "_h", None), None
        if h:
            lib().ref_voc_destroy(h)

    @property
    def n_words(self):
        return lib().ref_voc_words(self._h)

    def transform(self, desc):
        """Frame::ComputeBoW (levelsup = 4) -> (word_ids, word_vals, fv_nodes, fv_start, fv_feats), like oracle.Vocabulary."""
        desc = np.ascontiguousarray(desc, dtype=np.uint8).reshape(-1, 32)
        n = len(desc)
        wi, wv = np.zeros(max(n, 1), np.int32), np.zeros(max(n, 1), np.float64)
        fn, fs, ff = np.zeros(max(n, 1), np.int32), np.zeros(n + 1, np.int32), np.zeros(max(n, 1), np.int32)
        nw, nf = C.c_int(), C.c_int()
        lib().ref_compute_bow(self._h, _ptr(desc), n, _ptr(wi), _ptr(wv), C.addressof(nw), _ptr(fn), _ptr(fs), _ptr(ff),
                              C.addressof(nf))
        return wi[: nw.value], wv[: nw.value], fn[: nf.value], fs[: nf.value + 1], ff[: fs[nf.value]]


def _padded(frame):
    """A copy of an oracle FrameView's C struct whose scale-factor table has 16 entries (the reference's Frame gets a
    std::vector; the wrapper copies 16 floats)."""
    from .oracle import _OrcFrame

    sf = np.ones(16, np.float32)
    k = min(16, len(frame.scale_factors))
    sf[:k] = frame.scale_factors[:k]
    sf[k:] = frame.scale_factors[k - 1]
    c = _OrcFrame()
    C.memmove(C.byref(c), C.byref(frame.c), C.sizeof(_OrcFrame))
    c.scale_factors = sf.ctypes.data
    return c, sf


def search_by_projection(cur, last, has_point, world_pos, point_desc, observed, th, b_mono=False, check_orientation=True):
    """ORBmatcher::SearchByProjection(CurrentFrame, LastFrame, th, bMono), the reference's own src/ORBmatcher.cc:1348-1491 on
    its own Frame objects (cur / last: oracle.FrameView) -> (match_of_cur, nmatches), like oracle.search_by_projection."""
    from .oracle import _OrcFrame, _OrcMapPoints

    hp = np.ascontiguousarray(has_point, dtype=np.uint8)
    wp = np.ascontiguousarray(world_pos, dtype=np.float32).reshape(-1, 3)
    pd = np.ascontiguousarray(point_desc, dtype=np.uint8).reshape(-1, 32)
    ob = np.ascontiguousarray(observed, dtype=np.uint8)
    m = _OrcMapPoints(hp.ctypes.data, wp.ctypes.data, pd.ctypes.data, ob.ctypes.data)
    (cc, _k1), (lc, _k2) = _padded(cur), _padded(last)
    out = np.full(max(cur.c.n, 1), -1, np.int32)
    L = lib()
    L.ref_search_by_projection.argtypes = [C.POINTER(_OrcFrame), C.POINTER(_OrcFrame), C.POINTER(_OrcMapPoints), C.c_float,
                                           C.c_int, C.c_int, C.c_void_p]
    n = L.ref_search_by_projection(C.byref(cc), C.byref(lc), C.byref(m), th, int(b_mono), int(check_orientation), _ptr(out))
    return out[: cur.c.n], n


def search_local_points(frame, occupied, valid, proj, level, view_cos, point_desc, observed, th, nnratio=0.8):
    """ORBmatcher::SearchByProjection(F, vpMapPoints, th), the reference's own src/ORBmatcher.cc:45-127 -> (match_of_key, n)."""
    from .oracle import _OrcFrame, _OrcTrackPoints

    va = np.ascontiguousarray(valid, dtype=np.uint8)
    pr = np.ascontiguousarray(proj, dtype=np.float32).reshape(-1, 3)
    lv = np.ascontiguousarray(level, dtype=np.int32)
    vc = np.ascontiguousarray(view_cos, dtype=np.float32)
    pd = np.ascontiguousarray(point_desc, dtype=np.uint8).reshape(-1, 32)
    ob = np.ascontiguousarray(observed, dtype=np.uint8)
    oc = None if occupied is None else np.ascontiguousarray(occupied, dtype=np.uint8)
    t = _OrcTrackPoints(len(va), va.ctypes.data, pr.ctypes.data, lv.ctypes.data, vc.ctypes.data, pd.ctypes.data, ob.ctypes.data)
    fc, _keep = _padded(frame)
    out = np.full(max(frame.c.n, 1), -1, np.int32)
    L = lib()
    L.ref_search_local_points.argtypes = [C.POINTER(_OrcFrame), C.c_void_p, C.POINTER(_OrcTrackPoints), C.c_float, C.c_float,
                                          C.c_void_p]
    n = L.ref_search_local_points(C.byref(fc), None if oc is None else _ptr(oc), C.byref(t), th, nnratio, _ptr(out))
    return out[: frame.c.n], n


# ---- the reference's stereo Frame constructor (src/Frame.cc:80-152), on the CPU extractors and on the GPU shim -------------
_FRAME_ARGS = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_float, C.c_float,
               C.c_float, C.c_int] + [C.c_void_p] * 12


def _frame_stereo(fn, h_l, h_r, nlevels, left, right, K4, bf, th_depth, mb_preset, capacity):
    left, right = _u8(left), _u8(right)
    assert left.shape == right.shape and left.strides[0] == right.strides[0]
    K4 = np.ascontiguousarray(K4, np.float32)
    cap = int(capacity)
    keys, keys_r, keys_un = (np.zeros(cap, KP_DTYPE) for _ in range(3))
    desc, desc_r = np.zeros((cap, 32), np.uint8), np.zeros((cap, 32), np.uint8)
    ur, dp = np.zeros(cap, np.float32), np.zeros(cap, np.float32)
    gstart, gitems = np.zeros(64 * 48 + 1, np.int32), np.zeros(cap, np.int32)
    counts, misc, tabs = np.zeros(4, np.int32), np.zeros(16, np.float32), np.zeros(4 * nlevels, np.float32)
    n = fn(h_l, h_r, _ptr(left), _ptr(right), left.shape[0], left.shape[1], left.strides[0], _ptr(K4), bf, th_depth, mb_preset, cap,
           _ptr(keys), _ptr(keys_r), _ptr(keys_un), _ptr(desc), _ptr(desc_r), _ptr(ur), _ptr(dp), _ptr(gstart), _ptr(gitems),
           _ptr(counts), _ptr(misc), _ptr(tabs))
    if n < 0:
        raise RuntimeError(f"frame_stereo failed ({n})")
    nr = int(counts[1])
    names = ("mb", "mnMinX", "mnMaxX", "mnMinY", "mnMaxY", "mfGridElementWidthInv", "mfGridElementHeightInv", "fx", "fy", "cx", "cy",
             "invfx", "invfy", "mbf", "mThDepth", "mfScaleFactor")
    return dict(N=n, mvKeys=keys[:n].copy(), mvKeysRight=keys_r[:nr].copy(), mvKeysUn=keys_un[:n].copy(),
                mDescriptors=desc[:n].copy(), mDescriptorsRight=desc_r[:nr].copy(), mvuRight=ur[:n].copy(), mvDepth=dp[:n].copy(),
                mGridStart=gstart, mGridItems=gitems[: gstart[-1]].copy(), counts=counts,
                statics=dict(zip(names, (float(v) for v in misc))), scaleTables=tabs.reshape(4, nlevels))


def frame_stereo(ex_l, ex_r, left, right, K4, bf, th_depth, mb_preset, capacity=20000):
    """Frame::Frame(imLeft, imRight, ...) — the reference's stereo constructor, unchanged, on the reference's CPU extractors
    (ex_l / ex_r: RefExtractor). mb_preset: what Frame::mb holds when ComputeStereoMatches reads it (ref_frame_entry.inc)."""
    L = lib()
    L.ref_frame_stereo.argtypes = _FRAME_ARGS
    return _frame_stereo(L.ref_frame_stereo, ex_l._h, ex_r._h, ex_l.nlevels, left, right, K4, bf, th_depth, mb_preset, capacity)


_SO_GPU = os.path.join(_HERE, "_ref", "libref_frame_gpu.so")
_lib_gpu = None


def gpu_available():
    return os.path.exists(_SO_GPU) or os.path.exists(_REF_SRC)


def lib_gpu():
    """oracle/_ref/libref_frame_gpu.so: the reference's Frame.cc (INTEGRATION.md's patch, nothing else) on liborbx.so."""
    global _lib_gpu
    if _lib_gpu is None:
        if build() is None or not os.path.exists(_SO_GPU):
            raise RuntimeError("oracle/_ref/libref_frame_gpu.so is missing and /root/reference is not here to build it")
        import orb_slam2_2021_b200 as pkg

        pkg.lib()  # liborbx.so first: the drop-in library resolves its orbx_* symbols against the same loaded copy
        L = C.CDLL(_SO_GPU)
        L.refgpu_create.restype = C.c_void_p
        L.refgpu_create.argtypes = [C.c_int, C.c_float, C.c_int, C.c_int, C.c_int]
        L.refgpu_destroy.argtypes = [C.c_void_p]
        L.refgpu_extract.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_int]
        L.refgpu_frame_stereo.argtypes = _FRAME_ARGS
        _lib_gpu = L
    return _lib_gpu


class GpuShimExtractor:
    """ORB_SLAM2::ORBextractorGPU (include/ORBextractorGPU.h) as the reference's Frame.cc sees it: constructed with the
    reference's arguments, called through the ORBX_WITH_OPENCV overload with the reference's operator() signature."""

    def __init__(self, nfeatures, scale_factor, nlevels, ini_th_fast, min_th_fast):
        self.nfeatures, self.nlevels = nfeatures, nlevels
        self._h = lib_gpu().refgpu_create(nfeatures, scale_factor, nlevels, ini_th_fast, min_th_fast)
        if not self._h:
            raise RuntimeError("refgpu_create failed")

    def __del__(self):
        h, self._h = getattr(self, "_h", None), None
        if h:
            lib_gpu().refgpu_destroy(h)

    def __call__(self, image):
        img = _u8(image)
        cap = self.nfeatures + 64 * self.nlevels + 64
        kps, desc = np.zeros(cap, KP_DTYPE), np.zeros((cap, 32), np.uint8)
        n = lib_gpu().refgpu_extract(self._h, _ptr(img), img.shape[0], img.shape[1], img.strides[0], _ptr(kps), _ptr(desc), cap)
        if n < 0:
            raise RuntimeError(f"refgpu_extract failed ({n})")
        return kps[:n].copy(), desc[:n].copy()


def frame_stereo_gpu(ex_l, ex_r, left, right, K4, bf, th_depth, mb_preset, capacity=20000):
    """The same constructor of the same Frame.cc with INTEGRATION.md's patch, on GpuShimExtractors (needs a CUDA device)."""
    return _frame_stereo(lib_gpu().refgpu_frame_stereo, ex_l._h, ex_r._h, ex_l.nlevels, left, right, K4, bf, th_depth, mb_preset,
                         capacity)
