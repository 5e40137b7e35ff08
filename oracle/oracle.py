"""ctypes binding of the CPU oracle (oracle/liborb_oracle.so). TEST INFRASTRUCTURE ONLY.

Only tests/, __graft_entry__.smoke() and the cpu_baseline / --impl reference legs of bench.py may
import this module. The product package orb_slam2_2021_b200 never does.
"""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "liborb_oracle.so")

KP_DTYPE = np.dtype(
    [("x", "<f4"), ("y", "<f4"), ("size", "<f4"), ("angle", "<f4"), ("response", "<f4"), ("octave", "<i4"),
     ("class_id", "<i4")]
)
assert KP_DTYPE.itemsize == 28


def build(force=False):
    """Compile the oracle with the committed Makefile (g++, no external dependency)."""
    if force or not os.path.exists(_SO):
        subprocess.check_call(["make", "-C", _HERE, "-B" if force else "-s", "liborb_oracle.so"])
    return _SO


_lib = None


def lib():
    global _lib
    if _lib is None:
        build()
        L = C.CDLL(_SO)
        u8p, f32p, i32p, vp = C.POINTER(C.c_uint8), C.POINTER(C.c_float), C.POINTER(C.c_int), C.c_void_p
        L.orc_create.restype = vp
        L.orc_create.argtypes = [C.c_int, C.c_float, C.c_int, C.c_int, C.c_int]
        L.orc_destroy.argtypes = [vp]
        L.orc_tables.argtypes = [vp, f32p, f32p, f32p, f32p, i32p, i32p]
        L.orc_extract.argtypes = [vp, vp, C.c_int, C.c_int, C.c_int, vp, vp, C.c_int]
        L.orc_level_size.argtypes = [vp, C.c_int, i32p, i32p]
        L.orc_level_copy.argtypes = [vp, C.c_int, vp, C.c_int]
        L.orc_level_blur_copy.argtypes = [vp, C.c_int, vp, C.c_int]
        L.orc_level_candidates.argtypes = [vp, C.c_int, vp, C.c_int]
        L.orc_stereo_match.argtypes = [vp, vp, vp, vp, C.c_int, vp, vp, C.c_int, C.c_float, C.c_float, vp, vp]
        L.orc_resize_linear.argtypes = [vp, C.c_int, C.c_int, C.c_int, vp, C.c_int, C.c_int, C.c_int]
        L.orc_gaussian7.argtypes = [vp, C.c_int, C.c_int, C.c_int, vp, C.c_int]
        L.orc_fast9.argtypes = [vp, C.c_int, C.c_int, C.c_int, C.c_int, vp, C.c_int]
        L.orc_fast_atan2.restype = C.c_float
        L.orc_fast_atan2.argtypes = [C.c_float, C.c_float]
        L.orc_sincos.argtypes = [C.c_float, f32p, f32p]
        L.orc_sincos_vs_libm.restype = C.c_long
        L.orc_sincos_vs_libm.argtypes = [C.c_uint32, C.c_uint32, C.c_uint32]
        L.orc_hamming256.argtypes = [vp, vp]
        L.orc_reflect101.argtypes = [C.c_int, C.c_int]
        L.orc_distribute.argtypes = [vp, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, vp, C.c_int]
        _lib = L
    return _lib


def _ptr(a):
    return a.ctypes.data_as(C.c_void_p)


def _u8(img):
    img = np.ascontiguousarray(img, dtype=np.uint8)
    assert img.ndim == 2
    return img


class OracleExtractor:
    """Mirror of ORB_SLAM2::ORBextractor (include/ORBextractor.h:56-100) over the CPU oracle."""

    def __init__(self, nfeatures, scale_factor, nlevels, ini_th_fast, min_th_fast):
        self.nfeatures, self.nlevels = nfeatures, nlevels
        self._h = lib().orc_create(nfeatures, scale_factor, nlevels, ini_th_fast, min_th_fast)
        if not self._h:
            raise ValueError("orc_create failed")

    def __del__(self):
        if getattr(self, "_h", None):
            lib().orc_destroy(self._h)
            self._h = None

    def tables(self):
        n = self.nlevels
        f = [np.zeros(n, np.float32) for _ in range(4)]
        q = np.zeros(n, np.int32)
        um = np.zeros(16, np.int32)
        fp = [a.ctypes.data_as(C.POINTER(C.c_float)) for a in f]
        lib().orc_tables(self._h, *fp, q.ctypes.data_as(C.POINTER(C.c_int)), um.ctypes.data_as(C.POINTER(C.c_int)))
        return dict(scale=f[0], inv_scale=f[1], sigma2=f[2], inv_sigma2=f[3], quotas=q, umax=um)

    def __call__(self, image):
        """operator()(image, mask, keypoints, descriptors) -> (keypoints[KP_DTYPE], descriptors[N,32])."""
        img = _u8(image)
        cap = self.nfeatures + 64 * self.nlevels + 64
        kps = np.zeros(cap, KP_DTYPE)
        desc = np.zeros((cap, 32), np.uint8)
        n = lib().orc_extract(self._h, _ptr(img), img.shape[0], img.shape[1], img.strides[0], _ptr(kps), _ptr(desc), cap)
        if n < 0:
            raise RuntimeError(f"orc_extract error {n}")
        return kps[:n].copy(), desc[:n].copy()

    def level(self, l, blurred=False):
        r, c = C.c_int(), C.c_int()
        lib().orc_level_size(self._h, l, C.byref(r), C.byref(c))
        out = np.zeros((r.value, c.value), np.uint8)
        fn = lib().orc_level_blur_copy if blurred else lib().orc_level_copy
        fn(self._h, l, _ptr(out), out.strides[0])
        return out

    def candidates(self, l):
        n = lib().orc_level_candidates(self._h, l, None, 0)
        out = np.zeros(max(n, 1), KP_DTYPE)
        lib().orc_level_candidates(self._h, l, _ptr(out), n)
        return out[:n]


def stereo_match(ex_l, ex_r, kps_l, desc_l, kps_r, desc_r, mbf, mb):
    """Frame::ComputeStereoMatches -> (mvuRight, mvDepth)."""
    kps_l, kps_r = np.ascontiguousarray(kps_l), np.ascontiguousarray(kps_r)
    desc_l, desc_r = np.ascontiguousarray(desc_l), np.ascontiguousarray(desc_r)
    n = len(kps_l)
    ur = np.full(max(n, 1), -1, np.float32)
    dp = np.full(max(n, 1), -1, np.float32)
    lib().orc_stereo_match(ex_l._h, ex_r._h, _ptr(kps_l), _ptr(desc_l), n, _ptr(kps_r), _ptr(desc_r), len(kps_r),
                           mbf, mb, _ptr(ur), _ptr(dp))
    return ur[:n], dp[:n]


def resize_linear(src, dw, dh):
    src = _u8(src)
    dst = np.zeros((dh, dw), np.uint8)
    lib().orc_resize_linear(_ptr(src), src.shape[1], src.shape[0], src.strides[0], _ptr(dst), dw, dh, dst.strides[0])
    return dst


def gaussian7(src):
    src = _u8(src)
    dst = np.zeros_like(src)
    lib().orc_gaussian7(_ptr(src), src.shape[1], src.shape[0], src.strides[0], _ptr(dst), dst.strides[0])
    return dst


def fast9(roi, threshold):
    """cv::FAST(roi, kps, threshold, true) on a (possibly non-contiguous) 2-D uint8 view."""
    assert roi.dtype == np.uint8 and roi.strides[1] == 1
    cap = roi.shape[0] * roi.shape[1] // 4 + 16
    out = np.zeros(cap, KP_DTYPE)
    n = lib().orc_fast9(_ptr(roi), roi.shape[1], roi.shape[0], roi.strides[0], threshold, _ptr(out), cap)
    return out[:n]


def fast_atan2(y, x):
    return lib().orc_fast_atan2(float(y), float(x))


def sincos(rad):
    s, c = C.c_float(), C.c_float()
    lib().orc_sincos(float(rad), C.byref(s), C.byref(c))
    return s.value, c.value


def sincos_vs_libm(lo, hi, step=1):
    """Count of float bit patterns lo, lo+step, ... <= hi on which orbx_sincosf differs from the host's glibc sincosf."""
    return int(lib().orc_sincos_vs_libm(int(lo), int(hi), int(step)))


def hamming256(a, b):
    a, b = np.ascontiguousarray(a, np.uint8), np.ascontiguousarray(b, np.uint8)
    return lib().orc_hamming256(_ptr(a), _ptr(b))


def distribute(cand, min_x, max_x, min_y, max_y, n_target):
    cand = np.ascontiguousarray(cand)
    cap = len(cand) + 8
    out = np.zeros(cap, KP_DTYPE)
    n = lib().orc_distribute(_ptr(cand), len(cand), min_x, max_x, min_y, max_y, n_target, _ptr(out), cap)
    return out[:n]


# ---- rows right after the hot path: oracle/frame.cpp --------------------------------

class _OrcFrame(C.Structure):
    _fields_ = [("n", C.c_int), ("keys_un", C.c_void_p), ("desc", C.c_void_p), ("u_right", C.c_void_p),
                ("Tcw", C.c_float * 12), ("fx", C.c_float), ("fy", C.c_float), ("cx", C.c_float), ("cy", C.c_float),
                ("mbf", C.c_float), ("mb", C.c_float), ("min_x", C.c_float), ("max_x", C.c_float),
                ("min_y", C.c_float), ("max_y", C.c_float), ("scale_factors", C.c_void_p)]


class _OrcMapPoints(C.Structure):
    _fields_ = [("has_point", C.c_void_p), ("world_pos", C.c_void_p), ("point_desc", C.c_void_p),
                ("observed", C.c_void_p)]


class FrameView:
    """What SearchByProjection / GetFeaturesInArea read from a Frame (see orc_frame in orb_oracle.h)."""

    def __init__(self, keys_un, desc, u_right, Tcw, fx, fy, cx, cy, mbf, mb, bounds, scale_factors):
        self.keys_un = np.ascontiguousarray(keys_un, dtype=KP_DTYPE)
        self.desc = None if desc is None else np.ascontiguousarray(desc, dtype=np.uint8).reshape(-1, 32)
        self.u_right = None if u_right is None else np.ascontiguousarray(u_right, dtype=np.float32)
        self.Tcw = np.ascontiguousarray(np.asarray(Tcw, dtype=np.float32)[:3, :4])
        self.scale_factors = np.ascontiguousarray(scale_factors, dtype=np.float32)
        s = _OrcFrame()
        s.n = len(self.keys_un)
        s.keys_un = self.keys_un.ctypes.data
        s.desc = None if self.desc is None else self.desc.ctypes.data
        s.u_right = None if self.u_right is None else self.u_right.ctypes.data
        s.Tcw[:] = [float(v) for v in self.Tcw.ravel()]
        s.fx, s.fy, s.cx, s.cy, s.mbf, s.mb = fx, fy, cx, cy, mbf, mb
        s.min_x, s.max_x, s.min_y, s.max_y = bounds
        s.scale_factors = self.scale_factors.ctypes.data
        self.c = s


def _frame_sigs():
    L = lib()
    if getattr(L, "_frame_sigs_done", False):
        return L
    vp = C.c_void_p
    L.orc_stereo_from_rgbd.argtypes = [vp, vp, C.c_int, vp, C.c_int, C.c_int, C.c_int, C.c_float, vp, vp]
    L.orc_assign_features_to_grid.argtypes = [C.POINTER(_OrcFrame), vp, vp]
    L.orc_features_in_area.argtypes = [C.POINTER(_OrcFrame), C.c_float, C.c_float, C.c_float, C.c_int, C.c_int, vp, C.c_int]
    L.orc_project.argtypes = [vp, vp, vp]
    L.orc_tlc_z.argtypes = [vp, vp]
    L.orc_tlc_z.restype = C.c_float
    L.orc_search_by_projection.argtypes = [C.POINTER(_OrcFrame), C.POINTER(_OrcFrame), C.POINTER(_OrcMapPoints), C.c_float,
                                           C.c_int, C.c_int, vp]
    L._frame_sigs_done = True
    return L


def stereo_from_rgbd(keys, depth, mbf, keys_un_x=None):
    """Frame::ComputeStereoFromRGBD -> (mvuRight, mvDepth)."""
    keys = np.ascontiguousarray(keys, dtype=KP_DTYPE)
    depth = np.ascontiguousarray(depth, dtype=np.float32)
    n = len(keys)
    ux = None if keys_un_x is None else np.ascontiguousarray(keys_un_x, dtype=np.float32)
    ur, dp = np.zeros(max(n, 1), np.float32), np.zeros(max(n, 1), np.float32)
    _frame_sigs().orc_stereo_from_rgbd(_ptr(keys), None if ux is None else _ptr(ux), n, _ptr(depth), depth.shape[0],
                                       depth.shape[1], depth.strides[0] // 4, mbf, _ptr(ur), _ptr(dp))
    return ur[:n], dp[:n]


def assign_features_to_grid(frame):
    """Frame::AssignFeaturesToGrid -> (cell_start[64*48+1], items); cell (ix, iy) at ix*48+iy."""
    start = np.zeros(64 * 48 + 1, np.int32)
    items = np.zeros(max(frame.c.n, 1), np.int32)
    k = _frame_sigs().orc_assign_features_to_grid(C.byref(frame.c), _ptr(start), _ptr(items))
    return start, items[:k]


def features_in_area(frame, x, y, r, min_level=-1, max_level=-1):
    out = np.zeros(max(frame.c.n, 1), np.int32)
    k = _frame_sigs().orc_features_in_area(C.byref(frame.c), x, y, r, min_level, max_level, _ptr(out), len(out))
    assert k >= 0
    return out[:k]


def project(Tcw, xw):
    T = np.ascontiguousarray(np.asarray(Tcw, dtype=np.float32)[:3, :4])
    xw = np.ascontiguousarray(xw, dtype=np.float32)
    out = np.zeros(3, np.float32)
    _frame_sigs().orc_project(_ptr(T), _ptr(xw), _ptr(out))
    return out


def tlc_z(Tcw, Tlw):
    a = np.ascontiguousarray(np.asarray(Tcw, dtype=np.float32)[:3, :4])
    b = np.ascontiguousarray(np.asarray(Tlw, dtype=np.float32)[:3, :4])
    return float(_frame_sigs().orc_tlc_z(_ptr(a), _ptr(b)))


def search_by_projection(cur, last, has_point, world_pos, point_desc, observed, th, b_mono=False, check_orientation=True):
    """ORBmatcher::SearchByProjection(CurrentFrame, LastFrame, th, bMono) -> (match_of_cur[int32, cur.n], nmatches)."""
    hp = np.ascontiguousarray(has_point, dtype=np.uint8)
    wp = np.ascontiguousarray(world_pos, dtype=np.float32).reshape(-1, 3)
    pd = np.ascontiguousarray(point_desc, dtype=np.uint8).reshape(-1, 32)
    ob = np.ascontiguousarray(observed, dtype=np.uint8)
    assert len(hp) == len(wp) == len(pd) == len(ob) == last.c.n
    m = _OrcMapPoints(hp.ctypes.data, wp.ctypes.data, pd.ctypes.data, ob.ctypes.data)
    out = np.full(max(cur.c.n, 1), -1, np.int32)
    n = _frame_sigs().orc_search_by_projection(C.byref(cur.c), C.byref(last.c), C.byref(m), th, int(b_mono),
                                               int(check_orientation), _ptr(out))
    return out[:cur.c.n], n


# ---- input staging: oracle/input.cpp -------------------------------------------------------

def cvt_gray(img, rgb_order=False):
    """cvtColor(img, CV_BGR2GRAY | CV_RGB2GRAY | CV_BGRA2GRAY | CV_RGBA2GRAY) on uint8 [H,W,C]; C == 1 copies."""
    img = np.ascontiguousarray(img, dtype=np.uint8)
    if img.ndim == 2:
        img = img[:, :, None]
    h, w, c = img.shape
    out = np.zeros((h, w), np.uint8)
    L = lib()
    L.orc_cvt_gray.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_int]
    L.orc_cvt_gray(_ptr(img), h, w, img.strides[0], c, int(rgb_order), _ptr(out), out.strides[0])
    return out


def remap_linear(img, mapx, mapy):
    """cv::remap(img, mapx, mapy, INTER_LINEAR), CV_32FC1 maps, BORDER_CONSTANT 0; uint8 [H,W] or [H,W,C]."""
    img = np.ascontiguousarray(img, dtype=np.uint8)
    squeeze = img.ndim == 2
    if squeeze:
        img = img[:, :, None]
    mapx, mapy = np.ascontiguousarray(mapx, dtype=np.float32), np.ascontiguousarray(mapy, dtype=np.float32)
    assert mapx.shape == mapy.shape
    h, w, c = img.shape
    dh, dw = mapx.shape
    out = np.zeros((dh, dw, c), np.uint8)
    L = lib()
    L.orc_remap_linear.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int,
                                   C.c_void_p, C.c_int]
    L.orc_remap_linear(_ptr(img), h, w, img.strides[0], c, _ptr(mapx), _ptr(mapy), mapx.strides[0] // 4, dh, dw, _ptr(out), out.strides[0])
    return out[:, :, 0] if squeeze else out


# ---- bag of words: oracle/bow.cpp ------------------------------------------------------------

class Vocabulary:
    """The DBoW2 ORB vocabulary as loadFromTextFile builds it: entry i = node i+1 (parent, leaf flag, descriptor, weight)."""

    def __init__(self, L, parent, is_leaf, desc, weight):
        self.parent = np.ascontiguousarray(parent, dtype=np.int32)
        self.is_leaf = np.ascontiguousarray(is_leaf, dtype=np.uint8)
        self.desc = np.ascontiguousarray(desc, dtype=np.uint8).reshape(-1, 32)
        self.weight = np.ascontiguousarray(weight, dtype=np.float64)
        lb = lib()
        lb.orc_voc_create.restype = C.c_void_p
        lb.orc_voc_create.argtypes = [C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
        lb.orc_voc_destroy.argtypes = [C.c_void_p]
        lb.orc_voc_words.argtypes = [C.c_void_p]
        lb.orc_compute_bow.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_int] + [C.c_void_p] * 7
        self.L = L
        self._h = lb.orc_voc_create(L, len(self.parent), _ptr(self.parent), _ptr(self.is_leaf), _ptr(self.desc), _ptr(self.weight))
        if not self._h:
            raise ValueError("bad vocabulary")

    def __del__(self):
        h, self._h = getattr(self, "_h", None), None
        if h:
            lib().orc_voc_destroy(h)

    @property
    def n_words(self):
        return lib().orc_voc_words(self._h)

    def transform(self, desc, levelsup=4):
        """-> (word_ids, word_vals, fv_nodes, fv_start, fv_feats): BowVector and FeatureVector (CSR) of the features."""
        desc = np.ascontiguousarray(desc, dtype=np.uint8).reshape(-1, 32)
        n = len(desc)
        wi, wv = np.zeros(max(n, 1), np.int32), np.zeros(max(n, 1), np.float64)
        fn, fs, ff = np.zeros(max(n, 1), np.int32), np.zeros(n + 1, np.int32), np.zeros(max(n, 1), np.int32)
        nw, nf = C.c_int(), C.c_int()
        lib().orc_compute_bow(self._h, _ptr(desc), n, levelsup, _ptr(wi), _ptr(wv), C.addressof(nw), _ptr(fn), _ptr(fs), _ptr(ff),
                              C.addressof(nf))
        return wi[: nw.value], wv[: nw.value], fn[: nf.value], fs[: nf.value + 1], ff[: fs[nf.value]]


class _OrcTrackPoints(C.Structure):
    _fields_ = [("n", C.c_int), ("valid", C.c_void_p), ("proj", C.c_void_p), ("level", C.c_void_p), ("view_cos", C.c_void_p),
                ("point_desc", C.c_void_p), ("observed", C.c_void_p)]


def search_local_points(frame, occupied, valid, proj, level, view_cos, point_desc, observed, th, nnratio=0.8):
    """ORBmatcher::SearchByProjection(F, vpMapPoints, th) (src/ORBmatcher.cc:45-127) -> (match_of_key, nmatches)."""
    va = np.ascontiguousarray(valid, dtype=np.uint8)
    pr = np.ascontiguousarray(proj, dtype=np.float32).reshape(-1, 3)
    lv = np.ascontiguousarray(level, dtype=np.int32)
    vc = np.ascontiguousarray(view_cos, dtype=np.float32)
    pd = np.ascontiguousarray(point_desc, dtype=np.uint8).reshape(-1, 32)
    ob = np.ascontiguousarray(observed, dtype=np.uint8)
    oc = None if occupied is None else np.ascontiguousarray(occupied, dtype=np.uint8)
    t = _OrcTrackPoints(len(va), va.ctypes.data, pr.ctypes.data, lv.ctypes.data, vc.ctypes.data, pd.ctypes.data, ob.ctypes.data)
    L = _frame_sigs()
    L.orc_search_local_points.argtypes = [C.POINTER(_OrcFrame), C.c_void_p, C.POINTER(_OrcTrackPoints), C.c_float, C.c_float, C.c_void_p]
    out = np.full(max(frame.c.n, 1), -1, np.int32)
    n = L.orc_search_local_points(C.byref(frame.c), None if oc is None else _ptr(oc), C.byref(t), th, nnratio, _ptr(out))
    return out[: frame.c.n], n
