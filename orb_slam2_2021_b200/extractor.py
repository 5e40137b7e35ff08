"""Host-side mirror of the reference interface for the hot path, over the C-ABI of liborbx.so.

  ORBextractor            <-> ORB_SLAM2::ORBextractor            (include/ORBextractor.h:56-100)
  compute_stereo_matches  <-> Frame::ComputeStereoMatches        (src/Frame.cc:522-700)
  StereoFrameBatch        <-> Frame::Frame(stereo) over N frames (src/Frame.cc:80-152), the throughput path

Same names, argument meaning and error behaviour as the reference where it has any: an empty image
returns no keypoints without raising (src/ORBextractor.cc:1044-1045); everything else that the
reference leaves undefined raises OrbxError. All compute runs in the CUDA library; nothing here
touches oracle/.
"""
import ctypes as C

import numpy as np

from ._lib import KP_DTYPE, OrbxError, check, lib, ptr


def _as_u8_2d(image):
    img = np.asarray(image)
    if img.dtype != np.uint8 or img.ndim != 2:
        raise TypeError("image must be a 2-D uint8 array (CV_8UC1)")  # the reference asserts type()==CV_8UC1, :1048
    if img.size and img.strides[1] != 1:
        img = np.ascontiguousarray(img)
    return img


class ORBextractor:
    """ORBextractor(nfeatures, scaleFactor, nlevels, iniThFAST, minThFAST) on CUDA device `device`."""

    def __init__(self, nfeatures, scaleFactor, nlevels, iniThFAST, minThFAST, device=0):
        self._h = C.c_void_p()
        check(lib().orbx_create(nfeatures, scaleFactor, nlevels, iniThFAST, minThFAST, device, C.byref(self._h)))
        self.nfeatures, self.nlevels, self.scaleFactor = nfeatures, nlevels, float(np.float32(scaleFactor))
        self._shape = None

    def __del__(self):
        h, self._h = getattr(self, "_h", None), None
        if h:
            lib().orbx_destroy(h)

    # --- getters (include/ORBextractor.h:79-98) ---
    def GetLevels(self):
        return lib().orbx_get_levels(self._h)

    def GetScaleFactor(self):
        return self.scaleFactor

    def _tables(self):
        n = self.nlevels
        f = [np.zeros(n, np.float32) for _ in range(4)]
        q = np.zeros(n, np.int32)
        check(lib().orbx_get_tables(self._h, ptr(f[0]), ptr(f[1]), ptr(f[2]), ptr(f[3]), ptr(q)))
        return f, q

    def GetScaleFactors(self):
        return self._tables()[0][0]

    def GetInverseScaleFactors(self):
        return self._tables()[0][1]

    def GetScaleSigmaSquares(self):
        return self._tables()[0][2]

    def GetInverseScaleSigmaSquares(self):
        return self._tables()[0][3]

    def GetFeaturesPerLevel(self):  # mnFeaturesPerLevel (protected in the reference; exposed for tests)
        return self._tables()[1]

    # --- operator() ---
    def __call__(self, image, mask=None):
        """(keypoints[KP_DTYPE], descriptors[N,32] uint8). `mask` is ignored, as in the reference."""
        img = _as_u8_2d(image)
        if img.size == 0:
            return np.zeros(0, KP_DTYPE), np.zeros((0, 32), np.uint8)
        rows, cols = img.shape
        cap = lib().orbx_max_keypoints(self._h, rows, cols)
        if cap < 0:
            check(cap)
        kps = np.zeros(cap, KP_DTYPE)
        desc = np.zeros((cap, 32), np.uint8)
        n = C.c_int(0)
        check(lib().orbx_extract(self._h, ptr(img), rows, cols, img.strides[0], ptr(kps), ptr(desc), cap, C.byref(n)))
        self._shape = (rows, cols)
        return kps[: n.value].copy(), desc[: n.value].copy()

    @property
    def mvImagePyramid(self):
        """The levels of the last extract, copied from the device (list of 2-D uint8 arrays)."""
        out = []
        for l in range(self.nlevels):
            r, c = C.c_int(), C.c_int()
            check(lib().orbx_pyramid_level_size(self._h, l, C.byref(r), C.byref(c)))
            a = np.zeros((r.value, c.value), np.uint8)
            check(lib().orbx_pyramid_level(self._h, l, ptr(a), a.strides[0]))
            out.append(a)
        return out


def compute_stereo_matches(extractor_left, extractor_right, keys_left, desc_left, keys_right, desc_right, mbf, mb):
    """Frame::ComputeStereoMatches -> (mvuRight, mvDepth), float32[N], -1 where unmatched."""
    kl = np.ascontiguousarray(keys_left, dtype=KP_DTYPE)
    kr = np.ascontiguousarray(keys_right, dtype=KP_DTYPE)
    dl = np.ascontiguousarray(desc_left, dtype=np.uint8).reshape(-1, 32)
    dr = np.ascontiguousarray(desc_right, dtype=np.uint8).reshape(-1, 32)
    if len(kl) != len(dl) or len(kr) != len(dr):
        raise ValueError("keypoints and descriptors differ in length")
    n = len(kl)
    ur = np.full(n, -1, np.float32)
    dp = np.full(n, -1, np.float32)
    check(lib().orbx_stereo_match(extractor_left._h, extractor_right._h, ptr(kl), ptr(dl), n, ptr(kr), ptr(dr), len(kr),
                                  float(mbf), float(mb), ptr(ur), ptr(dp)))
    return ur, dp


def compute_stereo_matches_resident(extractor_left, extractor_right, n_left, n_right, mbf, mb):
    """Frame::ComputeStereoMatches on the keypoints and descriptors the two extractors' last calls left on the device
    (orbx_stereo_match_resident: nothing is uploaded) -> (mvuRight, mvDepth)."""
    ur = np.full(n_left, -1, np.float32)
    dp = np.full(n_left, -1, np.float32)
    check(lib().orbx_stereo_match_resident(extractor_left._h, extractor_right._h, int(n_left), int(n_right), float(mbf), float(mb),
                                           ptr(ur), ptr(dp)))
    return ur, dp


def stereo_frame(extractor_left, extractor_right, image_left, image_right, mbf, mb):
    """One stereo frame in one call (orbx_stereo_frame; src/Frame.cc:113-125: ExtractORB left + right, ComputeStereoMatches)
    -> (keys_left, desc_left, keys_right, desc_right, mvuRight, mvDepth)."""
    il, ir = _as_u8_2d(image_left), _as_u8_2d(image_right)
    if il.shape != ir.shape:
        raise ValueError("left and right images differ in size")
    if il.size == 0:
        z, d = np.zeros(0, KP_DTYPE), np.zeros((0, 32), np.uint8)
        return z, d, z.copy(), d.copy(), np.zeros(0, np.float32), np.zeros(0, np.float32)
    rows, cols = il.shape
    cap = lib().orbx_max_keypoints(extractor_left._h, rows, cols)
    if cap < 0:
        check(cap)
    kl, kr = np.zeros(cap, KP_DTYPE), np.zeros(cap, KP_DTYPE)
    dl, dr = np.zeros((cap, 32), np.uint8), np.zeros((cap, 32), np.uint8)
    ur, dp = np.zeros(cap, np.float32), np.zeros(cap, np.float32)
    nl, nr = C.c_int(0), C.c_int(0)
    check(lib().orbx_stereo_frame(extractor_left._h, extractor_right._h, ptr(il), ptr(ir), rows, cols, il.strides[0], ir.strides[0],
                                  float(mbf), float(mb), ptr(kl), ptr(dl), C.byref(nl), ptr(kr), ptr(dr), C.byref(nr), cap, ptr(ur),
                                  ptr(dp)))
    a, b = nl.value, nr.value
    return kl[:a].copy(), dl[:a].copy(), kr[:b].copy(), dr[:b].copy(), ur[:a].copy(), dp[:a].copy()


def compute_stereo_matches_static(extractor_left, extractor_right, keys_left, desc_left, keys_right, desc_right, mbf, mb):
    """compute_stereo_matches through orbx_stereo_match_static: same results, device scratch in a static array inside
    the library (no allocation; at most 8192 keypoints per side and 4096 rows; calls overlapping on one GPU race)."""
    kl = np.ascontiguousarray(keys_left, dtype=KP_DTYPE)
    kr = np.ascontiguousarray(keys_right, dtype=KP_DTYPE)
    dl = np.ascontiguousarray(desc_left, dtype=np.uint8).reshape(-1, 32)
    dr = np.ascontiguousarray(desc_right, dtype=np.uint8).reshape(-1, 32)
    if len(kl) != len(dl) or len(kr) != len(dr):
        raise ValueError("keypoints and descriptors differ in length")
    n = len(kl)
    ur = np.full(n, -1, np.float32)
    dp = np.full(n, -1, np.float32)
    check(lib().orbx_stereo_match_static(extractor_left._h, extractor_right._h, ptr(kl), ptr(dl), n, ptr(kr), ptr(dr), len(kr),
                                         float(mbf), float(mb), ptr(ur), ptr(dp)))
    return ur, dp


def set_tuning(key, value):
    """orbx_set_tuning: debug/experiment knobs of the library (include/orbx.h lists the keys); 0 = built-in default."""
    check(lib().orbx_set_tuning(str(key).encode(), int(value)))


def host_numa_node():
    """(node, bound) of the last orbx_host_alloc: the NUMA node of the CUDA device it aimed at (-1 unknown, -2 no
    allocation yet) and how it was placed (1 memory policy, 2 allocated from one of the node's cores, 0 plain)."""
    bound = C.c_int(0)
    node = lib().orbx_host_numa_node(C.byref(bound))
    return int(node), int(bound.value)


class PinnedArray:
    """numpy view over page-locked host memory from orbx_host_alloc (for end-to-end transfers); the pages sit on the
    NUMA node of the current CUDA device where the host permits (see include/orbx.h)."""

    def __init__(self, shape, dtype):
        self.nbytes = int(np.prod(shape)) * np.dtype(dtype).itemsize
        self._p = lib().orbx_host_alloc(max(self.nbytes, 1))
        if not self._p:
            raise OrbxError(-4, "orbx_host_alloc failed")
        buf = (C.c_uint8 * max(self.nbytes, 1)).from_address(self._p)
        self.array = np.frombuffer(buf, dtype=dtype, count=int(np.prod(shape))).reshape(shape)

    def __del__(self):
        p, self._p = getattr(self, "_p", None), None
        if p:
            self.array = None
            lib().orbx_host_free(p)


class StereoFrameBatch:
    """N frames per call: extract(left) + extract(right) + stereo match (mono: extract only)."""

    def __init__(self, nfeatures, scaleFactor, nlevels, iniThFAST, minThFAST, rows, cols, max_frames, stereo=True,
                 device=0):
        self._h = C.c_void_p()
        check(lib().orbx_batch_create(nfeatures, scaleFactor, nlevels, iniThFAST, minThFAST, rows, cols, max_frames,
                                      int(stereo), device, C.byref(self._h)))
        self.rows, self.cols, self.max_frames, self.stereo, self.nlevels = rows, cols, max_frames, stereo, nlevels
        self.pitch = lib().orbx_batch_max_keypoints(self._h)

    def __del__(self):
        h, self._h = getattr(self, "_h", None), None
        if h:
            lib().orbx_batch_destroy(h)

    @property
    def stream(self):
        return lib().orbx_batch_stream(self._h)

    @property
    def launches_per_run(self):
        return lib().orbx_batch_launches_per_run(self._h)

    @property
    def device_bytes(self):
        return lib().orbx_batch_device_bytes(self._h)

    def alloc_outputs(self, n_frames, pinned=False):
        P = self.pitch
        shapes = dict(n_left=((n_frames,), np.int32), kps_left=((n_frames, P), KP_DTYPE),
                      desc_left=((n_frames, P, 32), np.uint8))
        if self.stereo:
            shapes.update(n_right=((n_frames,), np.int32), kps_right=((n_frames, P), KP_DTYPE),
                          desc_right=((n_frames, P, 32), np.uint8), u_right=((n_frames, P), np.float32),
                          depth=((n_frames, P), np.float32))
        if pinned:
            pins = {k: PinnedArray(s, d) for k, (s, d) in shapes.items()}
            self._pins = getattr(self, "_pins", []) + [pins]  # every set stays alive as long as the batch (streaming uses two)
            return {k: v.array for k, v in pins.items()}
        return {k: np.zeros(s, d) for k, (s, d) in shapes.items()}

    def _check_imgs(self, a, n):
        a = np.asarray(a)
        if a.dtype != np.uint8 or a.shape != (n, self.rows, self.cols) or not a.flags.c_contiguous:
            raise TypeError(f"images must be C-contiguous uint8 [{n},{self.rows},{self.cols}]")
        return a

    def upload(self, left, right=None):
        n = len(left)
        left = self._check_imgs(left, n)
        right = self._check_imgs(right, n) if self.stereo else None
        check(lib().orbx_batch_upload(self._h, ptr(left), ptr(right), n))
        return n

    def set_rectification(self, side, map1, map2, crop_top=0):
        """CV_32FC1 maps of initUndistortRectifyMap for one side (0 left, 1 right); None clears them."""
        if map1 is None:
            check(lib().orbx_batch_set_rectification(self._h, side, None, None, 0, 0, 0, crop_top))
            return
        map1, map2 = np.ascontiguousarray(map1, dtype=np.float32), np.ascontiguousarray(map2, dtype=np.float32)
        if map1.shape != map2.shape or map1.ndim != 2:
            raise ValueError("maps must be two float32 arrays of one 2-D shape")
        check(lib().orbx_batch_set_rectification(self._h, side, ptr(map1), ptr(map2), map1.shape[0], map1.shape[1], map1.strides[0], crop_top))

    def upload_raw(self, left, right=None, rgb=False):
        """Raw camera frames uint8 [n,H,W] or [n,H,W,3|4]: rectified (if maps are set), cropped and converted to gray on
        the device, straight into level 0 (cv::remap + crop + cvtColor of the reference's input path)."""
        left = np.asarray(left)
        if left.dtype != np.uint8 or left.ndim not in (3, 4) or not left.flags.c_contiguous:
            raise TypeError("raw frames must be C-contiguous uint8 [n,H,W] or [n,H,W,C]")
        if self.stereo:
            right = np.asarray(right)
            if right.dtype != np.uint8 or right.shape != left.shape or not right.flags.c_contiguous:
                raise TypeError("right frames must match the left frames")
        n, h, w = left.shape[:3]
        ch = 1 if left.ndim == 3 else left.shape[3]
        check(lib().orbx_batch_upload_raw(self._h, ptr(left), ptr(right) if self.stereo else None, n, h, w, ch, int(rgb)))
        return n

    def upload_device(self, left_ptr, right_ptr, n_frames):
        check(lib().orbx_batch_upload_device(self._h, left_ptr, right_ptr, n_frames))

    def run(self, n_frames, mbf=0.0, mb=1.0, sync=True):
        check(lib().orbx_batch_run(self._h, n_frames, float(mbf), float(mb), int(sync)))

    def download(self, n_frames, out=None):
        out = out or self.alloc_outputs(n_frames)
        g = out.get
        check(lib().orbx_batch_download(self._h, n_frames, ptr(g("n_left")), ptr(g("kps_left")), ptr(g("desc_left")),
                                        ptr(g("n_right")), ptr(g("kps_right")), ptr(g("desc_right")),
                                        ptr(g("u_right")), ptr(g("depth"))))
        return out

    def process(self, left, right=None, mbf=0.0, mb=1.0, out=None):
        """End-to-end: host images in, host results out (H2D + kernels + D2H)."""
        n = len(left)
        left = self._check_imgs(left, n)
        right = self._check_imgs(right, n) if self.stereo else None
        out = out or self.alloc_outputs(n)
        g = out.get
        check(lib().orbx_batch_process(self._h, ptr(left), ptr(right), n, float(mbf), float(mb), ptr(g("n_left")),
                                       ptr(g("kps_left")), ptr(g("desc_left")), ptr(g("n_right")), ptr(g("kps_right")),
                                       ptr(g("desc_right")), ptr(g("u_right")), ptr(g("depth"))))
        return out

    def process_async(self, left, right, mbf, mb, out):
        """Streaming form of process(): enqueue and return; call wait() before reading `out` or reusing the buffers.

        `left`, `right` and every array of `out` should be pinned (PinnedArray / alloc_outputs(pinned=True)); the
        caller keeps them alive and untouched until wait() returns."""
        n = len(left)
        left = self._check_imgs(left, n)
        right = self._check_imgs(right, n) if self.stereo else None
        g = out.get
        check(lib().orbx_batch_process_async(self._h, ptr(left), ptr(right), n, float(mbf), float(mb), ptr(g("n_left")),
                                             ptr(g("kps_left")), ptr(g("desc_left")), ptr(g("n_right")), ptr(g("kps_right")),
                                             ptr(g("desc_right")), ptr(g("u_right")), ptr(g("depth"))))
        self._inflight = (left, right, out)  # keep the buffers alive
        return out

    def wait(self):
        check(lib().orbx_batch_wait(self._h))
        self._inflight = None

    def stats(self, n_frames):
        a, b, c = C.c_longlong(), C.c_longlong(), C.c_longlong()
        check(lib().orbx_batch_stats(self._h, n_frames, C.byref(a), C.byref(b), C.byref(c)))
        return a.value, b.value, c.value

    def set_profiling(self, on=True):
        check(lib().orbx_batch_set_profiling(self._h, int(on)))

    def stage_times(self):
        """({stage: total ms}, runs) accumulated since the last call (CUDA events on the engine stream)."""
        n = lib().orbx_batch_stage_count()
        ms = (C.c_double * n)()
        runs = C.c_int()
        check(lib().orbx_batch_stage_times(self._h, ms, C.byref(runs)))
        return {lib().orbx_batch_stage_name(i).decode(): ms[i] for i in range(n)}, runs.value

    # --- test taps ---
    def level(self, frame, side, level, blurred=False):
        r, c = C.c_int(), C.c_int()
        check(lib().orbx_batch_level_size(self._h, level, C.byref(r), C.byref(c)))
        a = np.zeros((r.value, c.value), np.uint8)
        check(lib().orbx_batch_level(self._h, frame, side, level, int(blurred), ptr(a), a.strides[0]))
        return a

    def candidates(self, frame, side, level):
        n = C.c_int()
        check(lib().orbx_batch_candidates(self._h, frame, side, level, None, 0, C.byref(n)))
        out = np.zeros(max(n.value, 1), KP_DTYPE)
        check(lib().orbx_batch_candidates(self._h, frame, side, level, ptr(out), n.value, C.byref(n)))
        return out[: n.value]
