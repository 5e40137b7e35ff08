"""Host-side mirror of the reference interface for the rows right after the hot path (include/orbx_frame.h).

  FrameView                       <-> what the functions below read from ORB_SLAM2::Frame (include/Frame.h)
  Matcher.ComputeStereoFromRGBD   <-> Frame::ComputeStereoFromRGBD   (src/Frame.cc:702-723)
  Matcher.AssignFeaturesToGrid    <-> Frame::AssignFeaturesToGrid    (src/Frame.cc:279-294)
  Matcher.GetFeaturesInArea       <-> Frame::GetFeaturesInArea       (src/Frame.cc:376-433)
  Matcher.SearchByProjection      <-> ORBmatcher::SearchByProjection(CurrentFrame, LastFrame, th, bMono)
                                                                      (src/ORBmatcher.cc:1348-1491)
All compute runs in liborbx.so on the GPU; nothing here touches oracle/.
"""
import ctypes as C

import numpy as np

from ._lib import KP_DTYPE, check, lib, ptr

GRID_COLS, GRID_ROWS = 64, 48


class _FrameStruct(C.Structure):  # orbx_frame_view
    _fields_ = [("n", C.c_int), ("keys_un", C.c_void_p), ("desc", C.c_void_p), ("u_right", C.c_void_p),
                ("Tcw", C.c_float * 12), ("fx", C.c_float), ("fy", C.c_float), ("cx", C.c_float), ("cy", C.c_float),
                ("mbf", C.c_float), ("mb", C.c_float), ("min_x", C.c_float), ("max_x", C.c_float),
                ("min_y", C.c_float), ("max_y", C.c_float), ("scale_factors", C.c_void_p), ("nlevels", C.c_int)]


class _MapPointStruct(C.Structure):  # orbx_mappoint_view
    _fields_ = [("has_point", C.c_void_p), ("world_pos", C.c_void_p), ("point_desc", C.c_void_p),
                ("observed", C.c_void_p)]


class _FrameSourceStruct(C.Structure):  # orbx_frame_source
    _fields_ = [("batch", C.c_void_p), ("frame", C.c_int), ("side", C.c_int)]


class FrameView:
    """mvKeysUn, mDescriptors, mvuRight, mTcw, intrinsics, image bounds and mvScaleFactors of one Frame."""

    def __init__(self, keys_un, desc, u_right, Tcw, fx, fy, cx, cy, mbf, mb, bounds, scale_factors):
        self.keys_un = np.ascontiguousarray(keys_un, dtype=KP_DTYPE)
        self.desc = None if desc is None else np.ascontiguousarray(desc, dtype=np.uint8).reshape(-1, 32)
        self.u_right = None if u_right is None else np.ascontiguousarray(u_right, dtype=np.float32)
        self.Tcw = np.ascontiguousarray(np.asarray(Tcw, dtype=np.float32)[:3, :4])
        self.scale_factors = np.ascontiguousarray(scale_factors, dtype=np.float32)
        if self.desc is not None and len(self.desc) != len(self.keys_un):
            raise ValueError("keypoints and descriptors differ in length")
        if self.u_right is not None and len(self.u_right) != len(self.keys_un):
            raise ValueError("keypoints and mvuRight differ in length")
        s = _FrameStruct()
        s.n = len(self.keys_un)
        s.keys_un = self.keys_un.ctypes.data
        s.desc = None if self.desc is None else self.desc.ctypes.data
        s.u_right = None if self.u_right is None else self.u_right.ctypes.data
        s.Tcw[:] = [float(v) for v in self.Tcw.ravel()]
        s.fx, s.fy, s.cx, s.cy, s.mbf, s.mb = fx, fy, cx, cy, mbf, mb
        s.min_x, s.max_x, s.min_y, s.max_y = bounds
        s.scale_factors = self.scale_factors.ctypes.data
        s.nlevels = len(self.scale_factors)
        self.c = s

    @property
    def N(self):
        return self.c.n


class MapPointView:
    """Per keypoint of the last frame: has a usable map point, its world position, descriptor and Observations() > 0."""

    def __init__(self, has_point, world_pos, point_desc, observed):
        self.has_point = np.ascontiguousarray(has_point, dtype=np.uint8)
        self.world_pos = np.ascontiguousarray(world_pos, dtype=np.float32).reshape(-1, 3)
        self.point_desc = np.ascontiguousarray(point_desc, dtype=np.uint8).reshape(-1, 32)
        self.observed = np.ascontiguousarray(observed, dtype=np.uint8)
        if not (len(self.has_point) == len(self.world_pos) == len(self.point_desc) == len(self.observed)):
            raise ValueError("map point arrays differ in length")
        self.c = _MapPointStruct(self.has_point.ctypes.data, self.world_pos.ctypes.data, self.point_desc.ctypes.data,
                                 self.observed.ctypes.data)


class _TrackPointStruct(C.Structure):  # orbx_trackpoint_view
    _fields_ = [("n", C.c_int), ("valid", C.c_void_p), ("proj", C.c_void_p), ("level", C.c_void_p), ("view_cos", C.c_void_p),
                ("point_desc", C.c_void_p), ("observed", C.c_void_p)]


class TrackPointView:
    """The local map as Tracking::SearchLocalPoints hands it to the matcher: per map point mbTrackInView && !isBad(),
    (mTrackProjX, mTrackProjY, mTrackProjXR), mnTrackScaleLevel, mTrackViewCos, descriptor, Observations() > 0."""

    def __init__(self, valid, proj, level, view_cos, point_desc, observed):
        self.valid = np.ascontiguousarray(valid, dtype=np.uint8)
        self.proj = np.ascontiguousarray(proj, dtype=np.float32).reshape(-1, 3)
        self.level = np.ascontiguousarray(level, dtype=np.int32)
        self.view_cos = np.ascontiguousarray(view_cos, dtype=np.float32)
        self.point_desc = np.ascontiguousarray(point_desc, dtype=np.uint8).reshape(-1, 32)
        self.observed = np.ascontiguousarray(observed, dtype=np.uint8)
        n = len(self.valid)
        if not all(len(a) == n for a in (self.proj, self.level, self.view_cos, self.point_desc, self.observed)):
            raise ValueError("map point arrays differ in length")
        self.c = _TrackPointStruct(n, self.valid.ctypes.data, self.proj.ctypes.data, self.level.ctypes.data, self.view_cos.ctypes.data,
                                   self.point_desc.ctypes.data, self.observed.ctypes.data)


class Matcher:
    """Owns the device scratch and the CUDA stream of the frame-level functions (orbx_matcher)."""

    def __init__(self, device=0, checkOrientation=True):
        self._h = C.c_void_p()
        self.mbCheckOrientation = bool(checkOrientation)  # ORBmatcher::ORBmatcher(nnratio, checkOri), src/ORBmatcher.cc:41
        check(lib().orbx_matcher_create(device, C.byref(self._h)))

    def __del__(self):
        h, self._h = getattr(self, "_h", None), None
        if h:
            lib().orbx_matcher_destroy(h)

    def ComputeStereoFromRGBD(self, keys, imDepth, mbf, keys_un_x=None):
        """-> (mvuRight, mvDepth) float32[N]; imDepth is the float32 depth map after mDepthMapFactor scaling."""
        keys = np.ascontiguousarray(keys, dtype=KP_DTYPE)
        d = np.asarray(imDepth)
        if d.dtype != np.float32 or d.ndim != 2 or d.strides[1] != 4:
            raise TypeError("imDepth must be a 2-D float32 array (CV_32F)")
        ux = None if keys_un_x is None else np.ascontiguousarray(keys_un_x, dtype=np.float32)
        n = len(keys)
        ur, dp = np.full(n, -1, np.float32), np.full(n, -1, np.float32)
        check(lib().orbx_stereo_from_rgbd(self._h, ptr(keys), ptr(ux), n, ptr(d), d.shape[0], d.shape[1], d.strides[0], float(mbf),
                                          ptr(ur), ptr(dp)))
        return ur, dp

    def RectifyGray(self, image, map1=None, map2=None, rgb=False, crop_top=0, crop_rows=None):
        """gray(crop(remap(image, map1, map2, INTER_LINEAR))) — cv::remap (Examples/Stereo/arducam_images.cpp:127-133)
        followed by cvtColor(..., CV_RGB2GRAY if rgb else CV_BGR2GRAY) (src/Tracking.cc:183-208). image: uint8 [H,W] or
        [H,W,3|4]; maps: float32 [h,w] or None."""
        img = np.asarray(image)
        if img.dtype != np.uint8 or img.ndim not in (2, 3):
            raise TypeError("image must be uint8 [H,W] or [H,W,C]")
        img = np.ascontiguousarray(img)
        ch = 1 if img.ndim == 2 else img.shape[2]
        if (map1 is None) != (map2 is None):
            raise ValueError("need both maps or none")
        if map1 is not None:
            map1, map2 = np.ascontiguousarray(map1, dtype=np.float32), np.ascontiguousarray(map2, dtype=np.float32)
            if map1.shape != map2.shape or map1.ndim != 2:
                raise ValueError("maps must be two float32 arrays of one 2-D shape")
        mrows, mcols = map1.shape if map1 is not None else img.shape[:2]
        rows = mrows - crop_top if crop_rows is None else crop_rows
        out = np.zeros((max(rows, 0), mcols), np.uint8)
        check(lib().orbx_rectify_gray(self._h, ptr(img), img.shape[0], img.shape[1], img.strides[0], ch, int(rgb), ptr(map1), ptr(map2),
                                      mrows, mcols, mcols * 4, crop_top, rows, ptr(out), out.strides[0] if out.size else mcols))
        return out

    def AssignFeaturesToGrid(self, frame):
        """-> (cell_start[64*48+1], cell_items): mGrid[ix][iy] = cell_items[cell_start[ix*48+iy] : cell_start[ix*48+iy+1]]."""
        start = np.zeros(GRID_COLS * GRID_ROWS + 1, np.int32)
        items = np.zeros(max(frame.N, 1), np.int32)
        n = C.c_int()
        check(lib().orbx_assign_features_to_grid(self._h, C.addressof(frame.c), ptr(start), ptr(items), C.byref(n)))
        return start, items[: n.value]

    def GetFeaturesInArea(self, frame, x, y, r, minLevel=-1, maxLevel=-1):
        out = np.zeros(max(frame.N, 1), np.int32)
        n = C.c_int()
        check(lib().orbx_get_features_in_area(self._h, C.addressof(frame.c), float(x), float(y), float(r), int(minLevel), int(maxLevel),
                                              ptr(out), frame.N, C.byref(n)))
        return out[: n.value]

    def SearchLocalPoints(self, F, points, th, occupied=None, nnratio=0.8, frame_src=None):
        """ORBmatcher(nnratio).SearchByProjection(F, vpMapPoints, th) (src/ORBmatcher.cc:45-127). One (frame, local map) pair or
        lists of pairs -> (match_of_key int32[N], nmatches) per pair. occupied: per keypoint, F.mvpMapPoints[i] holds a point
        with Observations() > 0 on entry."""
        single = isinstance(F, FrameView)
        fr = [F] if single else list(F)
        pts = [points] if single else list(points)
        occ = [occupied] if single else (list(occupied) if occupied is not None else [None] * len(fr))
        if not (len(fr) == len(pts) == len(occ)) or not fr:
            raise ValueError("need equally many frames, map point views and occupancy arrays")
        occ = [None if o is None else np.ascontiguousarray(o, dtype=np.uint8) for o in occ]
        for f, o in zip(fr, occ):
            if o is not None and len(o) != f.N:
                raise ValueError("occupied and frame differ in length")
        n = len(fr)
        fa, pa = (_FrameStruct * n)(*[f.c for f in fr]), (_TrackPointStruct * n)(*[p.c for p in pts])
        oa = (C.c_void_p * n)(*[None if o is None else o.ctypes.data for o in occ])
        outs = [np.full(max(f.N, 1), -1, np.int32) for f in fr]
        op = (C.c_void_p * n)(*[o.ctypes.data for o in outs])
        nm = np.zeros(n, np.int32)
        if frame_src is None:
            check(lib().orbx_search_local_points(self._h, fa, oa, pa, n, float(th), float(nnratio), op, ptr(nm)))
        else:  # per pair None or (StereoFrameBatch, frame, side): the frame's arrays are read on the device
            v = [frame_src] if single else list(frame_src)
            sa = (_FrameSourceStruct * n)(*[_FrameSourceStruct(None, 0, 0) if x is None else _FrameSourceStruct(x[0]._h, x[1], x[2]) for x in v])
            check(lib().orbx_search_local_points_resident(self._h, fa, sa, oa, pa, n, float(th), float(nnratio), op, ptr(nm)))
        res = [(o[: f.N], int(k)) for o, f, k in zip(outs, fr, nm)]
        return res[0] if single else res

    def SearchByProjection(self, CurrentFrame, LastFrame, lastPoints, th, bMono=False, cur_src=None, last_src=None):
        """One (current, last) pair, or lists of pairs -> (match_of_cur int32[N_cur], nmatches) per pair.
        cur_src / last_src: per pair None or (StereoFrameBatch, frame, side): the frame's keypoints, descriptors and uRight
        are read where the batch engine left them on the device (orbx_search_by_projection_resident) instead of from the
        view's host arrays."""
        single = isinstance(CurrentFrame, FrameView)
        cur = [CurrentFrame] if single else list(CurrentFrame)
        last = [LastFrame] if single else list(LastFrame)
        pts = [lastPoints] if single else list(lastPoints)
        if not (len(cur) == len(last) == len(pts)) or not cur:
            raise ValueError("need equally many current frames, last frames and map point views")
        for l, p in zip(last, pts):
            if len(p.has_point) != l.N:
                raise ValueError("map point view and last frame differ in length")
        n = len(cur)
        ca, la, pa = (_FrameStruct * n)(*[f.c for f in cur]), (_FrameStruct * n)(*[f.c for f in last]), (_MapPointStruct * n)(*[p.c for p in pts])
        outs = [np.full(max(f.N, 1), -1, np.int32) for f in cur]
        op = (C.c_void_p * n)(*[o.ctypes.data for o in outs])
        nm = np.zeros(n, np.int32)
        if cur_src is None and last_src is None:
            check(lib().orbx_search_by_projection(self._h, ca, la, pa, n, float(th), int(bMono), int(self.mbCheckOrientation), op, ptr(nm)))
        else:
            def srcs(v):
                v = [v] if single and (v is None or isinstance(v, tuple)) else (list(v) if v is not None else [None] * n)
                if len(v) != n:
                    raise ValueError("need one frame source (or None) per pair")
                return (_FrameSourceStruct * n)(*[_FrameSourceStruct(None, 0, 0) if x is None else _FrameSourceStruct(x[0]._h, x[1], x[2])
                                                  for x in v])
            check(lib().orbx_search_by_projection_resident(self._h, ca, srcs(cur_src), la, srcs(last_src), pa, n, float(th), int(bMono),
                                                           int(self.mbCheckOrientation), op, ptr(nm)))
        res = [(o[: f.N], int(k)) for o, f, k in zip(outs, cur, nm)]
        return res[0] if single else res
