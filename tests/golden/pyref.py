"""cv2-driven Python model of the hot path, used ONLY to generate and cross-check golden vectors.

Every pixel primitive is the real OpenCV kernel the reference would call (cv2 4.13.0: resize, FAST,
GaussianBlur, fastAtan2); the control flow around them follows /root/reference/src/ORBextractor.cc
and src/Frame.cc. The quadtree is written from SURVEY.md Appendix A's functional description (lists
and reversals), i.e. independently of the std::list restatement in oracle/distribute.cpp, so the
two cross-validate each other. cos/sin use float64 numpy rounded to float32 (correctly rounded with
overwhelming probability), which is what orbx_sincosf is specified to return.
"""
import math

import cv2
import numpy as np

KP_DTYPE = np.dtype(
    [("x", "<f4"), ("y", "<f4"), ("size", "<f4"), ("angle", "<f4"), ("response", "<f4"), ("octave", "<i4"),
     ("class_id", "<i4")]
)
f32 = np.float32


def load_pattern():
    import os
    import re

    here = os.path.dirname(os.path.abspath(__file__))
    txt = open(os.path.join(here, "..", "..", "orb_slam2_2021_b200", "csrc", "orbx_pattern.h")).read()
    txt = txt[txt.index("ORBX_PATTERN_VALUES") + len("ORBX_PATTERN_VALUES"):]
    nums = [int(v) for v in re.findall(r"-?\d+", txt)]
    assert len(nums) == 1024
    return np.array(nums, np.int32).reshape(256, 2, 2)  # [pair, point, (x,y)]


def cv_round(v):
    return int(np.rint(f32(v)))  # half to even


class PyRefExtractor:
    def __init__(self, nfeatures, scale_factor, nlevels, ini_th, min_th):
        self.nfeatures, self.nlevels, self.ini_th, self.min_th = nfeatures, nlevels, ini_th, min_th
        sfd = float(f32(scale_factor))
        self.scale = np.ones(nlevels, f32)
        for i in range(1, nlevels):
            self.scale[i] = f32(float(self.scale[i - 1]) * sfd)
        self.inv_scale = (f32(1.0) / self.scale).astype(f32)
        factor = f32(1.0 / sfd)
        desired = f32(f32(f32(nfeatures) * f32(f32(1) - factor)) / f32(f32(1) - f32(float(factor) ** nlevels)))
        self.quota = []
        for _ in range(nlevels - 1):
            self.quota.append(cv_round(desired))
            desired = f32(desired * factor)
        self.quota.append(max(nfeatures - sum(self.quota), 0))
        self.umax = [15, 15, 15, 15, 14, 14, 14, 13, 13, 12, 11, 10, 9, 8, 6, 3]
        self.pattern = load_pattern()
        self.fast = {t: cv2.FastFeatureDetector_create(t, True, cv2.FAST_FEATURE_DETECTOR_TYPE_9_16)
                     for t in (ini_th, min_th)}
        self.levels = []

    # ComputePyramid :1105-1135
    def pyramid(self, img):
        self.levels = []
        H, W = img.shape
        for l in range(self.nlevels):
            w = cv_round(f32(W) * self.inv_scale[l])
            h = cv_round(f32(H) * self.inv_scale[l])
            if l == 0:
                self.levels.append(img.copy())
            else:
                self.levels.append(cv2.resize(self.levels[l - 1], (w, h), interpolation=cv2.INTER_LINEAR))

    # cell loop :768-832
    def detect(self, im):
        h, w = im.shape
        minB, maxBX, maxBY = 16, w - 16, h - 16
        width, height = f32(maxBX - minB), f32(maxBY - minB)
        nCols, nRows = int(width / f32(30)), int(height / f32(30))
        wCell, hCell = math.ceil(width / f32(nCols)), math.ceil(height / f32(nRows))
        out = []
        for i in range(nRows):
            iniY = minB + i * hCell
            maxY = iniY + hCell + 6
            if iniY >= maxBY - 3:
                continue
            maxY = min(maxY, maxBY)
            for j in range(nCols):
                iniX = minB + j * wCell
                maxX = iniX + wCell + 6
                if iniX >= maxBX - 6:
                    continue
                maxX = min(maxX, maxBX)
                roi = im[iniY:maxY, iniX:maxX]
                kps = self.fast[self.ini_th].detect(roi)
                if not kps:
                    kps = self.fast[self.min_th].detect(roi)
                for k in kps:
                    out.append((k.pt[0] + j * wCell, k.pt[1] + i * hCell, k.response))
        return out

    def ic_angle(self, im, x, y):
        cx, cy = cv_round(x), cv_round(y)
        m10 = m01 = 0
        for v in range(-15, 16):
            d = self.umax[abs(v)]
            row = im[cy + v, cx - d:cx + d + 1].astype(np.int64)
            m10 += int((np.arange(-d, d + 1) * row).sum())
            m01 += v * int(row.sum())
        return cv2.fastAtan2(float(m01), float(m10))

    def descriptor(self, blur, x, y, angle):
        factor_pi = f32(np.pi / float(f32(180.0)))
        a_rad = f32(f32(angle) * factor_pi)
        a = f32(np.cos(np.float64(a_rad)))
        b = f32(np.sin(np.float64(a_rad)))
        px = self.pattern[..., 0].astype(f32)
        py = self.pattern[..., 1].astype(f32)
        ry = np.rint((px * b).astype(f32) + (py * a).astype(f32)).astype(np.int32)
        rx = np.rint((px * a).astype(f32) - (py * b).astype(f32)).astype(np.int32)
        vals = blur[cv_round(y) + ry, cv_round(x) + rx]  # [256, 2]
        bits = (vals[:, 0] < vals[:, 1]).astype(np.uint8)
        return np.packbits(bits.reshape(32, 8), axis=1, bitorder="little").reshape(32)

    def __call__(self, img):
        self.pyramid(img)
        kps, descs = [], []
        for l, im in enumerate(self.levels):
            h, w = im.shape
            cand = self.detect(im)
            kept = distribute(cand, 16, w - 16, 16, h - 16, self.quota[l])
            if not kept:
                continue
            blur = cv2.GaussianBlur(im.copy(), (7, 7), 2, sigmaY=2, borderType=cv2.BORDER_REFLECT_101)
            size = f32(int(f32(31) * self.scale[l]))
            for (x, y, r) in kept:
                x, y = f32(x + 16), f32(y + 16)
                ang = self.ic_angle(im, x, y)
                descs.append(self.descriptor(blur, x, y, ang))
                if l:
                    x, y = f32(x * self.scale[l]), f32(y * self.scale[l])
                kps.append((x, y, size, ang, r, l, -1))
        K = np.array(kps, KP_DTYPE) if kps else np.zeros(0, KP_DTYPE)
        D = np.array(descs, np.uint8).reshape(-1, 32)
        return K, D


# ---- DistributeOctTree, functional form of SURVEY.md Appendix A ------------------------------------
def _split(node, cand):
    x0, y0, x1, y1, keys = node
    mx = x0 + math.ceil(f32(x1 - x0) / f32(2))
    my = y0 + math.ceil(f32(y1 - y0) / f32(2))
    boxes = [(x0, y0, mx, my), (mx, y0, x1, my), (x0, my, mx, y1), (mx, my, x1, y1)]
    parts = [[], [], [], []]
    for k in keys:
        q = (0 if f32(cand[k][0]) < f32(mx) else 1) + (0 if f32(cand[k][1]) < f32(my) else 2)
        parts[q].append(k)
    return [(b[0], b[1], b[2], b[3], p) for b, p in zip(boxes, parts) if p]


def distribute(cand, minX, maxX, minY, maxY, N):
    nIni = int(np.round(f32(maxX - minX) / f32(maxY - minY)))
    hX = f32(f32(maxX - minX) / f32(nIni))
    ini = [[int(f32(hX * f32(i))), 0, int(f32(hX * f32(i + 1))), maxY - minY, []] for i in range(nIni)]
    for k, c in enumerate(cand):
        ini[int(f32(c[0]) / hX)][4].append(k)
    L = [tuple(n) for n in ini if n[4]]
    while True:
        multi = [n for n in L if len(n[4]) > 1]
        singles = [n for n in L if len(n[4]) == 1]
        children = []
        for n in multi:
            children += _split(n, cand)
        L2 = children[::-1] + singles
        V = [c for c in children if len(c[4]) > 1]  # creation order
        done = len(L2) >= N or len(L2) == len(L)
        L = L2
        if done:
            break
        if len(L) + 3 * len(V) > N:
            while True:
                prev = len(L)
                order = sorted(range(len(V)), key=lambda i: (len(V[i][4]), i))[::-1]  # largest, later-created first
                new_children, removed = [], set()
                size = len(L)
                for i in order:
                    ch = _split(V[i], cand)
                    new_children += ch
                    removed.add(id(V[i]))
                    size += len(ch) - 1
                    if size >= N:
                        break
                L = new_children[::-1] + [n for n in L if id(n) not in removed]
                V = [c for c in new_children if len(c[4]) > 1]
                if len(L) >= N or len(L) == prev:
                    break
            break
    out = []
    for n in L:
        best = n[4][0]
        for k in n[4][1:]:
            if cand[k][2] > cand[best][2]:
                best = k
        out.append(cand[best])
    return out


# ---- Frame::ComputeStereoMatches :522-700 ------------------------------------------------------------
def stereo_match(exL, exR, kL, dL, kR, dR, mbf, mb):
    n = len(kL)
    uR = np.full(n, -1, f32)
    depth = np.full(n, -1, f32)
    nRows = exL.levels[0].shape[0]
    rows = [[] for _ in range(nRows)]
    for iR in range(len(kR)):
        y = f32(kR["y"][iR])
        r = f32(f32(2.0) * exL.scale[kR["octave"][iR]])
        for yi in range(int(math.floor(f32(y - r))), int(math.ceil(f32(y + r))) + 1):
            rows[yi].append(iR)
    mbf, mb = f32(mbf), f32(mb)
    with np.errstate(divide="ignore", invalid="ignore"):
        maxD = f32(mbf / mb)
    pairs = []
    bits_r = np.unpackbits(dR, axis=1) if len(dR) else np.zeros((0, 256), np.uint8)
    for iL in range(n):
        uL, vL, lvl = f32(kL["x"][iL]), f32(kL["y"][iL]), int(kL["octave"][iL])
        cands = rows[int(vL)]
        if not cands:
            continue
        minU, maxU = f32(uL - maxD), uL
        if maxU < 0:
            continue
        best, bestR = 100, 0
        bl = np.unpackbits(dL[iL])
        for iR in cands:
            if abs(int(kR["octave"][iR]) - lvl) > 1:
                continue
            u = f32(kR["x"][iR])
            if u >= minU and u <= maxU:
                d = int((bl != bits_r[iR]).sum())
                if d < best:
                    best, bestR = d, iR
        if best >= 75:
            continue
        s = exL.inv_scale[lvl]

        def rnd(v):  # ::round, half away from zero
            v = float(f32(v))
            return f32(math.floor(abs(v) + 0.5) * (1 if v >= 0 else -1))

        su, sv, sr = rnd(f32(uL * s)), rnd(f32(vL * s)), rnd(f32(f32(kR["x"][bestR]) * s))
        IL, IR = exL.levels[lvl], exR.levels[lvl]
        if sr < 0 or sr + 11 >= IR.shape[1]:
            continue
        cu, cv, cr = int(su), int(sv), int(sr)
        pl = IL[cv - 5:cv + 6, cu - 5:cu + 6].astype(np.float32)
        pl = pl - pl[5, 5]
        dists = []
        for inc in range(-5, 6):
            pr = IR[cv - 5:cv + 6, cr + inc - 5:cr + inc + 6].astype(np.float32)
            pr = pr - pr[5, 5]
            dists.append(f32(cv2.norm(pl, pr, cv2.NORM_L1)))
        bestSad, bestInc = 2 ** 31 - 1, 0
        for inc in range(-5, 6):
            if dists[inc + 5] < bestSad:
                bestSad, bestInc = int(dists[inc + 5]), inc
        if bestInc in (-5, 5):
            continue
        d1, d2, d3 = dists[bestInc + 4], dists[bestInc + 5], dists[bestInc + 6]
        with np.errstate(divide="ignore", invalid="ignore"):
            delta = f32(f32(d1 - d3) / f32(f32(2.0) * f32(f32(d1 + d3) - f32(f32(2.0) * d2))))
        if delta < -1 or delta > 1:
            continue
        bestu = f32(exL.scale[lvl] * f32(f32(sr + f32(bestInc)) + delta))
        disp = f32(uL - bestu)
        if disp >= 0 and disp < maxD:
            if disp <= 0:
                disp = f32(0.01)
                bestu = f32(float(uL) - 0.01)
            depth[iL] = f32(mbf / disp)
            uR[iL] = bestu
            pairs.append((bestSad, iL))
    if pairs:
        pairs.sort()
        median = f32(pairs[len(pairs) // 2][0])
        th = f32(f32(f32(1.5) * f32(1.4)) * median)
        for sad, iL in pairs:
            if f32(sad) >= th:
                uR[iL] = -1
                depth[iL] = -1
    return uR, depth
