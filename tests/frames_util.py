"""Scenarios and a second, pure-Python restatement for the frame-level rows.

The Python functions below follow /root/reference/src/Frame.cc:279-294, 376-445 and
/root/reference/src/ORBmatcher.cc:1348-1491 line by line with numpy float32 scalars; they exist to cross-check
the C++ oracle (oracle/frame.cpp) on small cases — two independent restatements must agree.
"""
import math

import numpy as np

from oracle import oracle as O

f32 = np.float32
GC, GR = 64, 48


def c_round(x):
    """std::round on a float: half away from zero."""
    x = float(x)
    return int(math.floor(abs(x) + 0.5)) * (1 if x >= 0 else -1)


def py_grid(keys, bounds):
    min_x, max_x, min_y, max_y = [f32(b) for b in bounds]
    inv_w, inv_h = f32(GC) / (max_x - min_x), f32(GR) / (max_y - min_y)
    grid = [[[] for _ in range(GR)] for _ in range(GC)]
    for i, kp in enumerate(keys):
        px, py = c_round((f32(kp["x"]) - min_x) * inv_w), c_round((f32(kp["y"]) - min_y) * inv_h)
        if 0 <= px < GC and 0 <= py < GR:
            grid[px][py].append(i)
    return grid, inv_w, inv_h


def py_features_in_area(keys, grid, inv_w, inv_h, bounds, x, y, r, min_level=-1, max_level=-1):
    min_x, _, min_y, _ = [f32(b) for b in bounds]
    x, y, r = f32(x), f32(y), f32(r)
    out = []
    x0 = max(0, int(math.floor((x - min_x - r) * inv_w)))
    if x0 >= GC:
        return out
    x1 = min(GC - 1, int(math.ceil((x - min_x + r) * inv_w)))
    if x1 < 0:
        return out
    y0 = max(0, int(math.floor((y - min_y - r) * inv_h)))
    if y0 >= GR:
        return out
    y1 = min(GR - 1, int(math.ceil((y - min_y + r) * inv_h)))
    if y1 < 0:
        return out
    check = min_level > 0 or max_level >= 0
    for ix in range(x0, x1 + 1):
        for iy in range(y0, y1 + 1):
            for idx in grid[ix][iy]:
                kp = keys[idx]
                if check:
                    if kp["octave"] < min_level:
                        continue
                    if max_level >= 0 and kp["octave"] > max_level:
                        continue
                if abs(f32(kp["x"]) - x) < r and abs(f32(kp["y"]) - y) < r:
                    out.append(idx)
    return out


def hamming(a, b):
    return int(np.unpackbits(np.bitwise_xor(a, b)).sum())


def py_search_by_projection(cur, last, has_point, world_pos, point_desc, observed, th, b_mono, check_ori):
    """cur/last: dicts with keys_un, desc, u_right, Tcw, fx, fy, cx, cy, mbf, mb, bounds, scale_factors."""
    grid, inv_w, inv_h = py_grid(cur["keys_un"], cur["bounds"])
    tlcz = f32(O.tlc_z(cur["Tcw"], last["Tcw"]))  # OpenCV gemm arithmetic, pinned against cv2 separately
    forward = bool(tlcz > f32(cur["mb"])) and not b_mono
    backward = bool(-tlcz > f32(cur["mb"])) and not b_mono
    assigned = [-1] * len(cur["keys_un"])
    hist = [[] for _ in range(30)]
    factor = f32(1.0) / f32(30)
    nmatches = 0
    fx, fy, cx, cy, mbf = [f32(cur[k]) for k in ("fx", "fy", "cx", "cy", "mbf")]
    min_x, max_x, min_y, max_y = [f32(b) for b in cur["bounds"]]
    for i in range(len(last["keys_un"])):
        if not has_point[i]:
            continue
        xc = O.project(cur["Tcw"], world_pos[i])
        invzc = f32(1.0 / float(xc[2])) if xc[2] != 0 else f32(np.inf if not np.signbit(xc[2]) else -np.inf)
        if invzc < 0:
            continue
        u = f32(f32(f32(fx * xc[0]) * invzc) + cx)
        v = f32(f32(f32(fy * xc[1]) * invzc) + cy)
        if u < min_x or u > max_x or v < min_y or v > max_y:
            continue
        oct_ = int(last["keys_un"][i]["octave"])
        radius = f32(f32(th) * f32(cur["scale_factors"][oct_]))
        if forward:
            cand = py_features_in_area(cur["keys_un"], grid, inv_w, inv_h, cur["bounds"], u, v, radius, oct_, -1)
        elif backward:
            cand = py_features_in_area(cur["keys_un"], grid, inv_w, inv_h, cur["bounds"], u, v, radius, 0, oct_)
        else:
            cand = py_features_in_area(cur["keys_un"], grid, inv_w, inv_h, cur["bounds"], u, v, radius, oct_ - 1, oct_ + 1)
        if not cand:
            continue
        best, best_idx = 256, -1
        for i2 in cand:
            if assigned[i2] >= 0 and observed[assigned[i2]]:
                continue
            if cur["u_right"] is not None and cur["u_right"][i2] > 0:
                ur = f32(u - f32(mbf * invzc))
                if abs(f32(ur - f32(cur["u_right"][i2]))) > radius:
                    continue
            d = hamming(point_desc[i], cur["desc"][i2])
            if d < best:
                best, best_idx = d, i2
        if best <= 100:
            assigned[best_idx] = i
            nmatches += 1
            if check_ori:
                rot = f32(f32(last["keys_un"][i]["angle"]) - f32(cur["keys_un"][best_idx]["angle"]))
                if rot < 0.0:
                    rot = f32(rot + f32(360.0))
                b = c_round(f32(rot * factor))
                if b == 30:
                    b = 0
                hist[b].append(best_idx)
    if check_ori:
        ind = [-1, -1, -1]
        mx = [0, 0, 0]
        for i in range(30):
            s = len(hist[i])
            if s > mx[0]:
                mx = [s, mx[0], mx[1]]
                ind = [i, ind[0], ind[1]]
            elif s > mx[1]:
                mx = [mx[0], s, mx[1]]
                ind = [ind[0], i, ind[1]]
            elif s > mx[2]:
                mx[2], ind[2] = s, i
        if mx[1] < f32(0.1) * f32(mx[0]):
            ind[1] = ind[2] = -1
        elif mx[2] < f32(0.1) * f32(mx[0]):
            ind[2] = -1
        for i in range(30):
            if i not in ind:
                for idx in hist[i]:
                    assigned[idx] = -1
                    nmatches -= 1
    return np.array(assigned, np.int32), nmatches


def random_scenario(seed, n_last=300, n_extra=200, W=640, H=400, tz=0.0, yaw=0.01, clones=30, with_uright=True):
    """Random frame pair: last-frame map points in front of the camera, current keypoints near their projections
    (plus distractors and clones that project to the same spot, so that blocked keypoints occur)."""
    rng = np.random.default_rng(seed)
    fx = fy = 590.08
    cx, cy, bf = W / 2 - 3.3, H / 2 + 1.7, 47.2
    sf = (f32(1.2) ** np.arange(8)).astype(f32)
    c, s = math.cos(yaw), math.sin(yaw)
    Tcw = np.array([[c, 0, s, 0.03], [0, 1, 0, -0.01], [-s, 0, c, -tz]], f32)
    Tlw = np.hstack([np.eye(3, dtype=f32), np.zeros((3, 1), f32)])
    z = rng.uniform(2, 30, n_last).astype(f32)
    ul, vl = rng.uniform(0, W, n_last), rng.uniform(0, H, n_last)
    world = np.stack([(ul - cx) * z / fx, (vl - cy) * z / fy, z], 1).astype(f32)
    if clones:  # clones of earlier points: same world position, nearly the same descriptor
        src = rng.integers(0, n_last - clones, clones)
        world[n_last - clones:] = world[src] + rng.normal(0, 0.002, (clones, 3)).astype(f32)
    last_keys = np.zeros(n_last, O.KP_DTYPE)
    last_keys["x"], last_keys["y"] = ul, vl
    last_keys["octave"] = rng.integers(0, 8, n_last)
    last_keys["angle"] = rng.uniform(0, 360, n_last)
    pdesc = rng.integers(0, 256, (n_last, 32), dtype=np.uint8)
    if clones:
        last_keys["octave"][n_last - clones:] = last_keys["octave"][src]
        last_keys["angle"][n_last - clones:] = last_keys["angle"][src] + rng.normal(0, 2, clones)
        pdesc[n_last - clones:] = pdesc[src]
        flip = rng.integers(0, 256, (clones, 2))
        for k in range(clones):
            pdesc[n_last - clones + k, flip[k, 0] % 32] ^= np.uint8(1 << (flip[k, 1] % 8))
    # current keypoints: projections + jitter, octave +-1, descriptor with a few flipped bits, angle jitter
    xc = world @ Tcw[:, :3].T.astype(np.float64) + Tcw[:, 3]
    up, vp = fx * xc[:, 0] / xc[:, 2] + cx, fy * xc[:, 1] / xc[:, 2] + cy
    n_cur = n_last + n_extra
    cur_keys = np.zeros(n_cur, O.KP_DTYPE)
    cur_keys["x"][:n_last] = up + rng.normal(0, 2.5, n_last)
    cur_keys["y"][:n_last] = vp + rng.normal(0, 2.5, n_last)
    cur_keys["octave"][:n_last] = np.clip(last_keys["octave"] + rng.integers(-1, 2, n_last), 0, 7)
    cur_keys["angle"][:n_last] = (last_keys["angle"] + rng.normal(0, 4, n_last) + rng.choice([0, 0, 0, 40, 200], n_last)) % 360
    cdesc = np.empty((n_cur, 32), np.uint8)
    cdesc[:n_last] = pdesc
    nflip = rng.integers(0, 70, n_last)
    for k in range(n_last):
        bits = rng.choice(256, nflip[k], replace=False)
        for b in bits:
            cdesc[k, b >> 3] ^= np.uint8(1 << (b & 7))
    cur_keys["x"][n_last:] = rng.uniform(-5, W + 5, n_extra)
    cur_keys["y"][n_last:] = rng.uniform(-5, H + 5, n_extra)
    cur_keys["octave"][n_last:] = rng.integers(0, 8, n_extra)
    cur_keys["angle"][n_last:] = rng.uniform(0, 360, n_extra)
    cdesc[n_last:] = rng.integers(0, 256, (n_extra, 32), dtype=np.uint8)
    perm = rng.permutation(n_cur)
    cur_keys, cdesc = cur_keys[perm], cdesc[perm]
    u_right = None
    if with_uright:
        u_right = np.where(rng.random(n_cur) < 0.7, cur_keys["x"] - bf / rng.uniform(2, 30, n_cur), -1).astype(f32)
        zc = np.empty(n_cur)
        inv = np.argsort(perm)
        zc[:] = 10
        zc[inv[:n_last]] = xc[:, 2]
        good = rng.random(n_cur) < 0.5  # half of them consistent with the projected depth
        u_right = np.where(good & (u_right > 0), cur_keys["x"] - bf / zc + rng.normal(0, 1.0, n_cur), u_right).astype(f32)
    bounds = (0.0, float(W), 0.0, float(H))
    common = dict(fx=fx, fy=fy, cx=cx, cy=cy, mbf=bf, mb=bf / fx, bounds=bounds, scale_factors=sf)
    cur = dict(keys_un=cur_keys, desc=cdesc, u_right=u_right, Tcw=Tcw, **common)
    last = dict(keys_un=last_keys, desc=None, u_right=None, Tcw=Tlw, **common)
    has_point = (rng.random(n_last) < 0.9).astype(np.uint8)
    observed = (rng.random(n_last) < 0.6).astype(np.uint8)
    return cur, last, has_point, world, pdesc, observed


def view(cls, d):
    return cls(d["keys_un"], d["desc"], d["u_right"], d["Tcw"], d["fx"], d["fy"], d["cx"], d["cy"], d["mbf"], d["mb"], d["bounds"],
               d["scale_factors"])


def py_search_local_points(F, occupied, valid, proj, level, view_cos, point_desc, observed, th, nnratio):
    """Second restatement of ORBmatcher::SearchByProjection(F, vpMapPoints, th), src/ORBmatcher.cc:45-127."""
    grid, inv_w, inv_h = py_grid(F["keys_un"], F["bounds"])
    n = len(F["keys_un"])
    assigned = [-1] * n
    blocked = [bool(occupied[i]) if occupied is not None else False for i in range(n)]
    nmatches = 0
    for i in range(len(valid)):
        if not valid[i]:
            continue
        lvl = int(level[i])
        r = f32(2.5) if float(view_cos[i]) > 0.998 else f32(4.0)
        if th != 1.0:
            r = f32(r * f32(th))
        radius = f32(r * f32(F["scale_factors"][lvl]))
        cand = py_features_in_area(F["keys_un"], grid, inv_w, inv_h, F["bounds"], proj[i][0], proj[i][1], radius, lvl - 1, lvl)
        if not cand:
            continue
        best, best_l, best2, best_l2, best_idx = 256, -1, 256, -1, -1
        for idx in cand:
            if blocked[idx]:
                continue
            if F["u_right"] is not None and F["u_right"][idx] > 0:
                if abs(f32(f32(proj[i][2]) - f32(F["u_right"][idx]))) > radius:
                    continue
            d = hamming(point_desc[i], F["desc"][idx])
            if d < best:
                best2, best, best_l2, best_l, best_idx = best, d, best_l, int(F["keys_un"][idx]["octave"]), idx
            elif d < best2:
                best_l2, best2 = int(F["keys_un"][idx]["octave"]), d
        if best <= 100:
            if best_l == best_l2 and f32(best) > f32(f32(nnratio) * f32(best2)):
                continue
            assigned[best_idx] = i
            blocked[best_idx] = bool(observed[i])
            nmatches += 1
    return np.array(assigned, np.int32), nmatches


def local_map_scenario(seed, n_points=400, n_extra=200, W=640, H=400, clones=40, with_uright=True):
    """A frame and a local map: every map point projects near one keypoint (and near decoys of similar appearance, so
    the ratio test matters); clones compete for the same keypoint."""
    rng = np.random.default_rng(seed)
    sf = (f32(1.2) ** np.arange(8)).astype(f32)
    proj = np.stack([rng.uniform(0, W, n_points), rng.uniform(0, H, n_points), np.zeros(n_points)], 1).astype(f32)
    level = rng.integers(0, 8, n_points).astype(np.int32)
    pdesc = rng.integers(0, 256, (n_points, 32), dtype=np.uint8)
    if clones:
        src = rng.integers(0, n_points - clones, clones)
        proj[n_points - clones:] = proj[src] + rng.normal(0, 0.5, (clones, 3)).astype(f32)
        level[n_points - clones:] = level[src]
        pdesc[n_points - clones:] = pdesc[src]
        pdesc[n_points - clones:, 0] ^= rng.integers(0, 4, clones).astype(np.uint8)
    n_dec = n_points // 2
    n_cur = n_points + n_dec + n_extra
    keys = np.zeros(n_cur, O.KP_DTYPE)
    desc = np.empty((n_cur, 32), np.uint8)

    def noisy(d, lo, hi):
        out = d.copy()
        for k in range(len(out)):
            for b in rng.choice(256, int(rng.integers(lo, hi)), replace=False):
                out[k, b >> 3] ^= np.uint8(1 << (b & 7))
        return out

    keys["x"][:n_points] = proj[:, 0] + rng.normal(0, 1.5, n_points)
    keys["y"][:n_points] = proj[:, 1] + rng.normal(0, 1.5, n_points)
    keys["octave"][:n_points] = np.clip(level - rng.integers(0, 2, n_points), 0, 7)
    desc[:n_points] = noisy(pdesc, 0, 60)
    dsrc = rng.integers(0, n_points, n_dec)  # decoys: same place, similar descriptor, same or neighbouring level
    keys["x"][n_points:n_points + n_dec] = proj[dsrc, 0] + rng.normal(0, 2.0, n_dec)
    keys["y"][n_points:n_points + n_dec] = proj[dsrc, 1] + rng.normal(0, 2.0, n_dec)
    keys["octave"][n_points:n_points + n_dec] = np.clip(level[dsrc] - rng.integers(0, 2, n_dec), 0, 7)
    desc[n_points:n_points + n_dec] = noisy(pdesc[dsrc], 20, 80)
    keys["x"][n_points + n_dec:] = rng.uniform(-5, W + 5, n_extra)
    keys["y"][n_points + n_dec:] = rng.uniform(-5, H + 5, n_extra)
    keys["octave"][n_points + n_dec:] = rng.integers(0, 8, n_extra)
    desc[n_points + n_dec:] = rng.integers(0, 256, (n_extra, 32), dtype=np.uint8)
    perm = rng.permutation(n_cur)
    keys, desc = keys[perm], desc[perm]
    u_right = None
    if with_uright:
        u_right = np.where(rng.random(n_cur) < 0.6, keys["x"] - rng.uniform(1, 20, n_cur), -1).astype(f32)
        inv = np.argsort(perm)
        proj[:, 2] = u_right[inv[:n_points]] + rng.normal(0, 2.0, n_points)  # mTrackProjXR near the keypoint's uRight
    bounds = (0.0, float(W), 0.0, float(H))
    F = dict(keys_un=keys, desc=desc, u_right=u_right, Tcw=np.eye(4, dtype=f32)[:3], fx=500.0, fy=500.0, cx=W / 2, cy=H / 2, mbf=40.0,
             mb=0.08, bounds=bounds, scale_factors=sf)
    occupied = (rng.random(n_cur) < 0.1).astype(np.uint8)
    valid = (rng.random(n_points) < 0.9).astype(np.uint8)
    view_cos = np.where(rng.random(n_points) < 0.5, 0.9995, 0.97).astype(f32)
    observed = (rng.random(n_points) < 0.8).astype(np.uint8)
    return F, occupied, valid, proj, level, view_cos, pdesc, observed
