"""Image families for the randomised sweeps: the usual synthetic scene plus content chosen to produce TIES (equal FAST scores,
equal Hamming distances, equal SAD sums), empty pyramid levels and saturated pixels -- the places where an implementation that
agrees on ordinary images can still differ from the reference."""
import numpy as np

from orb_slam2_2021_b200 import synth

STYLES = ("scene", "noise", "checker", "quantised", "sparse", "saturated", "repeated")


def stereo_pair(style, W, H, seed):
    rng = np.random.default_rng(seed)
    if style == "scene":
        return synth.stereo_pair(W, H, seed)
    if style == "noise":  # white noise: corners everywhere, quotas and the quadtree saturate
        left = rng.integers(0, 256, (H, W), dtype=np.uint8)
        d = int(rng.integers(0, 40))
        right = np.roll(left, -d, axis=1)
        return left, np.ascontiguousarray(right)
    if style == "checker":  # exact repetition: every corner has the same score, descriptor and SAD
        c = int(rng.integers(3, 24))
        yy, xx = np.mgrid[0:H, 0:W]
        lo, hi = int(rng.integers(0, 100)), int(rng.integers(150, 256))
        left = np.where(((yy // c) + (xx // c)) % 2 == 0, lo, hi).astype(np.uint8)
        d = int(rng.integers(0, 2 * c))
        return left, np.ascontiguousarray(np.roll(left, -d, axis=1))
    if style == "quantised":  # few grey levels: many equal scores
        l, r = synth.stereo_pair(W, H, seed)
        q = int(rng.choice([16, 32, 64]))
        return (l // q * q).astype(np.uint8), (r // q * q).astype(np.uint8)
    if style == "sparse":  # flat with a few blobs: empty cells, empty levels, fewer keypoints than the quota
        left = np.full((H, W), int(rng.integers(0, 256)), np.uint8)
        for _ in range(int(rng.integers(0, 12))):
            x, y, s = int(rng.integers(0, W)), int(rng.integers(0, H)), int(rng.integers(2, 30))
            left[max(0, y - s):y + s, max(0, x - s):x + s] = int(rng.integers(0, 256))
        d = int(rng.integers(0, 30))
        return left, np.ascontiguousarray(np.roll(left, -d, axis=1))
    if style == "saturated":  # contrast stretched until most pixels clip at 0 / 255
        l, r = synth.stereo_pair(W, H, seed)
        g = float(rng.uniform(2, 8))

        def f(a):
            return np.clip((a.astype(np.float32) - 128) * g + 128, 0, 255).astype(np.uint8)

        return f(l), f(r)
    if style == "repeated":  # a small tile repeated: identical descriptors at many places (Hamming and SAD ties in the match)
        t = int(rng.integers(24, 80))
        tile = synth.synth(t + 8, t + 8, seed)[:t, :t]
        left = np.tile(tile, (H // t + 1, W // t + 1))[:H, :W]
        d = int(rng.integers(0, t))
        return np.ascontiguousarray(left), np.ascontiguousarray(np.roll(left, -d, axis=1))
    raise ValueError(style)
