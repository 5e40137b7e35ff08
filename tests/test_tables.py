"""A0: constructor tables, per-level quotas, level geometry, umax, pattern — closed-form values of SURVEY.md §8."""
import hashlib
import re
import os

import numpy as np
import pytest

import orb_slam2_2021_b200 as pkg
from oracle import oracle as O

QUOTAS = {
    1000: [217, 181, 151, 126, 105, 87, 73, 60],
    2000: [434, 362, 302, 251, 209, 175, 145, 122],
    1200: [261, 217, 181, 151, 126, 105, 87, 72],
    8000: [1737, 1448, 1207, 1005, 838, 698, 582, 485],
}


@pytest.mark.parametrize("nf", list(QUOTAS))
def test_quotas_and_scales(nf):
    o = O.OracleExtractor(nf, 1.2, 8, 20, 7)
    t = o.tables()
    assert list(t["quotas"]) == QUOTAS[nf]
    assert list(t["umax"]) == [15, 15, 15, 15, 14, 14, 14, 13, 13, 12, 11, 10, 9, 8, 6, 3]
    # product library computes the same tables without touching the GPU
    e = pkg.ORBextractor(nf, 1.2, 8, 20, 7)
    assert np.array_equal(e.GetScaleFactors(), t["scale"])
    assert np.array_equal(e.GetInverseScaleFactors(), t["inv_scale"])
    assert np.array_equal(e.GetScaleSigmaSquares(), t["sigma2"])
    assert np.array_equal(e.GetInverseScaleSigmaSquares(), t["inv_sigma2"])
    assert list(e.GetFeaturesPerLevel()) == QUOTAS[nf]
    assert e.GetLevels() == 8
    s = np.float32(1)
    for i in range(8):
        assert t["scale"][i] == s
        s = np.float32(np.float64(s) * np.float64(np.float32(1.2)))


def test_level_sizes_kitti():
    o = O.OracleExtractor(2000, 1.2, 8, 20, 7)
    o(np.zeros((376, 1241), np.uint8))
    sizes = [o.level(l).shape[::-1] for l in range(8)]
    assert sizes == [(1241, 376), (1034, 313), (862, 261), (718, 218), (598, 181), (499, 151), (416, 126), (346, 105)]


def test_max_keypoints_bound():
    e = pkg.ORBextractor(2000, 1.2, 8, 20, 7)
    n = pkg.lib().orbx_max_keypoints(e._h, 376, 1241)
    assert 2000 + 16 <= n <= 2000 + 8 * 16


def test_geometry_errors():
    e = pkg.ORBextractor(1000, 1.2, 8, 20, 7)
    assert pkg.lib().orbx_max_keypoints(e._h, 100, 100) == -3  # level 7 would be 28 px: reference divides by zero
    assert pkg.lib().orbx_max_keypoints(e._h, 2000, 300) == -3  # portrait: nIni == 0
    with pytest.raises(pkg.OrbxError):
        pkg.ORBextractor(1000, 1.0, 8, 20, 7)
    with pytest.raises(pkg.OrbxError):
        pkg.ORBextractor(1000, 1.2, 0, 20, 7)


def test_pattern_checksum():
    p = os.path.join(os.path.dirname(pkg.__file__), "csrc", "orbx_pattern.h")
    txt = open(p).read()
    txt = txt[txt.index("ORBX_PATTERN_VALUES") + len("ORBX_PATTERN_VALUES"):]
    nums = [int(v) for v in re.findall(r"-?\d+", txt)]
    assert len(nums) == 1024 and min(nums) == -13 and max(nums) <= 13
    assert hashlib.sha256(np.array(nums, np.int8).tobytes()).hexdigest() == \
        "2164181aea6ff9ac426ca512d5130d15e1f6e3cd47b1cbdd568bbe1e55d49023"
    assert max(np.hypot(nums[0::2], nums[1::2])) < 19  # samples stay inside the 19-px edge
