"""The oracle against the committed golden vectors (tests/golden/*.npz, produced by the cv2-driven
model in tests/golden/pyref.py via make_golden.py). Bit-exact on every field."""
import glob
import os

import numpy as np
import pytest

from oracle import oracle as O
from orb_slam2_2021_b200 import synth

GOLD = sorted(glob.glob(os.path.join(os.path.dirname(__file__), "golden", "*_seed*.npz")))


def test_fixtures_present():
    assert len(GOLD) >= 4


@pytest.mark.parametrize("path", GOLD, ids=[os.path.basename(p) for p in GOLD])
def test_oracle_matches_golden(path):
    g = np.load(path)
    name, seed = str(g["config"]), int(g["seed"])
    W, H, nf, nl, sf, ini, mn, fx, bf = synth.CONFIGS[name]
    left, right = synth.stereo_pair(W, H, seed)
    el = O.OracleExtractor(nf, sf, nl, ini, mn)
    kl, dl = el(left)
    assert kl.tobytes() == g["kps_left"].tobytes()
    assert np.array_equal(dl, g["desc_left"])
    if "kps_right" in g:
        er = O.OracleExtractor(nf, sf, nl, ini, mn)
        kr, dr = er(right)
        assert kr.tobytes() == g["kps_right"].tobytes()
        assert np.array_equal(dr, g["desc_right"])
        ur, dp = O.stereo_match(el, er, kl, dl, kr, dr, bf, bf / fx)
        assert np.array_equal(ur, g["u_right"]) and np.array_equal(dp, g["depth"])
        ur, dp = O.stereo_match(el, er, kl, dl, kr, dr, bf, 0.0)  # the reference's uninitialised-mb case
        assert np.array_equal(ur, g["u_right_mb0"]) and np.array_equal(dp, g["depth_mb0"])
