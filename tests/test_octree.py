"""DistributeOctTree: oracle (std::list restatement) vs the independent functional model of
SURVEY.md Appendix A (tests/golden/pyref.py), plus the structural properties of the result."""
import numpy as np
import pytest
from hypothesis import given, settings, strategies as st

from oracle import oracle as O
from tests.golden import pyref


def make_cands(rng, n, w, h, clustered=False, max_resp=60):
    if clustered:
        cx, cy = rng.integers(0, w, 6), rng.integers(0, h, 6)
        k = rng.integers(0, 6, n)
        x = np.clip(cx[k] + rng.integers(-9, 10, n), 0, w - 1)
        y = np.clip(cy[k] + rng.integers(-9, 10, n), 0, h - 1)
    else:
        x, y = rng.integers(0, w, n), rng.integers(0, h, n)
    xy = np.unique(np.stack([y, x], 1), axis=0)  # unique positions, row-major like FAST output
    c = np.zeros(len(xy), O.KP_DTYPE)
    c["x"], c["y"] = xy[:, 1], xy[:, 0]
    c["response"] = rng.integers(7, max_resp, len(xy))
    c["size"], c["angle"], c["class_id"] = 7, -1, -1
    return c


def run_both(c, w, h, N):
    got = O.distribute(c, 16, 16 + w, 16, 16 + h, N)
    ref = pyref.distribute([(float(a), float(b), float(r)) for a, b, r in zip(c["x"], c["y"], c["response"])],
                           16, 16 + w, 16, 16 + h, N)
    return got, ref


@pytest.mark.parametrize("seed", range(12))
@pytest.mark.parametrize("clustered", [False, True])
def test_oracle_matches_functional_model(seed, clustered):
    rng = np.random.default_rng(seed)
    w, h = int(rng.integers(60, 1300)), int(rng.integers(60, 400))
    if w < h // 2 + 1:
        w = h
    n = int(rng.integers(1, 3000))
    N = int(rng.integers(1, 600))
    c = make_cands(rng, n, w, h, clustered, max_resp=12 if seed % 2 else 200)  # few response values -> many ties
    got, ref = run_both(c, w, h, N)
    assert [(float(g["x"]), float(g["y"]), float(g["response"])) for g in got] == ref


@settings(max_examples=60, deadline=None)
@given(st.integers(0, 2 ** 31), st.integers(1, 400), st.integers(1, 120))
def test_properties(seed, n, N):
    rng = np.random.default_rng(seed)
    w, h = 400, 150
    c = make_cands(rng, n, w, h, clustered=bool(seed & 1))
    got = O.distribute(c, 16, 16 + w, 16, 16 + h, N)
    nIni = int(np.round(np.float32(w) / np.float32(h)))
    assert len(got) <= max(N + 2, 4 * nIni)
    assert len(got) <= len(c) and (len(got) > 0) == (len(c) > 0)
    # (the loop may stop early when a pass does not grow the list, :672, so len(got) < min(n, N) is legal)
    keys = {(float(g["x"]), float(g["y"])) for g in got}
    assert len(keys) == len(got)  # one key per leaf, no duplicates
    cand = {(float(a), float(b)): float(r) for a, b, r in zip(c["x"], c["y"], c["response"])}
    assert all(cand[(float(g["x"]), float(g["y"]))] == float(g["response"]) for g in got)


def test_empty_and_single():
    c = np.zeros(0, O.KP_DTYPE)
    assert len(O.distribute(c, 16, 416, 16, 166, 100)) == 0
    c = make_cands(np.random.default_rng(0), 1, 400, 150)
    assert len(O.distribute(c, 16, 416, 16, 166, 100)) == 1
