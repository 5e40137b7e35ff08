"""Writes tests/golden/frame_search.npz: outputs of the oracle's SearchByProjection on seeded random scenarios
(tests/frames_util.py). Run from the repo root:  python -m tests.golden.make_frame_golden"""
import os

import numpy as np

from oracle import oracle as O
from tests import frames_util as U

CASES = [(11, 0.0, 7, 0, 1), (12, 0.5, 7, 0, 1), (13, -0.5, 15, 0, 1), (14, 0.5, 15, 1, 0), (15, 0.0, 7, 0, 0)]


def main():
    out = {"n_cases": np.int32(len(CASES))}
    for k, (seed, tz, th, mono, ori) in enumerate(CASES):
        cur, last, has, world, pdesc, obs = U.random_scenario(seed, tz=tz)
        m, n = O.search_by_projection(U.view(O.FrameView, cur), U.view(O.FrameView, last), has, world, pdesc, obs, th, bool(mono), bool(ori))
        for name, v in (("seed", seed), ("tz", tz), ("th", th), ("mono", mono), ("ori", ori), ("nmatches", n)):
            out[f"case{k}_{name}"] = np.float64(v)
        out[f"case{k}_match"] = m
        print(f"case {k}: seed {seed} tz {tz} th {th} mono {mono} ori {ori} -> {n} matches")
    np.savez_compressed(os.path.join(os.path.dirname(__file__), "frame_search.npz"), **out)


if __name__ == "__main__":
    main()
