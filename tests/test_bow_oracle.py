"""CPU tests of the ComputeBoW oracle against an independent pure-Python restatement."""
import numpy as np
import pytest

from oracle import oracle as O

from . import bow_util as B


@pytest.mark.parametrize("k,L,levelsup,ragged", [(10, 3, 1, False), (4, 5, 4, False), (10, 2, 4, False), (7, 4, 2, True), (2, 6, 4, False)])
def test_transform_against_python(k, L, levelsup, ragged):
    parent, is_leaf, desc, weight = B.synthetic_vocabulary(k, L, seed=k * 10 + L, ragged=ragged)
    voc = O.Vocabulary(L, parent, is_leaf, desc, weight)
    assert voc.n_words == int(is_leaf.sum())
    for n in (0, 1, 300):
        feats = B.features_near_words(desc, is_leaf, n, seed=n) if n else np.zeros((0, 32), np.uint8)
        wi, wv, fn, fs, ff = voc.transform(feats, levelsup)
        ids, vals, fv = B.py_transform(L, parent, is_leaf, desc, weight, feats, levelsup)
        assert list(wi) == ids and list(wv) == vals  # doubles bit for bit
        assert list(fn) == list(fv)
        for m, node in enumerate(fn):
            assert list(ff[fs[m]:fs[m + 1]]) == fv[node]
        if n == 300:
            assert abs(wv.sum() - 1.0) < 1e-12 and len(wi) < 300  # L1 normalised, repeated words merged
            if levelsup >= L:
                assert list(fn) == [0]  # nid_level <= 0: everything hangs off the root
