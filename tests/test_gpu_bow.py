"""GPU parity for SURVEY.md §8f N2 (ComputeBoW) through include/orbx_bow.h: word ids, node ids, feature lists and the
double-precision values bit for bit against the oracle."""
import numpy as np
import pytest

import orb_slam2_2021_b200 as pkg
from oracle import oracle as O

from . import bow_util as B

pytestmark = pytest.mark.gpu


def same(a, b):
    assert len(a) == len(b) == 5
    for x, y in zip(a, b):
        assert x.dtype == y.dtype and np.array_equal(x, y)


@pytest.mark.parametrize("k,L,levelsup,ragged", [(10, 3, 1, False), (4, 5, 4, False), (10, 2, 4, False), (7, 4, 2, True), (2, 6, 4, False),
                                                 (10, 5, 4, False), (20, 3, 2, True)])
def test_compute_bow_parity(k, L, levelsup, ragged):
    voc_arrays = B.synthetic_vocabulary(k, L, seed=k * 10 + L, ragged=ragged)
    ov, gv = O.Vocabulary(L, *voc_arrays), pkg.ORBVocabulary(L, *voc_arrays)
    assert gv.n_words == ov.n_words
    sets = [B.features_near_words(voc_arrays[2], voc_arrays[1], n, seed=n) if n else np.zeros((0, 32), np.uint8)
            for n in (0, 1, 33, 500, 2000, 3000, 8192)]  # 3000: 48 KB of dynamic shared memory exactly
    for d in sets:  # one frame per call
        same(gv.transform(d, levelsup), ov.transform(d, levelsup))
    for got, d in zip(gv.transform(sets, levelsup), sets):  # all frames in one call (ragged sizes)
        same(got, ov.transform(d, levelsup))


def test_bow_of_extracted_descriptors():
    """Frame::ComputeBoW on real ORB descriptors from the extractor (levelsup = 4 as in src/Frame.cc:452)."""
    from orb_slam2_2021_b200 import synth

    W, H, nf, nl, sf, ini, mn, fx, bf = synth.CONFIGS["kitti"]
    _, desc = pkg.ORBextractor(nf, sf, nl, ini, mn)(synth.synth(W, H, 3))
    arrays = B.synthetic_vocabulary(10, 5, seed=1)
    got = pkg.ORBVocabulary(5, *arrays).transform(desc, 4)
    same(got, O.Vocabulary(5, *arrays).transform(desc, 4))
    assert abs(got[1].sum() - 1.0) < 1e-12


def test_bow_errors():
    p, l, d, w = B.synthetic_vocabulary(3, 2, seed=0)
    bad = p.copy()
    bad[0] = 5
    with pytest.raises(pkg.OrbxError):
        pkg.ORBVocabulary(2, bad, l, d, w)
    l2 = l.copy()
    l2[0] = 1  # an inner node flagged as leaf
    with pytest.raises(pkg.OrbxError):
        pkg.ORBVocabulary(2, p, l2, d, w)
    v = pkg.ORBVocabulary(2, p, l, d, w)
    with pytest.raises(pkg.OrbxError):
        v.transform(np.zeros((9000, 32), np.uint8))


def test_bow_on_device_resident_descriptors():
    """orbx_compute_bow_resident (the descriptors read where the batch engine left them) == orbx_compute_bow on the
    downloaded descriptors == the oracle, left and right side, a frame range in the middle of the batch."""
    from orb_slam2_2021_b200 import synth

    W, H, nf, nl, sf, ini, mn, fx, bf = synth.CONFIGS["euroc"]
    n = 5
    left, right = synth.stereo_batch(W, H, n, seed0=70)
    batch = pkg.StereoFrameBatch(nf, sf, nl, ini, mn, H, W, n, stereo=True)
    out = batch.process(left, right, bf, bf / fx)
    arrays = B.synthetic_vocabulary(10, 4, seed=2)
    gv, ov = pkg.ORBVocabulary(4, *arrays), O.Vocabulary(4, *arrays)
    for side, key in ((0, "left"), (1, "right")):
        cnt = out["n_" + key][1:4]
        got = gv.transform_resident(batch, cnt, first_frame=1, side=side)
        for f, g in zip(range(1, 4), got):
            d = out["desc_" + key][f, : out["n_" + key][f]]
            same(g, gv.transform(d, 4))
            same(g, ov.transform(d, 4))
            assert len(g[0]) > 100
    with pytest.raises(pkg.OrbxError):
        gv.transform_resident(batch, out["n_left"], first_frame=3)  # runs past the batch
