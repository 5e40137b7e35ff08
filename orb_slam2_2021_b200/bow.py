"""Host-side mirror of ORBVocabulary / Frame::ComputeBoW over include/orbx_bow.h (src/Frame.cc:447-454,
Thirdparty/DBoW2/DBoW2/TemplatedVocabulary.h:1140-1272). All compute runs in liborbx.so on the GPU."""
import ctypes as C

import numpy as np

from ._lib import check, lib, ptr


class ORBVocabulary:
    """ORBVocabulary(L, parent, is_leaf, desc, weight): the arrays loadFromTextFile reads, one entry per node after the root."""

    def __init__(self, L, parent, is_leaf, desc, weight, device=0):
        parent = np.ascontiguousarray(parent, dtype=np.int32)
        is_leaf = np.ascontiguousarray(is_leaf, dtype=np.uint8)
        desc = np.ascontiguousarray(desc, dtype=np.uint8).reshape(-1, 32)
        weight = np.ascontiguousarray(weight, dtype=np.float64)
        if not (len(parent) == len(is_leaf) == len(desc) == len(weight)):
            raise ValueError("vocabulary arrays differ in length")
        self._h = C.c_void_p()
        check(lib().orbx_vocabulary_create(device, L, len(parent), ptr(parent), ptr(is_leaf), ptr(desc), ptr(weight), C.byref(self._h)))

    def __del__(self):
        h, self._h = getattr(self, "_h", None), None
        if h:
            lib().orbx_vocabulary_destroy(h)

    @property
    def n_words(self):
        return lib().orbx_vocabulary_words(self._h)

    def transform(self, descriptors, levelsup=4):
        """One descriptor array [n,32] or a list of them -> (word_ids, word_vals, fv_nodes, fv_start, fv_feats) per frame:
        the BowVector sorted by word id and the FeatureVector as CSR sorted by node id."""
        single = isinstance(descriptors, np.ndarray)
        ds = [np.ascontiguousarray(d, dtype=np.uint8).reshape(-1, 32) for d in ([descriptors] if single else descriptors)]
        nf = len(ds)
        n = np.array([len(d) for d in ds], np.int32)
        wi = [np.zeros(max(k, 1), np.int32) for k in n]
        wv = [np.zeros(max(k, 1), np.float64) for k in n]
        fn = [np.zeros(max(k, 1), np.int32) for k in n]
        fs = [np.zeros(k + 1, np.int32) for k in n]
        ff = [np.zeros(max(k, 1), np.int32) for k in n]
        arr = lambda xs: (C.c_void_p * nf)(*[x.ctypes.data for x in xs])  # noqa: E731
        nw, nfv = np.zeros(nf, np.int32), np.zeros(nf, np.int32)
        check(lib().orbx_compute_bow(self._h, arr(ds), ptr(n), nf, int(levelsup), arr(wi), arr(wv), ptr(nw), arr(fn), arr(fs), arr(ff),
                                     ptr(nfv)))
        res = [(wi[f][: nw[f]], wv[f][: nw[f]], fn[f][: nfv[f]], fs[f][: nfv[f] + 1], ff[f][: fs[f][nfv[f]]]) for f in range(nf)]
        return res[0] if single else res

    def transform_resident(self, batch, n, first_frame=0, side=0, levelsup=4):
        """transform() on descriptors that are still on the device: frames first_frame.. of `batch` (a StereoFrameBatch after
        run / process), n[f] = the frames' keypoint counts -> the same tuples as transform(), one per frame."""
        n = np.ascontiguousarray(n, dtype=np.int32)
        nf = len(n)
        wi = [np.zeros(max(k, 1), np.int32) for k in n]
        wv = [np.zeros(max(k, 1), np.float64) for k in n]
        fn = [np.zeros(max(k, 1), np.int32) for k in n]
        fs = [np.zeros(k + 1, np.int32) for k in n]
        ff = [np.zeros(max(k, 1), np.int32) for k in n]
        arr = lambda xs: (C.c_void_p * nf)(*[x.ctypes.data for x in xs])  # noqa: E731
        nw, nfv = np.zeros(nf, np.int32), np.zeros(nf, np.int32)
        check(lib().orbx_compute_bow_resident(self._h, batch._h, int(first_frame), int(side), ptr(n), nf, int(levelsup), arr(wi), arr(wv),
                                              ptr(nw), arr(fn), arr(fs), arr(ff), ptr(nfv)))
        return [(wi[f][: nw[f]], wv[f][: nw[f]], fn[f][: nfv[f]], fs[f][: nfv[f] + 1], ff[f][: fs[f][nfv[f]]]) for f in range(nf)]
