"""ctypes binding of oracle/_ref/libref_orbextractor.so: the REFERENCE'S OWN ORB_SLAM2::ORBextractor
(/root/reference/src/ORBextractor.cc compiled where it lies, see oracle/Makefile and oracle/ref_wrap.cpp) over the
OpenCV stand-in of oracle/cvshim/. TEST INFRASTRUCTURE ONLY: tests/, smoke() and bench.py's CPU legs may import it.

/root/reference exists only in the build container; the built library is git-ignored but travels to the GPU box.
`available()` says whether the library can be used here; `build()` (re)builds it where the sources exist.
"""
import ctypes as C
import os
import subprocess

import numpy as np

from .oracle import KP_DTYPE, _ptr, _u8

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "_ref", "libref_orbextractor.so")
_REF_SRC = "/root/reference/src/ORBextractor.cc"


def build(force=False):
    """Compile the reference translation unit against the shim (only where /root/reference exists)."""
    if os.path.exists(_REF_SRC) and (force or not os.path.exists(_SO)):
        subprocess.check_call(["make", "-C", _HERE, "-B" if force else "-s", "_ref"])
    return _SO if os.path.exists(_SO) else None


def available():
    return os.path.exists(_SO) or os.path.exists(_REF_SRC)


_lib = None


def lib():
    global _lib
    if _lib is None:
        if build() is None:
            raise RuntimeError("oracle/_ref/libref_orbextractor.so is missing and /root/reference is not here to build it")
        L = C.CDLL(_SO)
        vp, f32p, i32p = C.c_void_p, C.POINTER(C.c_float), C.POINTER(C.c_int)
        L.ref_create.restype = vp
        L.ref_create.argtypes = [C.c_int, C.c_float, C.c_int, C.c_int, C.c_int]
        L.ref_destroy.argtypes = [vp]
        L.ref_tables.argtypes = [vp, f32p, f32p, f32p, f32p]
        L.ref_extract.argtypes = [vp, vp, C.c_int, C.c_int, C.c_int, vp, vp, C.c_int]
        L.ref_level_size.argtypes = [vp, C.c_int, i32p, i32p]
        L.ref_level_copy.argtypes = [vp, C.c_int, vp, C.c_int]
        _lib = L
    return _lib


class RefExtractor:
    """ORB_SLAM2::ORBextractor itself (include/ORBextractor.h:56-100): same constructor, operator() as __call__."""

    def __init__(self, nfeatures, scale_factor, nlevels, ini_th_fast, min_th_fast):
        self.nfeatures, self.nlevels = nfeatures, nlevels
        self._h = lib().ref_create(nfeatures, scale_factor, nlevels, ini_th_fast, min_th_fast)

    def __del__(self):
        if getattr(self, "_h", None):
            lib().ref_destroy(self._h)
            self._h = None

    def tables(self):
        f = [np.zeros(self.nlevels, np.float32) for _ in range(4)]
        lib().ref_tables(self._h, *[a.ctypes.data_as(C.POINTER(C.c_float)) for a in f])
        return dict(scale=f[0], inv_scale=f[1], sigma2=f[2], inv_sigma2=f[3])

    def __call__(self, image):
        img = _u8(image)
        cap = self.nfeatures + 64 * self.nlevels + 64
        kps = np.zeros(cap, KP_DTYPE)
        desc = np.zeros((cap, 32), np.uint8)
        n = lib().ref_extract(self._h, _ptr(img), img.shape[0], img.shape[1], img.strides[0], _ptr(kps), _ptr(desc), cap)
        if n < 0:
            raise RuntimeError("ref_extract: capacity too small")
        return kps[:n].copy(), desc[:n].copy()

    def level(self, l):
        r, c = C.c_int(), C.c_int()
        if lib().ref_level_size(self._h, l, C.byref(r), C.byref(c)) != 0:
            raise IndexError(l)
        out = np.zeros((r.value, c.value), np.uint8)
        lib().ref_level_copy(self._h, l, _ptr(out), out.strides[0])
        return out
