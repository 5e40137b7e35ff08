"""The C-ABI library loads and exports every symbol include/orbx.h and include/orbx_frame.h declare; error behaviour that
needs no GPU. No compute calls here."""
import ctypes
import os
import re

import numpy as np
import pytest

import orb_slam2_2021_b200 as pkg
from orb_slam2_2021_b200 import _lib

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_symbols():
    syms = set()
    for name in ("orbx.h", "orbx_frame.h", "orbx_input.h", "orbx_bow.h"):  # the C-ABI headers (ORBextractorGPU.h is a C++ shim over them)
        hdr = open(os.path.join(ROOT, "include", name)).read()
        hdr = re.sub(r"/\*.*?\*/", "", hdr, flags=re.S)
        syms |= set(re.findall(r"\b(orbx_[a-z0-9_]+)\s*\(", hdr))
    return sorted(syms)


def test_header_symbols_exported():
    syms = declared_symbols()
    assert len(syms) >= 25
    L = ctypes.CDLL(pkg.SO_PATH)
    for s in syms:
        assert hasattr(L, s), f"{s} declared in include/*.h but not exported by liborbx.so"
    assert set(syms) == set(_lib.SIGNATURES), set(syms) ^ set(_lib.SIGNATURES)


def test_no_oracle_in_product():
    """The product path must not import or link the oracle (or cv2)."""
    for dirpath, _, files in os.walk(os.path.join(ROOT, "orb_slam2_2021_b200")):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h", ".cpp")) and f != "synth.py":
                txt = open(os.path.join(dirpath, f), errors="replace").read()
                assert "import cv2" not in txt, f
                assert not re.search(r"^\s*(from|import)\s+oracle", txt, flags=re.M), f
                assert '#include "../../oracle' not in txt and "orb_oracle" not in txt, f


def test_keypoint_layout():
    assert pkg.KP_DTYPE.itemsize == 28
    assert [pkg.KP_DTYPE.fields[n][1] for n in ("x", "y", "size", "angle", "response", "octave", "class_id")] == \
        [0, 4, 8, 12, 16, 20, 24]


def test_empty_image_is_silent():
    e = pkg.ORBextractor(1000, 1.2, 8, 20, 7)
    k, d = e(np.zeros((0, 0), np.uint8))
    assert len(k) == 0 and d.shape == (0, 32)
    n = ctypes.c_int(-7)
    assert pkg.lib().orbx_extract(e._h, None, 0, 0, 0, None, None, 0, ctypes.byref(n)) == 0
    assert n.value == 0  # src/ORBextractor.cc:1044-1045 returns silently; an out-count is set to 0 rather than left undefined


def test_bad_arguments():
    e = pkg.ORBextractor(1000, 1.2, 8, 20, 7)
    with pytest.raises(TypeError):
        e(np.zeros((10, 10), np.float32))
    L = pkg.lib()
    assert L.orbx_pyramid_level_size(e._h, 0, ctypes.byref(ctypes.c_int()), ctypes.byref(ctypes.c_int())) == _lib.ERR_STATE
    assert b"no image" in L.orbx_last_error()
    assert L.orbx_get_tables(None, None, None, None, None, None) == _lib.ERR_ARG


def test_fails_loudly_without_gpu(has_gpu):
    if has_gpu:
        pytest.skip("GPU present")
    e = pkg.ORBextractor(1000, 1.2, 8, 20, 7)
    with pytest.raises(pkg.OrbxError) as ei:
        e(np.zeros((480, 640), np.uint8))
    assert ei.value.code == _lib.ERR_CUDA
    with pytest.raises(pkg.OrbxError):
        pkg.StereoFrameBatch(1000, 1.2, 8, 20, 7, 480, 640, 2)


def test_quota_beyond_the_quadtree_capacity_is_a_clean_error():
    """Per-level quotas whose node list does not fit one CTA's shared memory (~2600 nodes) are served from a global-memory list;
    only a list beyond the 16-bit positions (65000 nodes per level) is refused, with ORBX_ERR_ARG and a text."""
    assert pkg.lib().orbx_max_keypoints(pkg.ORBextractor(20000, 1.2, 8, 20, 7)._h, 2160, 3840) > 20000
    assert pkg.lib().orbx_max_keypoints(pkg.ORBextractor(3000, 1.2, 1, 20, 7)._h, 480, 640) > 3000
    e = pkg.ORBextractor(400000, 1.2, 8, 20, 7)
    rc = pkg.lib().orbx_max_keypoints(e._h, 2160, 3840)
    assert rc == pkg._lib.ERR_ARG
    assert b"quota" in pkg.lib().orbx_last_error()


@pytest.mark.gpu
def test_oversized_quota_fails_cleanly_on_the_device_too():
    """nfeatures = 400000 at 8 levels: orbx_extract reports ORBX_ERR_ARG with the quota text instead of a launch failure,
    and quotas above the shared-memory list (20000 features) run from the global-memory list."""
    from orb_slam2_2021_b200 import synth

    e = pkg.ORBextractor(400000, 1.2, 8, 20, 7)
    with pytest.raises(pkg.OrbxError) as ei:
        e(synth.synth(1280, 720, 1))
    assert ei.value.code == pkg._lib.ERR_ARG and "quota" in str(ei.value)
    k, d = pkg.ORBextractor(20000, 1.2, 8, 20, 7)(synth.synth(1280, 720, 1))
    assert len(k) > 3000 and d.shape == (len(k), 32)
    ok = pkg.ORBextractor(12000, 1.2, 8, 20, 7)
    k, d = ok(synth.synth(1280, 720, 1))
    assert len(k) > 3000 and d.shape == (len(k), 32)
