"""ctypes loader for liborbx.so (the C-ABI of include/orbx.h). There is no fallback of any kind:
a missing library or a missing CUDA device raises."""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
SO_PATH = os.path.join(_HERE, "liborbx.so")
CSRC = os.path.join(_HERE, "csrc")

KP_DTYPE = np.dtype(
    [("x", "<f4"), ("y", "<f4"), ("size", "<f4"), ("angle", "<f4"), ("response", "<f4"), ("octave", "<i4"),
     ("class_id", "<i4")]
)
assert KP_DTYPE.itemsize == 28

OK, ERR_ARG, ERR_CAPACITY, ERR_GEOMETRY, ERR_CUDA, ERR_STATE = 0, -1, -2, -3, -4, -5


class OrbxError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__(f"orbx error {code}: {msg}")
        self.code = code


def build(force=False):
    """nvcc -gencode arch=compute_100a,code=sm_100a build of liborbx.so (in-tree, via csrc/Makefile)."""
    if force:
        subprocess.check_call(["make", "-C", CSRC, "clean"])
    subprocess.check_call(["make", "-C", CSRC, "-j8", "-s"])
    return SO_PATH


_lib = None

# name -> (restype, argtypes); the complete export list of include/orbx.h, orbx_frame.h, orbx_input.h and orbx_bow.h
_vp, _i, _f, _sz, _ll = C.c_void_p, C.c_int, C.c_float, C.c_size_t, C.c_longlong
_ip = C.POINTER(C.c_int)
SIGNATURES = {
    "orbx_last_error": (C.c_char_p, []),
    "orbx_device_count": (_i, []),
    "orbx_set_tuning": (_i, [C.c_char_p, _i]),
    "orbx_create": (_i, [_i, _f, _i, _i, _i, _i, C.POINTER(_vp)]),
    "orbx_destroy": (None, [_vp]),
    "orbx_get_levels": (_i, [_vp]),
    "orbx_get_tables": (_i, [_vp, _vp, _vp, _vp, _vp, _vp]),
    "orbx_max_keypoints": (_i, [_vp, _i, _i]),
    "orbx_extract": (_i, [_vp, _vp, _i, _i, _sz, _vp, _vp, _i, _ip]),
    "orbx_pyramid_level_size": (_i, [_vp, _i, _ip, _ip]),
    "orbx_pyramid_level": (_i, [_vp, _i, _vp, _sz]),
    "orbx_stereo_match": (_i, [_vp, _vp, _vp, _vp, _i, _vp, _vp, _i, _f, _f, _vp, _vp]),
    "orbx_stereo_match_resident": (_i, [_vp, _vp, _i, _i, _f, _f, _vp, _vp]),
    "orbx_stereo_frame": (_i, [_vp, _vp, _vp, _vp, _i, _i, _sz, _sz, _f, _f, _vp, _vp, _ip, _vp, _vp, _ip, _i, _vp, _vp]),
    "orbx_stereo_match_static": (_i, [_vp, _vp, _vp, _vp, _i, _vp, _vp, _i, _f, _f, _vp, _vp]),
    "orbx_batch_create": (_i, [_i, _f, _i, _i, _i, _i, _i, _i, _i, _i, C.POINTER(_vp)]),
    "orbx_batch_destroy": (None, [_vp]),
    "orbx_batch_max_keypoints": (_i, [_vp]),
    "orbx_batch_upload": (_i, [_vp, _vp, _vp, _i]),
    "orbx_batch_upload_device": (_i, [_vp, _vp, _vp, _i]),
    "orbx_batch_run": (_i, [_vp, _i, _f, _f, _i]),
    "orbx_batch_download": (_i, [_vp, _i, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp]),
    "orbx_batch_stats": (_i, [_vp, _i, C.POINTER(_ll), C.POINTER(_ll), C.POINTER(_ll)]),
    "orbx_batch_process": (_i, [_vp, _vp, _vp, _i, _f, _f, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp]),
    "orbx_batch_process_async": (_i, [_vp, _vp, _vp, _i, _f, _f, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp]),
    "orbx_batch_wait": (_i, [_vp]),
    "orbx_batch_stream": (_vp, [_vp]),
    "orbx_batch_launches_per_run": (_i, [_vp]),
    "orbx_batch_device_bytes": (_ll, [_vp]),
    "orbx_batch_set_profiling": (_i, [_vp, _i]),
    "orbx_batch_stage_count": (_i, []),
    "orbx_batch_stage_name": (C.c_char_p, [_i]),
    "orbx_batch_stage_times": (_i, [_vp, C.POINTER(C.c_double), _ip]),
    "orbx_host_alloc": (_vp, [_sz]),
    "orbx_host_free": (None, [_vp]),
    "orbx_host_numa_node": (_i, [_ip]),
    "orbx_batch_level_size": (_i, [_vp, _i, _ip, _ip]),
    "orbx_batch_level": (_i, [_vp, _i, _i, _i, _i, _vp, _sz]),
    "orbx_batch_candidates": (_i, [_vp, _i, _i, _i, _vp, _i, _ip]),
    # include/orbx_frame.h
    "orbx_matcher_create": (_i, [_i, C.POINTER(_vp)]),
    "orbx_matcher_destroy": (None, [_vp]),
    "orbx_stereo_from_rgbd": (_i, [_vp, _vp, _vp, _i, _vp, _i, _i, _sz, _f, _vp, _vp]),
    "orbx_assign_features_to_grid": (_i, [_vp, _vp, _vp, _vp, _ip]),
    "orbx_get_features_in_area": (_i, [_vp, _vp, _f, _f, _f, _i, _i, _vp, _i, _ip]),
    "orbx_search_by_projection": (_i, [_vp, _vp, _vp, _vp, _i, _f, _i, _i, _vp, _vp]),
    "orbx_search_by_projection_resident": (_i, [_vp, _vp, _vp, _vp, _vp, _vp, _i, _f, _i, _i, _vp, _vp]),
    "orbx_search_local_points": (_i, [_vp, _vp, _vp, _vp, _i, _f, _f, _vp, _vp]),
    "orbx_search_local_points_resident": (_i, [_vp, _vp, _vp, _vp, _vp, _i, _f, _f, _vp, _vp]),
    # include/orbx_input.h
    "orbx_rectify_gray": (_i, [_vp, _vp, _i, _i, _sz, _i, _i, _vp, _vp, _i, _i, _sz, _i, _i, _vp, _sz]),
    "orbx_batch_set_rectification": (_i, [_vp, _i, _vp, _vp, _i, _i, _sz, _i]),
    "orbx_batch_upload_raw": (_i, [_vp, _vp, _vp, _i, _i, _i, _i, _i]),
    # include/orbx_bow.h
    "orbx_vocabulary_create": (_i, [_i, _i, _i, _vp, _vp, _vp, _vp, C.POINTER(_vp)]),
    "orbx_vocabulary_destroy": (None, [_vp]),
    "orbx_vocabulary_words": (_i, [_vp]),
    "orbx_compute_bow": (_i, [_vp, _vp, _vp, _i, _i, _vp, _vp, _vp, _vp, _vp, _vp, _vp]),
    "orbx_compute_bow_resident": (_i, [_vp, _vp, _i, _i, _vp, _i, _i, _vp, _vp, _vp, _vp, _vp, _vp, _vp]),
}


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(SO_PATH):
            raise ImportError(
                f"{SO_PATH} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
                "(nvcc, sm_100a). orb_slam2_2021_b200 has no CPU or PyTorch fallback."
            )
        L = C.CDLL(SO_PATH)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(L, name)  # AttributeError if the library does not export the symbol
            fn.restype = res
            fn.argtypes = args
        _lib = L
    return _lib


def check(rc):
    if rc != 0:
        raise OrbxError(rc, lib().orbx_last_error().decode(errors="replace"))


def ptr(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)
