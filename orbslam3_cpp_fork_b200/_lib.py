"""ctypes binding of libcuorb.so — exactly the entry points include/cuorb.h declares.

There is no CPU fallback: if the shared library is missing this module raises, and every call that
needs a GPU fails with the library's own error when no CUDA device is usable.
"""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

_PKG = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_PKG, "libcuorb.so")

KP_DTYPE = np.dtype(
    [("x", "<f4"), ("y", "<f4"), ("size", "<f4"), ("angle", "<f4"), ("response", "<f4"),
     ("octave", "<i4"), ("class_id", "<i4")]
)
assert KP_DTYPE.itemsize == 28  # cv::KeyPoint

E_EMPTY, E_BADARG, E_CAPACITY, E_CUDA, E_UNSUPPORTED, E_STATE = -1, -2, -3, -4, -5, -6
MEM_HOST, MEM_DEVICE = 0, 1
STAGES = ("h2d", "pyramid", "fast", "octree", "blur", "describe", "stereo")
MP_IN_VIEW, MP_BAD, MP_HAS_OBS, MP_IN_VIEW_R = 1, 2, 4, 8


class CuorbError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__(f"libcuorb error {code}: {msg}")
        self.code = code


class Config(C.Structure):
    _fields_ = [("nfeatures", C.c_int32), ("scale_factor", C.c_float), ("nlevels", C.c_int32),
                ("ini_th_fast", C.c_int32), ("min_th_fast", C.c_int32), ("max_width", C.c_int32),
                ("max_height", C.c_int32), ("max_batch", C.c_int32), ("device", C.c_int32)]


class FrustumCamera(C.Structure):
    """cuorb_frustum_camera: the Frame members Frame::isInFrustum reads."""

    _fields_ = [("rcw", C.c_float * 9), ("tcw", C.c_float * 3), ("ow", C.c_float * 3),
                ("fx", C.c_float), ("fy", C.c_float), ("cx", C.c_float), ("cy", C.c_float), ("mbf", C.c_float),
                ("min_x", C.c_float), ("max_x", C.c_float), ("min_y", C.c_float), ("max_y", C.c_float),
                ("log_scale_factor", C.c_float), ("n_levels", C.c_int32)]


class FrameView(C.Structure):
    _fields_ = [
        ("n", C.c_int32), ("keys_un", C.c_void_p), ("descriptors", C.c_void_p),
        ("u_right", C.c_void_p), ("bound_with_obs", C.c_void_p),
        ("min_x", C.c_float), ("min_y", C.c_float), ("max_x", C.c_float), ("max_y", C.c_float),
        ("grid_w_inv", C.c_float), ("grid_h_inv", C.c_float), ("mb", C.c_float), ("mbf", C.c_float),
        ("nlevels", C.c_int32), ("scale_factors", C.c_void_p),
    ]


# name -> (restype, argtypes); also the list tests check against include/cuorb.h
SIGNATURES = {
    "cuorb_last_error": (C.c_char_p, []),
    "cuorb_version": (C.c_int, []),
    "cuorb_device_count": (C.c_int, []),
    "cuorb_host_alloc": (C.c_void_p, [C.c_size_t]),
    "cuorb_host_free": (None, [C.c_void_p]),
    "cuorb_extractor_create": (C.c_void_p, [C.POINTER(Config)]),
    "cuorb_extractor_destroy": (None, [C.c_void_p]),
    "cuorb_extractor_get_tables": (C.c_int, [C.c_void_p] + [C.c_void_p] * 5),
    "cuorb_extractor_capacity": (C.c_int, [C.c_void_p]),
    "cuorb_extract": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_size_t, C.c_int, C.c_int,
                                C.c_void_p, C.c_void_p, C.c_int, C.POINTER(C.c_int)]),
    "cuorb_extract_batch_async": (C.c_int, [C.c_void_p, C.c_int, C.c_void_p, C.c_int, C.c_int, C.c_size_t,
                                            C.c_size_t, C.c_int, C.c_int, C.c_int]),
    "cuorb_extract_fetch": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p]),
    "cuorb_extract_fetch_async": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p]),
    "cuorb_sync": (C.c_int, [C.c_void_p]),
    "cuorb_extractor_status": (C.c_int, [C.c_void_p]),
    "cuorb_extractor_stream": (C.c_void_p, [C.c_void_p]),
    "cuorb_extractor_pyramid": (C.c_int, [C.c_void_p, C.c_int, C.c_int, C.POINTER(C.c_void_p),
                                          C.POINTER(C.c_int), C.POINTER(C.c_int), C.POINTER(C.c_size_t)]),
    "cuorb_extractor_device_features": (C.c_int, [C.c_void_p, C.c_int] + [C.POINTER(C.c_void_p)] * 4),
    "cuorb_extractor_download_plane": (C.c_int64, [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_void_p,
                                                   C.c_size_t, C.POINTER(C.c_int), C.POINTER(C.c_int)]),
    "cuorb_extractor_download_candidates": (C.c_int, [C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_void_p,
                                                      C.c_void_p, C.c_int]),
    "cuorb_stereo_match_batch_async": (C.c_int, [C.c_void_p, C.c_void_p, C.c_float, C.c_float]),
    "cuorb_stereo_fetch": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p]),
    "cuorb_stereo_fetch_async": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int]),
    "cuorb_stereo_match": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_int,
                                     C.c_void_p, C.c_void_p, C.c_float, C.c_float, C.c_void_p, C.c_void_p]),
    "cuorb_extract_stereo": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_size_t, C.c_float,
                                      C.c_float, C.c_void_p, C.c_void_p, C.POINTER(C.c_int), C.c_void_p, C.c_void_p,
                                      C.POINTER(C.c_int), C.c_int, C.c_void_p, C.c_void_p]),
    "cuorb_hamming256": (C.c_int, [C.c_void_p, C.c_void_p]),
    "cuorb_hamming_batch": (C.c_int, [C.c_int, C.c_void_p, C.c_int, C.c_void_p, C.c_int, C.c_void_p,
                                      C.c_void_p, C.c_int, C.c_void_p]),
    "cuorb_search_by_projection": (C.c_int, [C.c_int, C.POINTER(FrameView), C.c_int] + [C.c_void_p] * 8 +
                                   [C.c_float, C.c_int, C.c_float, C.c_float, C.c_void_p]),
    "cuorb_search_by_projection_frame": (C.c_int, [C.c_int, C.POINTER(FrameView), C.c_int] + [C.c_void_p] * 7 +
                                         [C.c_float, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p]),
    "cuorb_search_by_projection_rig": (C.c_int, [C.c_int, C.POINTER(FrameView), C.POINTER(FrameView), C.c_void_p, C.c_void_p, C.c_int] +
                                       [C.c_void_p] * 11 + [C.c_float, C.c_int, C.c_float, C.c_float, C.c_void_p, C.c_void_p]),
    "cuorb_search_by_projection_frame_rig": (C.c_int, [C.c_int, C.POINTER(FrameView), C.POINTER(FrameView), C.c_int] + [C.c_void_p] * 9 +
                                             [C.c_float, C.c_int, C.c_int, C.c_int] + [C.c_void_p] * 5),
    "cuorb_search_by_bow_rig": (C.c_int, [C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_void_p,
                                          C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_float, C.c_int, C.c_void_p,
                                          C.c_void_p]),
    "cuorb_localmap_create": (C.c_void_p, [C.c_int, C.c_int]),
    "cuorb_localmap_destroy": (None, [C.c_void_p]),
    "cuorb_localmap_update": (C.c_int, [C.c_void_p, C.c_int, C.c_int, C.c_int] + [C.c_void_p] * 6),
    "cuorb_localmap_count": (C.c_int, [C.c_void_p]),
    "cuorb_localmap_search": (C.c_int, [C.c_void_p, C.c_void_p, C.POINTER(FrameView), C.c_int, C.c_void_p, C.c_int, C.c_void_p,
                                        C.c_float, C.c_float, C.c_int, C.c_float, C.c_float, C.c_void_p, C.c_void_p, C.c_int,
                                        C.c_void_p, C.c_void_p, C.c_void_p]),
    "cuorb_localmap_search_resident": (C.c_int, [C.c_void_p, C.c_void_p, C.POINTER(FrameView), C.c_void_p, C.c_void_p, C.c_void_p,
                                                 C.c_void_p, C.c_int, C.c_void_p, C.c_int, C.c_void_p, C.c_float, C.c_float,
                                                 C.c_int, C.c_float, C.c_float, C.c_void_p, C.c_void_p, C.c_int, C.c_void_p,
                                                 C.c_void_p, C.c_void_p]),
    "cuorb_preproc_create": (C.c_void_p, [C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p,
                                          C.c_void_p, C.c_int]),
    "cuorb_preproc_destroy": (None, [C.c_void_p]),
    "cuorb_preprocess": (C.c_int, [C.c_void_p, C.c_int, C.c_void_p, C.c_size_t, C.c_size_t, C.c_int, C.c_int,
                                   C.c_void_p, C.c_int]),
    "cuorb_extract_batch_async_pre": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_size_t,
                                                C.c_size_t, C.c_int, C.c_int, C.c_int, C.c_int]),
    "cuorb_undistort_keypoints": (C.c_int, [C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]),
    "cuorb_stereo_from_rgbd": (C.c_int, [C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int,
                                         C.c_size_t, C.c_int, C.c_float, C.c_void_p, C.c_void_p]),
    "cuorb_is_in_frustum": (C.c_int, [C.c_int, C.c_void_p, C.c_int] + [C.c_void_p] * 4 + [C.c_float] + [C.c_void_p] * 7),
    "cuorb_is_in_frustum_noalloc": (C.c_int, [C.c_int, C.c_void_p, C.c_int] + [C.c_void_p] * 4 + [C.c_float] +
                                    [C.c_void_p] * 7),
    "cuorb_knn_match2": (C.c_int, [C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p,
                                   C.c_void_p, C.c_void_p]),
    "cuorb_distinctive_descriptors": (C.c_int, [C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p]),
    "cuorb_search_by_bow": (C.c_int, [C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_void_p,
                                      C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p,
                                      C.c_int, C.c_float, C.c_int, C.c_void_p]),
    "cuorb_search_by_projection_reloc": (C.c_int, [C.c_int, C.POINTER(FrameView), C.c_int] + [C.c_void_p] * 6 +
                                         [C.c_float, C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p]),
    "cuorb_search_by_projection_sim3": (C.c_int, [C.c_int, C.POINTER(FrameView), C.c_int] + [C.c_void_p] * 5 +
                                        [C.c_int, C.c_float, C.c_void_p]),
    "cuorb_fuse": (C.c_int, [C.c_int, C.POINTER(FrameView), C.c_void_p, C.c_int] + [C.c_void_p] * 6 +
                   [C.c_float, C.c_int, C.c_void_p]),
    "cuorb_search_by_sim3": (C.c_int, [C.c_int, C.POINTER(FrameView), C.POINTER(FrameView)] + [C.c_void_p] * 10 +
                             [C.c_float, C.c_void_p]),
    "cuorb_search_for_initialization": (C.c_int, [C.c_int, C.POINTER(FrameView), C.POINTER(FrameView), C.c_void_p,
                                                  C.c_int, C.c_float, C.c_int, C.c_void_p]),
    "cuorb_search_for_triangulation": (C.c_int, [C.c_int, C.POINTER(FrameView), C.POINTER(FrameView), C.c_int] +
                                       [C.c_void_p] * 7 + [C.c_int, C.c_int, C.c_int, C.c_void_p]),
    "cuorb_synth_stereo": (C.c_int, [C.c_int, C.c_uint64, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p,
                                    C.c_void_p]),
    "cuorb_kernel_launch_count": (C.c_uint64, []),
    "cuorb_set_profiling": (C.c_int, [C.c_void_p, C.c_int]),
    "cuorb_stage_ms": (C.c_int, [C.c_void_p, C.c_void_p]),
}

_lib = None


def load() -> C.CDLL:
    """dlopen libcuorb.so; raises if it has not been built (no fallback of any kind)."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise ImportError(
                f"{LIB_PATH} not found: build it with `python -m orbslam3_cpp_fork_b200.build` "
                "(libcuorb is the only implementation; there is no CPU fallback)")
        L = C.CDLL(LIB_PATH)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(L, name)  # AttributeError if the library lacks a declared symbol
            fn.restype = res
            fn.argtypes = args
        _lib = L
    return _lib


def last_error() -> str:
    return load().cuorb_last_error().decode(errors="replace")


def check(rc: int) -> int:
    if rc < 0:
        raise CuorbError(rc, last_error())
    return rc


def ptr(a):
    """numpy array / int address / None -> c_void_p"""
    if a is None:
        return None
    if isinstance(a, np.ndarray):
        return a.ctypes.data_as(C.c_void_p)
    return C.c_void_p(int(a))
