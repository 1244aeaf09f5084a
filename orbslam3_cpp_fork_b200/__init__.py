"""orbslam3_cpp_fork_b200 — B200-native ORB front-end (libcuorb) behind the reference's operator API.

Only the hot path of SURVEY.md §8 lives here: csrc/ (sm_100a CUDA kernels + the C ABI of
include/cuorb.h), the ctypes binding, and the host-side mirrors of ORBextractor / ORBmatcher.
"""
from ._lib import KP_DTYPE, CuorbError, load  # noqa: F401
from .extractor import ORBextractor  # noqa: F401
from .matcher import Frame, MapPointSoA, ORBmatcher, RigFrame  # noqa: F401
from .preproc import FramePreprocessor  # noqa: F401
from .frame_ops import ComputeStereoFromRGBD, LocalMap, UndistortKeyPoints, isInFrustum  # noqa: F401
from .bfmatch import ComputeDistinctiveDescriptors, knnMatch2  # noqa: F401
