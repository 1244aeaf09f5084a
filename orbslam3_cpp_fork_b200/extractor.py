"""Host-side mirror of class ORB_SLAM3::ORBextractor (reference include/ORBextractor.h:41-119)
over the libcuorb C ABI.  Same constructor arguments, same operator() semantics (returns monoIndex,
-1 for an empty image, mask ignored, keypoints replaced, descriptors n x 32 u8 or empty), same
getters, and `mvImagePyramid` access.  All compute happens in libcuorb's CUDA kernels.
"""
from __future__ import annotations

import ctypes as C

import numpy as np

from . import _lib
from ._lib import KP_DTYPE, CuorbError, check, ptr


def _as_device_ptr(x):
    """torch CUDA tensor -> (ptr, shape, strides-in-bytes) without importing torch at module load."""
    if hasattr(x, "data_ptr") and hasattr(x, "is_cuda"):
        if not x.is_cuda:
            raise ValueError("torch tensor must be a CUDA tensor (or pass a numpy array for host memory)")
        es = x.element_size()
        return x.data_ptr(), tuple(x.shape), tuple(s * es for s in x.stride())
    return None


class ORBextractor:
    """ORBextractor(nfeatures, scaleFactor, nlevels, iniThFAST, minThFAST) — include/ORBextractor.h:48.

    Extra keyword arguments size the device arena: max_width/max_height (largest image),
    max_batch (images per batched call), device (CUDA ordinal).
    """

    def __init__(self, nfeatures=1000, scaleFactor=1.2, nlevels=8, iniThFAST=20, minThFAST=7, *,
                 max_width=752, max_height=480, max_batch=1, device=0):
        self._L = _lib.load()
        self.nfeatures, self.scaleFactor, self.nlevels = int(nfeatures), float(scaleFactor), int(nlevels)
        self.iniThFAST, self.minThFAST = int(iniThFAST), int(minThFAST)
        self.max_batch, self.device = int(max_batch), int(device)
        cfg = _lib.Config(self.nfeatures, self.scaleFactor, self.nlevels, self.iniThFAST, self.minThFAST,
                          int(max_width), int(max_height), self.max_batch, self.device)
        self._h = self._L.cuorb_extractor_create(C.byref(cfg))
        if not self._h:
            raise CuorbError(_lib.E_CUDA, _lib.last_error())
        n = self.nlevels
        self._sf, self._isf, self._s2, self._is2 = (np.zeros(n, np.float32) for _ in range(4))
        self.mnFeaturesPerLevel = np.zeros(n, np.int32)
        check(self._L.cuorb_extractor_get_tables(self._h, ptr(self._sf), ptr(self._isf), ptr(self._s2),
                                                 ptr(self._is2), ptr(self.mnFeaturesPerLevel)))
        self._n_last = 0
        self._keep = None  # keeps the last input array alive while the async copy runs

    def close(self):
        if getattr(self, "_h", None):
            self._L.cuorb_extractor_destroy(self._h)
            self._h = None

    __del__ = close

    @property
    def handle(self):
        return self._h

    # ---- getters (include/ORBextractor.h:63-85) ----
    def GetLevels(self): return self.nlevels
    def GetScaleFactor(self): return self.scaleFactor
    def GetScaleFactors(self): return self._sf.copy()
    def GetInverseScaleFactors(self): return self._isf.copy()
    def GetScaleSigmaSquares(self): return self._s2.copy()
    def GetInverseScaleSigmaSquares(self): return self._is2.copy()

    @property
    def capacity(self) -> int:
        return check(self._L.cuorb_extractor_capacity(self._h))

    # ---- operator() (include/ORBextractor.h:55-61, src/ORBextractor.cc:1052-1138) ----
    def __call__(self, image, mask=None, vLappingArea=(0, 0)):
        """Returns (monoIndex, keypoints[KP_DTYPE], descriptors[n, 32] u8).  monoIndex == -1 and
        empty outputs for an empty image (the reference returns -1 before touching its outputs)."""
        if image is None or getattr(image, "size", 0) == 0:
            return -1, np.zeros(0, KP_DTYPE), np.zeros((0, 32), np.uint8)
        image = np.asarray(image)
        if image.dtype != np.uint8 or image.ndim != 2:
            raise ValueError("image must be 8-bit single channel (CV_8UC1), as the reference asserts")
        if image.strides[1] != 1:
            image = np.ascontiguousarray(image)
        h, w = image.shape
        cap = self.capacity
        kps = np.zeros(cap, KP_DTYPE)
        desc = np.zeros((cap, 32), np.uint8)
        n = C.c_int(0)
        mono = check(self._L.cuorb_extract(self._h, ptr(image), w, h, image.strides[0], int(vLappingArea[0]),
                                           int(vLappingArea[1]), ptr(kps), ptr(desc), cap, C.byref(n)))
        self._n_last = 1
        return mono, kps[: n.value].copy(), desc[: n.value].copy()

    # ---- batched / asynchronous form ----
    def extract_batch_async(self, images, vLappingArea=(0, 0)):
        """images: numpy u8 [n, h, w] in host memory (pinned for true asynchrony) or a torch CUDA
        uint8 tensor [n, h, w].  Enqueues the whole pipeline on the handle's stream."""
        dev = _as_device_ptr(images)
        if dev is not None:
            p, shape, strides = dev
            where = _lib.MEM_DEVICE
        else:
            images = np.asarray(images)
            if images.dtype != np.uint8:
                raise ValueError("images must be uint8")
            p, shape, strides, where = images.ctypes.data, images.shape, images.strides, _lib.MEM_HOST
        if len(shape) != 3 or strides[2] != 1:
            raise ValueError("images must be [n, h, w] with unit pixel stride")
        n, h, w = shape
        self._keep = images
        check(self._L.cuorb_extract_batch_async(self._h, n, C.c_void_p(p), w, h, strides[1], strides[0], where,
                                                int(vLappingArea[0]), int(vLappingArea[1])))
        self._n_last = n

    def fetch(self, kps=None, desc=None, counts=None, mono=None, sync=True):
        """Copy the last batch's results to host.  Returns (counts, monoIndex, kps[n, cap], desc[n, cap, 32])."""
        n, cap = self._n_last, self.capacity
        kps = np.zeros((n, cap), KP_DTYPE) if kps is None else kps
        desc = np.zeros((n, cap, 32), np.uint8) if desc is None else desc
        counts = np.zeros(n, np.int32) if counts is None else counts
        mono = np.zeros(n, np.int32) if mono is None else mono
        fn = self._L.cuorb_extract_fetch if sync else self._L.cuorb_extract_fetch_async
        check(fn(self._h, ptr(kps), ptr(desc), kps.shape[1], ptr(counts), ptr(mono)))
        return counts, mono, kps, desc

    @property
    def stream(self) -> int:
        """cudaStream_t of the handle (integer address), e.g. for torch.cuda.ExternalStream."""
        return int(self._L.cuorb_extractor_stream(self._h) or 0)

    def sync(self):
        check(self._L.cuorb_sync(self._h))

    # ---- mvImagePyramid (include/ORBextractor.h:87) ----
    def pyramid_device(self, level, image_index=0):
        """(device pointer to pixel (0,0), width, height, stride) of a level of the last batch."""
        p, w, h, s = C.c_void_p(), C.c_int(), C.c_int(), C.c_size_t()
        check(self._L.cuorb_extractor_pyramid(self._h, image_index, level, C.byref(p), C.byref(w), C.byref(h), C.byref(s)))
        return p.value, w.value, h.value, s.value

    def download_plane(self, level, plane=1, image_index=0):
        """plane 0: level with its 19-px border, 1: level, 2: blurred level (numpy u8 2-D)."""
        _, w, h, _ = self.pyramid_device(level, image_index)
        shape = (h + 38, w + 38) if plane == 0 else (h, w)
        out = np.zeros(shape, np.uint8)
        r = self._L.cuorb_extractor_download_plane(self._h, image_index, level, plane, ptr(out), out.size, None, None)
        check(int(r))
        return out

    @property
    def mvImagePyramid(self):
        return [self.download_plane(l, 1, 0) for l in range(self.nlevels)]

    def download_candidates(self, level, image_index=0, cap=400000):
        xs, ys, rs = (np.zeros(cap, np.int32) for _ in range(3))
        n = check(self._L.cuorb_extractor_download_candidates(self._h, image_index, level, ptr(xs), ptr(ys), ptr(rs), cap))
        assert n <= cap
        return xs[:n].copy(), ys[:n].copy(), rs[:n].copy()

    # ---- instrumentation ----
    def set_profiling(self, on=True):
        check(self._L.cuorb_set_profiling(self._h, int(on)))

    def stage_ms(self):
        out = np.zeros(len(_lib.STAGES), np.float32)
        check(self._L.cuorb_stage_ms(self._h, ptr(out)))
        return dict(zip(_lib.STAGES, out.tolist()))
