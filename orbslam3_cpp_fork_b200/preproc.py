"""Host-side mirror of the frame preparation the reference runs before ORBextractor::operator()
(SURVEY.md §8f row N1): System::TrackStereo/TrackRGBD/TrackMonocular's cv::remap / cv::resize
(src/System.cc:280-300, 384-386, 476) followed by Tracking::GrabImage*'s cv::cvtColor to gray
(src/Tracking.cc:1339-1357, 1449-1462, 1520-1533).  All compute is libcuorb's k_preproc kernel;
without the CUDA library construction raises.
"""
from __future__ import annotations

import ctypes as C

import numpy as np

from . import _lib
from ._lib import check, ptr
from .extractor import _as_device_ptr

PIX_GRAY, PIX_RGB, PIX_BGR, PIX_RGBA, PIX_BGRA = range(5)
PRE_NONE, PRE_REMAP, PRE_RESIZE = range(3)
_CHANNELS = {PIX_GRAY: 1, PIX_RGB: 3, PIX_BGR: 3, PIX_RGBA: 4, PIX_BGRA: 4}


class FramePreprocessor:
    """mode PRE_REMAP: maps = (M1, M2) float32 [dst_h, dst_w] from cv::initUndistortRectifyMap
    (settings->M1l()/M2l()); mode PRE_RESIZE: dsize = settings->newImSize() as (w, h)."""

    def __init__(self, src_size, mode=PRE_NONE, maps=None, dsize=None, max_batch=1, device=0):
        self._L = _lib.load()
        self.src_w, self.src_h = int(src_size[0]), int(src_size[1])
        self.mode, self.max_batch, self.device = int(mode), int(max_batch), int(device)
        mx = my = None
        if self.mode == PRE_REMAP:
            mx = np.ascontiguousarray(maps[0], dtype=np.float32)
            my = np.ascontiguousarray(maps[1], dtype=np.float32)
            if mx.shape != my.shape or mx.ndim != 2:
                raise ValueError("maps must be two float32 arrays [dst_h, dst_w]")
            self.dst_h, self.dst_w = mx.shape
        elif self.mode == PRE_RESIZE:
            self.dst_w, self.dst_h = int(dsize[0]), int(dsize[1])
        else:
            self.dst_w, self.dst_h = self.src_w, self.src_h
        self._h = self._L.cuorb_preproc_create(self.device, self.mode, self.src_w, self.src_h, self.dst_w,
                                               self.dst_h, ptr(mx), ptr(my), self.max_batch)
        if not self._h:
            raise _lib.CuorbError(-4, _lib.last_error())

    def close(self):
        if getattr(self, "_h", None):
            self._L.cuorb_preproc_destroy(self._h)
            self._h = None

    __del__ = close

    @property
    def handle(self):
        return self._h

    def _describe(self, frames, fmt):
        ch = _CHANNELS[int(fmt)]
        dev = _as_device_ptr(frames)
        if dev is not None:
            p, shape, strides = dev
            where = _lib.MEM_DEVICE
        else:
            frames = np.asarray(frames)
            if frames.dtype != np.uint8:
                raise ValueError("frames must be uint8")
            p, shape, strides, where = frames.ctypes.data, frames.shape, frames.strides, _lib.MEM_HOST
        want = 3 if ch == 1 else 4
        if len(shape) != want or (ch > 1 and shape[3] != ch):
            raise ValueError(f"frames must be [n, h, w{', ' + str(ch) if ch > 1 else ''}]")
        if shape[1] != self.src_h or shape[2] != self.src_w or strides[2] != ch or (ch > 1 and strides[3] != 1):
            raise ValueError("frames must be src_h x src_w with interleaved channels")
        return frames, p, shape[0], strides[1], strides[0], where

    def __call__(self, frames, fmt=PIX_GRAY):
        """frames [n, h, w(, c)] (numpy host or torch CUDA) -> numpy gray [n, dst_h, dst_w]."""
        keep, p, n, stride, istride, where = self._describe(frames, fmt)
        out = np.empty((n, self.dst_h, self.dst_w), np.uint8)
        check(self._L.cuorb_preprocess(self._h, n, C.c_void_p(p), stride, istride, where, int(fmt), ptr(out),
                                       _lib.MEM_HOST))
        return out

    def extract_batch_async(self, extractor, frames, fmt=PIX_GRAY, vLappingArea=(0, 0)):
        """preprocess + ORBextractor::operator() on the extractor's stream; fetch from the extractor."""
        keep, p, n, stride, istride, where = self._describe(frames, fmt)
        extractor._keep = keep
        check(self._L.cuorb_extract_batch_async_pre(extractor.handle, self._h, n, C.c_void_p(p), stride, istride,
                                                    where, int(fmt), int(vLappingArea[0]), int(vLappingArea[1])))
        extractor._n_last = n
