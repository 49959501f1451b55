"""Host-side mirror of YDORBSLAM::OrbExtractor (src/orbExtractor.hpp:31-73) over the C ABI.

Same constructor arguments, same outputs (keypoints level-major as cv::KeyPoint records, N x 32 uint8
descriptor matrix), same scale getters, plus access to the padded pyramid (the reference's public
m_v_imagePyramid).  All compute happens in libydorb.so on the GPU.
"""
import ctypes as C

import numpy as np

from . import _lib
from ._lib import KP_DTYPE, check, lib, ptr


class OrbExtractor:
    def __init__(self, n_features, scale_factor, n_levels, ini_th_fast, min_th_fast, *, max_width=1920, max_height=1080,
                 max_batch=1, device=0, reference_min_th_quirk=True):
        """Arguments as src/orbExtractor.hpp:33-34.  reference_min_th_quirk=True reproduces the reference ctor,
        which initialises m_int_minFastThd with iniThFAST (src/orbExtractor.cpp:318)."""
        self.params = _lib.ExtractorParams(n_features, scale_factor, n_levels, ini_th_fast, min_th_fast,
                                           1 if reference_min_th_quirk else 0)
        self._h = C.c_void_p()
        check(lib().ydorb_extractor_create(C.byref(self.params), max_width, max_height, max_batch, device, C.byref(self._h)))
        self.n_levels = n_levels
        self.max_batch = max_batch
        self.device = device
        q = np.zeros(n_levels, np.int32)
        check(lib().ydorb_extractor_quotas(self._h, ptr(q)))
        self.quotas = q
        self.capacity = max(int(q.sum()), 1)
        self._last_shape = None

    def close(self):
        if self._h:
            lib().ydorb_extractor_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # ---- getters (src/orbExtractor.hpp:42-48)
    def getLevelsNum(self):
        return self.n_levels

    def getScaleFactor(self):
        return float(self.params.scale_factor)

    def _tables(self):
        t = [np.zeros(self.n_levels, np.float32) for _ in range(4)]
        check(lib().ydorb_extractor_scale_tables(self._h, *[ptr(a) for a in t]))
        return t

    def getScaleFactors(self):
        return self._tables()[0]

    def getInvScaleFactors(self):
        return self._tables()[1]

    def getScaleFactorSquares(self):
        return self._tables()[2]

    def getInvScaleFactorSquares(self):
        return self._tables()[3]

    @property
    def stream(self):
        return lib().ydorb_extractor_stream(self._h)

    def sync(self):
        check(lib().ydorb_extractor_sync(self._h))

    def set_stream(self, cuda_stream):
        """Enqueue on a caller-owned cudaStream_t (int); None restores the handle's own stream."""
        check(lib().ydorb_extractor_set_stream(self._h, C.c_void_p(cuda_stream) if cuda_stream else None))

    def set_profiling(self, enable=True):
        check(lib().ydorb_extractor_set_profiling(self._h, 1 if enable else 0))

    def stage_ms(self):
        """CUDA-event durations of the last call: [pyramid, fast, quadtree, describe] in ms."""
        ms = (C.c_float * 4)()
        check(lib().ydorb_extractor_stage_ms(self._h, ms))
        return [float(v) for v in ms]

    # ---- extractAndCompute (src/orbExtractor.cpp:355-399)
    def extractAndCompute(self, image):
        """image: HxW uint8.  Returns (keypoints KP_DTYPE[N], descriptors uint8[N,32])."""
        if image is None or image.size == 0:
            return np.zeros(0, KP_DTYPE), np.zeros((0, 32), np.uint8)
        assert image.dtype == np.uint8 and image.ndim == 2, "CV_8UC1 expected (src/orbExtractor.cpp:361)"
        k, d, n = self.extract_batch(image[None])
        return k[0, : n[0]].copy(), d[0, : n[0]].copy()

    def extract_batch(self, images):
        """images: BxHxW uint8 (host).  Returns (kps[B,cap], desc[B,cap,32], n[B])."""
        images = np.ascontiguousarray(images, dtype=np.uint8)
        b, h, w = images.shape
        cap = self.capacity
        kps = np.zeros((b, cap), KP_DTYPE)
        desc = np.zeros((b, cap, 32), np.uint8)
        n = np.zeros(b, np.int32)
        check(lib().ydorb_extract_batch(self._h, ptr(images), w, h, w, C.c_size_t(h * w), b, ptr(kps), ptr(desc), cap, ptr(n)))
        self._last_shape = (b, h, w)
        return kps, desc, n

    def extract_batch_device(self, d_images, w, h, batch, d_kps, d_desc, capacity, d_nout, stride=None, frame_stride=None):
        """All pointers are raw device addresses (ints); asynchronous on self.stream."""
        stride = stride or w
        frame_stride = frame_stride or stride * h
        check(lib().ydorb_extract_batch_device(self._h, ptr(d_images), w, h, stride, C.c_size_t(frame_stride), batch,
                                               ptr(d_kps), ptr(d_desc), capacity, ptr(d_nout)))
        self._last_shape = (batch, h, w)

    # ---- pyramid access (public m_v_imagePyramid, src/orbExtractor.hpp:50)
    def pyramid_level_device(self, frame, level):
        p, w, h, st = C.c_void_p(), C.c_int(), C.c_int(), C.c_int()
        check(lib().ydorb_pyramid_level(self._h, frame, level, C.byref(p), C.byref(w), C.byref(h), C.byref(st)))
        return p.value, w.value, h.value, st.value

    def pyramid_level(self, frame, level, padded=False):
        _, w, h, _ = self.pyramid_level_device(frame, level)
        if padded:
            w, h = w + 38, h + 38
        out = np.zeros((h, w), np.uint8)
        check(lib().ydorb_pyramid_download(self._h, frame, level, 1 if padded else 0, ptr(out), w))
        return out

    def debug_candidates(self, frame, level):
        n = C.c_int32()
        check(lib().ydorb_debug_candidates(self._h, frame, level, None, None, 0, C.byref(n)))
        xy = np.zeros((max(n.value, 1), 2), np.int32)
        resp = np.zeros(max(n.value, 1), np.float32)
        check(lib().ydorb_debug_candidates(self._h, frame, level, ptr(xy), ptr(resp), n.value, C.byref(n)))
        return xy[: n.value], resp[: n.value]
