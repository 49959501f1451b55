"""ctypes loader for libydorb.so.  There is no fallback: if the CUDA library is missing, importing fails."""
import ctypes as C
import pathlib

import numpy as np

HERE = pathlib.Path(__file__).resolve().parent
LIB_PATH = HERE / "libydorb.so"
MAX_LEVELS = 16

KP_DTYPE = np.dtype(
    [("x", "<f4"), ("y", "<f4"), ("size", "<f4"), ("angle", "<f4"), ("response", "<f4"), ("octave", "<i4"), ("class_id", "<i4")]
)
TOP2_DTYPE = np.dtype([("best_idx", "<i4"), ("best_dist", "<i4"), ("second_dist", "<i4"), ("second_idx", "<i4")])
MPPROJ_DTYPE = np.dtype(
    [("proj_x", "<f4"), ("proj_y", "<f4"), ("proj_right_x", "<f4"), ("view_cos", "<f4"), ("level", "<i4"), ("active", "<i4")]
)
WINDOW_QUERY_DTYPE = np.dtype(
    [("x", "<f4"), ("y", "<f4"), ("r", "<f4"), ("min_level", "<i4"), ("max_level", "<i4"), ("oct_lo", "<i4"), ("oct_hi", "<i4"),
     ("right_x", "<f4"), ("right_r", "<f4"), ("angle", "<f4"), ("flags", "<i4")]
)
Q_ACTIVE, Q_BLOCKS = 1, 2


class WindowParams(C.Structure):
    _fields_ = [
        ("max_dist", C.c_int32),
        ("strict", C.c_int32),
        ("exclusive", C.c_int32),
        ("check_orientation", C.c_int32),
        ("chi2_gate", C.c_int32),
        ("inv_sf2", C.c_float * MAX_LEVELS),
    ]


class FeatureVector(C.Structure):
    """DBoW3::FeatureVector in CSR form (include/ydorb.h YdFeatureVector)."""

    _fields_ = [("node_ids", C.c_void_p), ("node_start", C.c_void_p), ("items", C.c_void_p), ("n_nodes", C.c_int32)]

    @classmethod
    def from_arrays(cls, node_ids, node_start, items):
        fv = cls()
        fv._keep = (np.ascontiguousarray(node_ids, np.int32), np.ascontiguousarray(node_start, np.int32),
                    np.ascontiguousarray(items, np.int32))
        fv.node_ids, fv.node_start, fv.items = (a.ctypes.data for a in fv._keep)
        fv.n_nodes = len(fv._keep[0])
        return fv


class ExtractorParams(C.Structure):
    _fields_ = [
        ("n_features", C.c_int32),
        ("scale_factor", C.c_float),
        ("n_levels", C.c_int32),
        ("ini_th_fast", C.c_int32),
        ("min_th_fast", C.c_int32),
        ("reference_min_th_quirk", C.c_int32),
    ]


class FrameGeom(C.Structure):
    _fields_ = [
        ("min_x", C.c_float),
        ("min_y", C.c_float),
        ("max_x", C.c_float),
        ("max_y", C.c_float),
        ("n_levels", C.c_int32),
        ("scale_factors", C.c_float * MAX_LEVELS),
    ]


class YdorbError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__(f"libydorb error {code}: {msg}")
        self.code = code


_lib = None


def lib():
    global _lib
    if _lib is None:
        if not LIB_PATH.exists():
            raise ImportError(
                f"{LIB_PATH} is missing: build it with `python -m ydorbslam_b200.build` (nvcc, sm_100a). "
                "There is no CPU fallback."
            )
        _lib = C.CDLL(str(LIB_PATH))
        _lib.ydorb_last_error.restype = C.c_char_p
        _lib.ydorb_extractor_stream.restype = C.c_void_p
        _lib.ydorb_matcher_stream.restype = C.c_void_p
        _lib.ydorb_kernel_launches.restype = C.c_longlong
        for name in ("ydorb_extractor_destroy", "ydorb_matcher_destroy", "ydorb_frame_destroy", "ydorb_three_maxima",
                     "ydorb_vocabulary_destroy", "ydorb_host_free"):
            getattr(_lib, name).restype = None
        _lib.ydorb_host_free.argtypes = [C.c_void_p]
    return _lib


def check(rc):
    if rc != 0:
        raise YdorbError(rc, lib().ydorb_last_error().decode())


def ptr(a):
    if a is None:
        return None
    if isinstance(a, np.ndarray):
        return a.ctypes.data_as(C.c_void_p)
    return C.c_void_p(int(a))  # raw device pointer


class PinnedArray:
    """numpy view of page-locked host memory from ydorb_host_alloc (include/ydorb.h); frees it on close()/GC."""

    def __init__(self, shape, dtype=np.uint8):
        self._p = C.c_void_p()
        dtype = np.dtype(dtype)
        nbytes = int(np.prod(shape)) * dtype.itemsize
        check(lib().ydorb_host_alloc(C.c_size_t(nbytes), C.byref(self._p)))
        buf = (C.c_uint8 * max(nbytes, 1)).from_address(self._p.value)
        self.array = np.frombuffer(buf, dtype=dtype, count=int(np.prod(shape))).reshape(shape)

    def close(self):
        if self._p is not None and self._p.value:
            self.array = None
            lib().ydorb_host_free(self._p)
            self._p = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass
