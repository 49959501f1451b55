"""Host mirror of the map-point side of the front-end (SURVEY 8f "next" rows) over the C ABI of libydorb.so:
MapPoint::computeDistinctiveDescriptors (src/mapPoint.cpp:169-218) and Frame::isInCameraFrustum with
MapPoint::predictScaleLevel (src/frame.cpp:295-326, src/mapPoint.cpp:265-277), batched over map points."""
import ctypes as C

import numpy as np

from ._lib import MAX_LEVELS, MPPROJ_DTYPE, check, lib, ptr

MAPPOINT3D_DTYPE = np.dtype([("pos", "<f4", 3), ("normal", "<f4", 3), ("max_distance", "<f4"), ("min_distance", "<f4")])


class Camera(C.Structure):
    """include/ydorb.h YdCamera: the Frame members isInCameraFrustum reads."""

    _fields_ = [("R", C.c_float * 9), ("t", C.c_float * 3), ("origin", C.c_float * 3), ("fx", C.c_float), ("fy", C.c_float),
                ("cx", C.c_float), ("cy", C.c_float), ("min_x", C.c_float), ("max_x", C.c_float), ("min_y", C.c_float),
                ("max_y", C.c_float), ("bf", C.c_float), ("log_scale_factor", C.c_float), ("n_levels", C.c_int32)]

    @classmethod
    def make(cls, R, t, fx, fy, cx, cy, bounds, bf, scale_factor, n_levels):
        """R, t: camera-from-world rotation / translation (m_cvMat_R_c2w, m_cvMat_t_c2w); origin = -R^T t in float32 as
        Frame::updatePoseMatrices computes it (src/frame.cpp:273-284)."""
        assert n_levels <= MAX_LEVELS
        R = np.asarray(R, np.float32).reshape(3, 3)
        t = np.asarray(t, np.float32).reshape(3)
        c = cls()
        c.R[:] = R.ravel().tolist()
        c.t[:] = t.tolist()
        Rw = R.T.copy()
        origin = np.zeros(3, np.float32)
        for i in range(3):  # -R^T * t with the float rounding of cv::gemm's 3x3 path
            origin[i] = -np.float32(np.float32(np.float32(Rw[i, 0] * t[0]) + np.float32(Rw[i, 1] * t[1])) + np.float32(Rw[i, 2] * t[2]))
        c.origin[:] = origin.tolist()
        c.fx, c.fy, c.cx, c.cy = fx, fy, cx, cy
        c.min_x, c.min_y, c.max_x, c.max_y = bounds
        c.bf = bf
        c.log_scale_factor = float(np.float32(np.log(float(np.float32(scale_factor)))))  # src/frame.cpp:79
        c.n_levels = n_levels
        return c


def computeDistinctiveDescriptors(matcher, descriptors, start):
    """descriptors: (total, 32) uint8, start: (n_points+1,) int32 CSR offsets.  Returns best[n_points] (relative row, -1
    for points without observations)."""
    d = np.ascontiguousarray(descriptors, np.uint8).reshape(-1, 32)
    s = np.ascontiguousarray(start, np.int32)
    best = np.zeros(len(s) - 1, np.int32)
    check(lib().ydorb_distinctive_descriptors(matcher._h, ptr(d), ptr(s), len(best), ptr(best)))
    return best


def isInCameraFrustum(matcher, camera, map_points, view_cos_limit=0.5):
    """map_points: MAPPOINT3D_DTYPE array.  Returns an MPPROJ_DTYPE array (the map points' tracking state)."""
    mp = np.ascontiguousarray(map_points, MAPPOINT3D_DTYPE)
    out = np.zeros(len(mp), MPPROJ_DTYPE)
    check(lib().ydorb_frustum(matcher._h, C.byref(camera), ptr(mp), len(mp), C.c_float(view_cos_limit), ptr(out)))
    return out
