"""Host mirror of the pre-/post-processing the reference runs around extractAndCompute, over the
C ABI of libydorb.so: cv::cvtColor to gray (src/tracking.cpp:73-89, :97-107), Frame::undistortKeyPoints /
computeImageBounds (src/frame.cpp:193-213, :224-248) and Frame::computeStereoFromRGBD fused with the depth-map
convertTo (src/frame.cpp:214-223, src/tracking.cpp:109-111).  A matcher handle supplies the device and the stream."""
import ctypes as C

import numpy as np

from ._lib import KP_DTYPE, check, lib, ptr

RGB2GRAY, BGR2GRAY, RGBA2GRAY, BGRA2GRAY = range(4)


def cvtColor(matcher, image, code):
    """image: (h, w, 3|4) uint8 -> (h, w) uint8, bit-identical to cv2.cvtColor(image, COLOR_*2GRAY)."""
    a = np.ascontiguousarray(image, np.uint8)
    h, w, ch = a.shape
    if ch != (4 if code >= RGBA2GRAY else 3):
        raise ValueError("channel count does not match the conversion code")
    out = np.zeros((h, w), np.uint8)
    check(lib().ydorb_cvt_gray(matcher._h, ptr(a), w, h, w * ch, int(code), ptr(out), w))
    return out


def undistortPoints(matcher, xy, K, dist, P):
    """cv::undistortPoints(xy, K, dist, R = I, P): (n, 2) float32 in, (n, 2) float32 out."""
    pts = np.ascontiguousarray(xy, np.float32).reshape(-1, 2)
    Kf, Pf, df = (np.ascontiguousarray(v, np.float32).ravel() for v in (K, P, dist))
    out = np.zeros_like(pts)
    check(lib().ydorb_undistort_points(matcher._h, ptr(pts), len(pts), ptr(Kf), ptr(df), len(df), ptr(Pf), ptr(out)))
    return out


def undistortKeyPoints(matcher, keypoints, K, dist, P):
    """Frame::undistortKeyPoints: returns a copy of `keypoints` with pt.x / pt.y undistorted; unchanged when dist[0] == 0
    (the reference's guard, src/frame.cpp:194)."""
    k = np.array(keypoints, dtype=KP_DTYPE, copy=True)
    d = np.ascontiguousarray(dist, np.float32).ravel()
    if len(k) and d[0] != 0.0:
        xy = undistortPoints(matcher, np.stack([k["x"], k["y"]], 1), K, d, P)
        k["x"], k["y"] = xy[:, 0], xy[:, 1]
    return k


def computeImageBounds(matcher, width, height, K, dist):
    """Frame::computeImageBounds (src/frame.cpp:224-248): (minX, maxX, minY, maxY)."""
    d = np.ascontiguousarray(dist, np.float32).ravel()
    if d[0] == 0.0:
        return 0.0, float(width), 0.0, float(height)
    c = undistortPoints(matcher, np.array([[0, 0], [width, 0], [0, height], [width, height]], np.float32), K, d, K)
    return (float(min(c[0, 0], c[2, 0])), float(max(c[1, 0], c[3, 0])), float(min(c[0, 1], c[1, 1])), float(max(c[2, 1], c[3, 1])))


def computeStereoFromRGBD(matcher, keypoints, depth_map, depth_map_factor, bf):
    """keypoints: undistorted key points; depth_map: (h, w) uint16 or float32 as read from disk (NOT yet scaled);
    returns (right_x, depth), -1 where the map has no depth."""
    k = np.ascontiguousarray(keypoints, KP_DTYPE)
    d = np.ascontiguousarray(depth_map)
    if d.dtype not in (np.uint16, np.float32):
        raise ValueError("depth map must be uint16 or float32")
    rx, dz = np.zeros(len(k), np.float32), np.zeros(len(k), np.float32)
    check(lib().ydorb_stereo_from_rgbd(matcher._h, ptr(k), len(k), ptr(d), 0 if d.dtype == np.uint16 else 1, d.shape[1], d.shape[0],
                                       d.strides[0], C.c_float(depth_map_factor), C.c_float(bf), ptr(rx), ptr(dz)))
    return rx, dz
