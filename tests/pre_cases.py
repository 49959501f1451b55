"""Shared inputs of the pre-processing tests."""
import numpy as np

from ydorbslam_b200 import synth
from ydorbslam_b200._lib import KP_DTYPE

# TUM1-like intrinsics with 5 distortion coefficients, a 4-coefficient set, a strong barrel set that drives far points
# into the icdist < 0 branch, and 8 / 12 coefficient (rational / thin-prism) sets
K = np.array([[517.306408, 0, 318.643040], [0, 516.469215, 255.313989], [0, 0, 1]], np.float32)
P_OTHER = np.array([[500.0, 0.5, 320.0], [0, 505.0, 240.0], [0, 0, 1]], np.float32)
DISTS = {
    "tum5": np.array([0.262383, -0.953104, -0.005358, 0.002628, 1.163314], np.float32),
    "four": np.array([0.262383, -0.953104, -0.005358, 0.002628], np.float32),
    "barrel": np.array([-0.4, 0.1, 0.001, -0.002, -0.3], np.float32),
    "rational8": np.array([0.1, -0.2, 0.001, 0.002, 0.05, 0.02, -0.01, 0.003], np.float32),
    "prism12": np.array([0.1, -0.2, 0.001, 0.002, 0.05, 0.02, -0.01, 0.003, 1e-3, -2e-3, 5e-4, 1e-3], np.float32),
}


def color_image(w, h, ch, seed):
    return synth.rand_bytes(seed, 7, w * h * ch).reshape(h, w, ch)


def points(n, seed, spread=1.0):
    u = synth.rand_u64(seed, 8, 2 * n).astype(np.float64) / 2.0**64
    xy = (u.reshape(n, 2) - 0.5) * spread + 0.5
    return (xy * [640, 480]).astype(np.float32)


def depth_map_u16(w, h, seed):
    d = (synth.rand_u64(seed, 9, w * h) % np.uint64(40000)).astype(np.uint16).reshape(h, w)
    d[(synth.rand_u64(seed, 10, w * h) % np.uint64(5) == 0).reshape(h, w)] = 0  # holes: no depth
    return d


def keypoints_at(xy):
    k = np.zeros(len(xy), KP_DTYPE)
    k["x"], k["y"] = xy[:, 0], xy[:, 1]
    k["octave"] = 0
    k["size"] = 31
    k["class_id"] = -1
    return k
