"""Host-side mirror of YDORBSLAM::OrbMatcher (src/orbMatcher.hpp:24-66) and the Frame hooks it relies on
(src/frame.hpp:83,96,166) over the C ABI.  All compute happens in libydorb.so on the GPU.
"""
import ctypes as C

import numpy as np

from . import _lib
from ._lib import KP_DTYPE, MPPROJ_DTYPE, TOP2_DTYPE, check, lib, ptr

HIGH_THD = 100  # OrbMatcher::m_int_highThd  src/orbMatcher.cpp:7
LOW_THD = 50    # OrbMatcher::m_int_lowThd   src/orbMatcher.cpp:8
HIST_LEN = 30   # OrbMatcher::m_int_histLen  src/orbMatcher.cpp:9


class Frame:
    """Feature state of a Frame on the device: keypoints, descriptors, rightX, 64x48 grid
    (Frame::assignKeyPointsToGrid, src/frame.cpp:249-264)."""

    def __init__(self, matcher, keypoints, descriptors, right_x=None, *, bounds, scale_factors):
        kps = np.ascontiguousarray(keypoints, dtype=KP_DTYPE)
        desc = np.ascontiguousarray(descriptors, dtype=np.uint8)
        assert desc.shape == (len(kps), 32)
        g = _lib.FrameGeom()
        g.min_x, g.min_y, g.max_x, g.max_y = (float(v) for v in bounds)
        g.n_levels = len(scale_factors)
        for i, s in enumerate(scale_factors):
            g.scale_factors[i] = float(s)
        rx = None if right_x is None else np.ascontiguousarray(right_x, dtype=np.float32)
        self.n = len(kps)
        self._h = C.c_void_p()
        self._matcher = matcher
        check(lib().ydorb_frame_create(matcher._h, C.byref(g), ptr(kps), ptr(desc), ptr(rx), self.n, C.byref(self._h)))

    def close(self):
        if self._h:
            lib().ydorb_frame_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def getKeyPointsInArea(self, x, y, r, min_level=-1, max_level=-1):
        """Frame::getKeyPointsInArea (src/frame.cpp:337-361): reference predicate and order."""
        out = np.zeros(max(self.n, 1), np.int32)
        n = C.c_int32()
        check(lib().ydorb_frame_keypoints_in_area(self._h, C.c_float(x), C.c_float(y), C.c_float(r), min_level, max_level,
                                                  ptr(out), len(out), C.byref(n)))
        return out[: n.value].copy()


class OrbMatcher:
    def __init__(self, best_second_ratio=0.6, check_orientation=True, *, device=0):
        """Arguments as src/orbMatcher.hpp:26."""
        self.ratio = float(best_second_ratio)
        self.check_orientation = bool(check_orientation)
        self.device = device
        self._h = C.c_void_p()
        check(lib().ydorb_matcher_create(device, C.byref(self._h)))

    def close(self):
        if self._h:
            lib().ydorb_matcher_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    @property
    def stream(self):
        return lib().ydorb_matcher_stream(self._h)

    def sync(self):
        check(lib().ydorb_matcher_sync(self._h))

    # ---- M0 computeDescriptorsDistance (src/orbMatcher.cpp:11-23), batched
    def computeDescriptorsDistance(self, a, b):
        a = np.ascontiguousarray(a, dtype=np.uint8).reshape(-1, 32)
        b = np.ascontiguousarray(b, dtype=np.uint8).reshape(-1, 32)
        assert a.shape == b.shape
        d = np.zeros(len(a), np.int32)
        check(lib().ydorb_hamming_pairs(self._h, ptr(a), ptr(b), len(a), ptr(d)))
        return d if len(d) != 1 else int(d[0])

    # ---- dense top-2 (inner loops of searchByBow*, src/orbMatcher.cpp:319-334)
    def top2(self, queries, db, idx_base=0):
        q = np.ascontiguousarray(queries, dtype=np.uint8).reshape(-1, 32)
        d = np.ascontiguousarray(db, dtype=np.uint8).reshape(-1, 32)
        out = np.zeros(len(q), TOP2_DTYPE)
        check(lib().ydorb_hamming_top2(self._h, ptr(q), len(q), ptr(d), len(d), idx_base, ptr(out)))
        return out

    def top2_device(self, d_q, nq, d_db, ndb, idx_base, d_out):
        check(lib().ydorb_hamming_top2_device(self._h, ptr(d_q), nq, ptr(d_db), ndb, idx_base, ptr(d_out)))

    def set_stream(self, cuda_stream):
        check(lib().ydorb_matcher_set_stream(self._h, C.c_void_p(cuda_stream) if cuda_stream else None))

    def top2_frames_device(self, d_desc, d_n, capacity, d_q_frame, d_db_frame, npairs, d_out, d_accept=None, ratio=None,
                           th=LOW_THD):
        """Batched dense top-2 between frames of one descriptor block (device pointers, asynchronous)."""
        check(lib().ydorb_hamming_top2_frames_device(self._h, ptr(d_desc), ptr(d_n), capacity, ptr(d_q_frame), ptr(d_db_frame),
                                                     npairs, ptr(d_out), ptr(d_accept),
                                                     C.c_float(self.ratio if ratio is None else ratio), th))

    def match_frames(self, extractor, q_frame, db_frame, ratio=None, th=LOW_THD):
        """Host API: top-2 + ratio test between frames of the extractor's last host-API batch (descriptors stay on
        the device).  Returns (top2[npairs, capacity], accept[npairs, capacity])."""
        qf = np.ascontiguousarray(q_frame, dtype=np.int32)
        df = np.ascontiguousarray(db_frame, dtype=np.int32)
        cap = extractor.capacity
        out = np.zeros((len(qf), cap), TOP2_DTYPE)
        acc = np.zeros((len(qf), cap), np.uint8)
        check(lib().ydorb_match_frames(self._h, extractor._h, ptr(qf), ptr(df), len(qf),
                                       C.c_float(self.ratio if ratio is None else ratio), th, ptr(out), ptr(acc)))
        return out, acc

    def top2_merge_device(self, d_shards, nshards, nq, d_out):
        check(lib().ydorb_top2_merge_device(self._h, ptr(d_shards), nshards, nq, ptr(d_out)))

    def ratio_accept_device(self, d_top2, nq, d_accept, ratio=None, th=LOW_THD):
        check(lib().ydorb_ratio_accept_device(self._h, ptr(d_top2), nq, C.c_float(self.ratio if ratio is None else ratio), th,
                                              ptr(d_accept)))

    # ---- Frame hooks
    def make_frame(self, keypoints, descriptors, right_x=None, *, bounds, scale_factors):
        return Frame(self, keypoints, descriptors, right_x, bounds=bounds, scale_factors=scale_factors)

    # ---- M1 searchByProjectionInFrameAndMapPoint (src/orbMatcher.cpp:24-64)
    def searchByProjectionInFrameAndMapPoint(self, frame, map_points, mp_descriptors, mp_obs_positive, thd=3.0,
                                             kp_occupied=None):
        """map_points: MPPROJ_DTYPE[n].  Returns (match count, match_of_keypoint int32[n_kp] with -1 = none)."""
        mps = np.ascontiguousarray(map_points, dtype=MPPROJ_DTYPE)
        md = np.ascontiguousarray(mp_descriptors, dtype=np.uint8).reshape(-1, 32)
        obs = np.ascontiguousarray(mp_obs_positive, dtype=np.uint8)
        occ = None if kp_occupied is None else np.ascontiguousarray(kp_occupied, dtype=np.uint8)
        out = np.full(max(frame.n, 1), -1, np.int32)
        n = C.c_int32()
        check(lib().ydorb_search_by_projection(frame._h, ptr(mps), ptr(md), ptr(obs), len(mps), ptr(occ), C.c_float(thd),
                                               C.c_float(self.ratio), ptr(out), C.byref(n)))
        return n.value, out[: frame.n]

    # ---- S1 Frame::computeStereoMatches (src/frame.cpp:362-472)
    def computeStereoMatches(self, left_ex, left_frame, right_ex, right_frame, kp_left, desc_left, kp_right, desc_right, bf,
                             max_d, mode=0):
        kl = np.ascontiguousarray(kp_left, dtype=KP_DTYPE)
        kr = np.ascontiguousarray(kp_right, dtype=KP_DTYPE)
        dl = np.ascontiguousarray(desc_left, dtype=np.uint8)
        dr = np.ascontiguousarray(desc_right, dtype=np.uint8)
        rx = np.zeros(max(len(kl), 1), np.float32)
        dp = np.zeros(max(len(kl), 1), np.float32)
        check(lib().ydorb_stereo_match(self._h, left_ex._h, left_frame, right_ex._h, right_frame, ptr(kl), ptr(dl), len(kl),
                                       ptr(kr), ptr(dr), len(kr), C.c_float(bf), C.c_float(max_d), mode, ptr(rx), ptr(dp)))
        return rx[: len(kl)], dp[: len(kl)]
