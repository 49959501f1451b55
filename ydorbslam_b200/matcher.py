"""Host-side mirror of YDORBSLAM::OrbMatcher (src/orbMatcher.hpp:24-66) and the Frame hooks it relies on
(src/frame.hpp:83,96,166) over the C ABI.  All compute happens in libydorb.so on the GPU.
"""
import ctypes as C

import numpy as np

from . import _lib
from . import queries as Q
from ._lib import KP_DTYPE, MPPROJ_DTYPE, TOP2_DTYPE, WINDOW_QUERY_DTYPE, FeatureVector, check, lib, ptr

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

    def searchByProjectionInFrameAndMapPoint_device(self, frame, d_map_points, d_mp_descriptors, d_mp_obs_positive, n_map_points,
                                                    d_match_out, thd=3.0, d_kp_occupied=None):
        """Raw device addresses (ints); d_match_out int32[n_kp] on the device.  Returns the match count."""
        n = C.c_int32()
        check(lib().ydorb_search_by_projection_device(frame._h, ptr(d_map_points), ptr(d_mp_descriptors), ptr(d_mp_obs_positive),
                                                      n_map_points, ptr(d_kp_occupied), C.c_float(thd), C.c_float(self.ratio),
                                                      ptr(d_match_out), C.byref(n)))
        return n.value

    # ---- M2/M3/M4: everything after the projection (include/ydorb.h ydorb_search_windows)
    def searchWindows(self, frame, queries, q_descriptors, params, kp_occupied=None):
        """Returns (n_matches, best_of[nq], match_of_keypoint[n_kp]); match codes: >=0 query, -1 untouched, -2 cleared."""
        q = np.ascontiguousarray(queries, dtype=WINDOW_QUERY_DTYPE)
        qd = np.ascontiguousarray(q_descriptors, dtype=np.uint8).reshape(-1, 32)
        assert len(qd) == len(q)
        occ = None if kp_occupied is None else np.ascontiguousarray(kp_occupied, dtype=np.uint8)
        best = np.full(max(len(q), 1), -1, np.int32)
        match = np.full(max(frame.n, 1), -1, np.int32)
        n = C.c_int32()
        check(lib().ydorb_search_windows(frame._h, ptr(q), ptr(qd), len(q), C.byref(params), ptr(occ), ptr(best), ptr(match),
                                         C.byref(n)))
        return n.value, best[: len(q)], match[: frame.n]

    def searchForInitialization(self, frame2, kps1, desc1, prev_matched, window_size=100, nn_ratio=None, th_low=LOW_THD,
                                area_mode=0):
        """EXTRA: `north_star` names searchForInitialization, the reference has none (src/orbMatcher.hpp:24-66, SURVEY
        F12).  ORB-SLAM2's published algorithm on this reference's Frame hooks (ydorb_match_windowed_top2); no reference
        parity claim.  Returns (n_matches, matches12, prev_matched updated)."""
        k1 = np.ascontiguousarray(kps1, dtype=KP_DTYPE)
        d1 = np.ascontiguousarray(desc1, dtype=np.uint8)
        prev = np.array(prev_matched, dtype=np.float32).reshape(len(k1), 2).copy()
        m12 = np.full(max(len(k1), 1), -1, np.int32)
        n = C.c_int32()
        check(lib().ydorb_match_windowed_top2(frame2._h, ptr(k1), ptr(d1), len(k1), ptr(prev), int(window_size),
                                              C.c_float(self.ratio if nn_ratio is None else nn_ratio), int(th_low),
                                              int(self.check_orientation), int(area_mode), ptr(m12), C.byref(n)))
        return n.value, m12[: len(k1)], prev

    def searchByProjectionInLastAndCurrentFrame(self, cur_frame, proj_x, proj_y, proj_right_x, octave, angle, active,
                                                obs_positive, q_descriptors, scale_factors, thd, motion="none",
                                                kp_occupied=None):
        """src/orbMatcher.cpp:65-155.  kp_occupied[i]: the current frame's slot i holds a map point with observations>0."""
        q, prm = Q.last_and_current_frame(proj_x, proj_y, proj_right_x, octave, angle, active, obs_positive, scale_factors,
                                          thd, motion, self.check_orientation)
        n, _, match = self.searchWindows(cur_frame, q, q_descriptors, prm, kp_occupied)
        return n, match

    def searchByProjectionInKeyFrameAndCurrentFrame(self, cur_frame, proj_x, proj_y, predicted_level, angle, active,
                                                    q_descriptors, scale_factors, thd, orb_dist, kp_occupied=None):
        """src/orbMatcher.cpp:156-239.  kp_occupied[i]: the current frame's slot i is non-null."""
        q, prm = Q.keyframe_and_current_frame(proj_x, proj_y, predicted_level, angle, active, scale_factors, thd, orb_dist,
                                              self.check_orientation)
        n, _, match = self.searchWindows(cur_frame, q, q_descriptors, prm, kp_occupied)
        return n, match

    def searchByProjectionInSim(self, kf_frame, proj_x, proj_y, predicted_level, active, q_descriptors, scale_factors, thd,
                                matched=None):
        """src/orbMatcher.cpp:240-302.  matched[i]: _vSptrMatchedMapPoints[i] non-null on entry."""
        q, prm = Q.in_sim(proj_x, proj_y, predicted_level, active, scale_factors, thd)
        n, _, match = self.searchWindows(kf_frame, q, q_descriptors, prm, matched)
        return n, match

    def searchBySim3(self, first_frame, second_frame, first_to_second, second_to_first, scale_factors, thd):
        """src/orbMatcher.cpp:566-681.  first_to_second = (proj_x, proj_y, predicted_level, active, descriptors) of the
        first key frame's map points projected into the second (active already excludes :583-590's matched points), and
        vice versa.  Returns (foundNum, second index per first index or -1)."""
        q12, p12 = Q.sim3_direction(*first_to_second[:4], scale_factors, thd)
        q21, p21 = Q.sim3_direction(*second_to_first[:4], scale_factors, thd)
        _, best12, _ = self.searchWindows(second_frame, q12, first_to_second[4], p12)
        _, best21, _ = self.searchWindows(first_frame, q21, second_to_first[4], p21)
        return Q.sim3_agreement(best12, best21)

    def fuseByProjectionCandidates(self, kf_frame, proj_x, proj_y, proj_right_x, predicted_level, active, q_descriptors,
                                   scale_factors, inv_sf2, thd):
        """Candidate search of fuseByProjection (src/orbMatcher.cpp:682-727): best key point per map point or -1."""
        q, prm = Q.fuse_by_projection(proj_x, proj_y, proj_right_x, predicted_level, active, scale_factors, inv_sf2, thd)
        n, best, _ = self.searchWindows(kf_frame, q, q_descriptors, prm)
        return n, best

    def fuseBySim3Candidates(self, kf_frame, proj_x, proj_y, predicted_level, active, q_descriptors, scale_factors, thd):
        """Candidate search of fuseBySim3 (src/orbMatcher.cpp:746-790)."""
        q, prm = Q.fuse_by_sim3(proj_x, proj_y, predicted_level, active, scale_factors, thd)
        n, best, _ = self.searchWindows(kf_frame, q, q_descriptors, prm)
        return n, best

    # ---- M5 BoW matchers (src/orbMatcher.cpp:303-462); feature vectors as (node_ids, node_start, items) CSR triples
    def _bow(self, mode, kps1, desc1, valid1, fv1, kps2, desc2, valid2, fv2):
        k1 = np.ascontiguousarray(kps1, dtype=KP_DTYPE)
        k2 = np.ascontiguousarray(kps2, dtype=KP_DTYPE)
        d1 = np.ascontiguousarray(desc1, dtype=np.uint8).reshape(-1, 32)
        d2 = np.ascontiguousarray(desc2, dtype=np.uint8).reshape(-1, 32)
        v1 = np.ascontiguousarray(valid1, dtype=np.uint8)
        v2 = None if valid2 is None else np.ascontiguousarray(valid2, dtype=np.uint8)
        f1, f2 = FeatureVector.from_arrays(*fv1), FeatureVector.from_arrays(*fv2)
        out = np.full(max(len(k2) if mode == 0 else len(k1), 1), -1, np.int32)
        n = C.c_int32()
        check(lib().ydorb_search_by_bow(self._h, ptr(k1), ptr(d1), ptr(v1), len(k1), C.byref(f1), ptr(k2), ptr(d2), ptr(v2),
                                        len(k2), C.byref(f2), mode, C.c_float(self.ratio), int(self.check_orientation),
                                        ptr(out), C.byref(n)))
        return n.value, out[: (len(k2) if mode == 0 else len(k1))]

    def searchByBowInKeyFrameAndFrame(self, kf_kps, kf_desc, kf_valid, kf_fv, fr_kps, fr_desc, fr_fv):
        """src/orbMatcher.cpp:303-379: returns (matchNum, key-frame index per frame key point or -1)."""
        return self._bow(0, kf_kps, kf_desc, kf_valid, kf_fv, fr_kps, fr_desc, None, fr_fv)

    def searchByBowInTwoKeyFrames(self, kps1, desc1, valid1, fv1, kps2, desc2, valid2, fv2):
        """src/orbMatcher.cpp:380-462: returns (matchNum, second index per first key point or -1)."""
        return self._bow(1, kps1, desc1, valid1, fv1, kps2, desc2, valid2, fv2)

    # ---- M6 searchForTriangulation (src/orbMatcher.cpp:463-565)
    def searchForTriangulation(self, kps1, desc1, right_x1, has_mp1, fv1, kps2, desc2, right_x2, has_mp2, fv2, F, epipole,
                               scale_factors, scale_factor_squares, stereo_only=False):
        """Returns (matchNum, pairs[k, 2]) with pairs in ascending first index as :555-563 emits them."""
        k1 = np.ascontiguousarray(kps1, dtype=KP_DTYPE)
        k2 = np.ascontiguousarray(kps2, dtype=KP_DTYPE)
        d1 = np.ascontiguousarray(desc1, dtype=np.uint8).reshape(-1, 32)
        d2 = np.ascontiguousarray(desc2, dtype=np.uint8).reshape(-1, 32)
        r1 = np.ascontiguousarray(right_x1, dtype=np.float32)
        r2 = np.ascontiguousarray(right_x2, dtype=np.float32)
        h1 = np.ascontiguousarray(has_mp1, dtype=np.uint8)
        h2 = np.ascontiguousarray(has_mp2, dtype=np.uint8)
        f1, f2 = FeatureVector.from_arrays(*fv1), FeatureVector.from_arrays(*fv2)
        Fm = np.ascontiguousarray(F, dtype=np.float32).reshape(9)
        sf = np.ascontiguousarray(scale_factors, dtype=np.float32)
        sf2 = np.ascontiguousarray(scale_factor_squares, dtype=np.float32)
        out = np.full(max(len(k1), 1), -1, np.int32)
        n = C.c_int32()
        check(lib().ydorb_search_for_triangulation(self._h, ptr(k1), ptr(d1), ptr(r1), ptr(h1), len(k1), C.byref(f1), ptr(k2),
                                                   ptr(d2), ptr(r2), ptr(h2), len(k2), C.byref(f2), ptr(Fm),
                                                   C.c_float(epipole[0]), C.c_float(epipole[1]), ptr(sf), ptr(sf2), len(sf),
                                                   int(stereo_only), int(self.check_orientation), ptr(out), C.byref(n)))
        out = out[: len(k1)]
        first = np.nonzero(out >= 0)[0]
        return n.value, np.stack([first, out[first]], axis=1).astype(np.int32)

    @staticmethod
    def computeThreeMaxima(sizes):
        """src/orbMatcher.cpp:827-854 on the histogram bin sizes."""
        sz = np.ascontiguousarray(sizes, dtype=np.int32)
        a, b, c = C.c_int32(), C.c_int32(), C.c_int32()
        lib().ydorb_three_maxima(ptr(sz), len(sz), C.byref(a), C.byref(b), C.byref(c))
        return a.value, b.value, c.value

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
