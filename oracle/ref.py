"""ctypes binding for oracle/_ref/libydorb_ref.so: the REFERENCE's own translation units, compiled unmodified
(oracle/refshim/Makefile) against the OpenCV stand-in headers.  TEST INFRASTRUCTURE -- the top of the parity chain:

    _ref (reference code)  ==  oracle (restatement)  ==  tests/golden fixtures;   CUDA == oracle == _ref

Only tests/, __graft_entry__ and bench.py's cpu_baseline / --impl reference legs may import this.  The library is built
in the container that has /root/reference; the GPU box receives the prebuilt .so with the snapshot (never the sources).
"""
import ctypes as C
import os
import pathlib
import subprocess

import numpy as np

from . import orc

_HERE = pathlib.Path(__file__).resolve().parent
_SO = _HERE / "_ref" / "libydorb_ref.so"
REFERENCE = pathlib.Path(os.environ.get("YDORB_REFERENCE", "/root/reference"))

KP_DTYPE = orc.KP_DTYPE
CAND_DTYPE = orc.CAND_DTYPE
PAD = orc.PAD


def can_build():
    return (REFERENCE / "src" / "orbExtractor.cpp").exists()


def build():
    """Build (or refresh) oracle/_ref when the reference checkout is present; otherwise use the prebuilt library."""
    orc.build()
    if can_build():
        subprocess.check_call(["make", "-C", str(_HERE / "refshim"), f"REF={REFERENCE}"], stdout=subprocess.DEVNULL)
    return _SO


def available():
    return _SO.exists() or can_build()


_lib = None


def lib():
    global _lib
    if _lib is None:
        build()
        if not _SO.exists():
            raise ImportError(f"{_SO} is missing and {REFERENCE} is not present to build it from")
        _lib = C.CDLL(str(_SO))
    return _lib


def _p(a):
    return a.ctypes.data_as(C.c_void_p) if a is not None else None


def ctor_tables(n_features=1000, scale=1.2, n_levels=8, ini_th=20, min_th=7):
    """Tables of OrbExtractor::OrbExtractor (src/orbExtractor.cpp:315-354) as the compiled reference produces them."""
    sf, isf, sf2, isf2 = (np.zeros(n_levels, np.float32) for _ in range(4))
    quota = np.zeros(n_levels, np.int32)
    umax = np.zeros(64, np.int32)
    th = np.zeros(2, np.int32)
    n = lib().ref_ctor_tables(
        n_features, C.c_float(scale), n_levels, ini_th, min_th, _p(sf), _p(isf), _p(sf2), _p(isf2), _p(quota), _p(umax), 64, _p(th)
    )
    return dict(sf=sf, inv_sf=isf, sf2=sf2, inv_sf2=isf2, quota=quota, umax=umax[:n], ini_th=int(th[0]), min_th=int(th[1]))


def extract(img, n_features=1000, scale=1.2, n_levels=8, ini_th=20, min_th=7, cap=None, want_pyramid=False,
            want_candidates=False):
    """extractAndCompute on a fresh reference OrbExtractor.  Returns (kps, desc, per_level_n[, pyramid][, candidates])."""
    img = np.ascontiguousarray(img, dtype=np.uint8)
    h, w = img.shape
    cap = cap or (n_features + 64)
    kps = np.zeros(cap, KP_DTYPE)
    desc = np.zeros((cap, 32), np.uint8)
    per = np.zeros(n_levels, np.int32)
    pyr = off = None
    pyr_cap = 0
    if want_pyramid:
        pyr_cap = int(sum((int(w * scale ** -l) + 40) * (int(h * scale ** -l) + 40) for l in range(n_levels)))
        pyr = np.zeros(pyr_cap, np.uint8)
        off = np.zeros(n_levels, np.int64)
    cand = start = None
    cand_cap = 0
    if want_candidates:
        cand_cap = w * h // 4 + 1024
        cand = np.zeros(cand_cap, CAND_DTYPE)
        start = np.zeros(n_levels + 1, np.int32)
    n = lib().ref_extract(
        n_features, C.c_float(scale), n_levels, ini_th, min_th, _p(img), w, h, w, _p(kps), _p(desc), cap, _p(per), _p(pyr),
        _p(off), C.c_int64(pyr_cap), _p(cand), cand_cap, _p(start),
    )
    assert 0 <= n <= cap, n
    out = [kps[:n], desc[:n], per]
    if want_pyramid:
        levels = []
        for l in range(n_levels):
            end = off[l + 1] if l + 1 < n_levels else None
            nxt = int(end) if end is not None else None
            levels.append((int(off[l]), nxt))
        out.append((pyr, levels))
    if want_candidates:
        assert start[-1] <= cand_cap
        out.append([cand[start[l] : start[l + 1]] for l in range(n_levels)])
    return tuple(out)


def pyramid_levels(pyr_and_levels, level_sizes):
    """Split the flat pyramid buffer returned by extract(want_pyramid=True) into padded (lh+38, lw+38) arrays."""
    pyr, levels = pyr_and_levels
    out = []
    for (o, _), (lw, lh) in zip(levels, level_sizes):
        out.append(pyr[o : o + (lw + 2 * PAD) * (lh + 2 * PAD)].reshape(lh + 2 * PAD, lw + 2 * PAD))
    return out


def quadtree(cand, rootw, rooth, want):
    """distributeQuadTree (src/orbExtractor.cpp:455-544) of the compiled reference; returns indices into cand."""
    cand = np.ascontiguousarray(cand, dtype=CAND_DTYPE)
    sel = np.zeros(max(2 * want + 8, len(cand) + 8), np.int32)
    m = lib().ref_quadtree(_p(cand), len(cand), rootw, rooth, want, _p(sel))
    return sel[:m]


def extract_batch(imgs, threads, n_features=1000, scale=1.2, n_levels=8, ini_th=20, min_th=7, cap=None):
    imgs = np.ascontiguousarray(imgs, dtype=np.uint8)
    b, h, w = imgs.shape
    cap = cap or (n_features + 64)
    kps = np.zeros((b, cap), KP_DTYPE)
    desc = np.zeros((b, cap, 32), np.uint8)
    n = np.zeros(b, np.int32)
    lib().ref_extract_batch(
        n_features, C.c_float(scale), n_levels, ini_th, min_th, _p(imgs), w, h, w, C.c_int64(h * w), b, _p(kps), _p(desc), cap,
        _p(n), threads,
    )
    return kps, desc, n


# ---- Frame / OrbMatcher of the compiled reference (oracle/refshim/ref_frame_driver.cpp) -----------------------------

def descriptor_distance(a, b):
    """OrbMatcher::computeDescriptorsDistance (src/orbMatcher.cpp:11-23) of the compiled reference."""
    a = np.ascontiguousarray(a, dtype=np.uint8)
    b = np.ascontiguousarray(b, dtype=np.uint8)
    assert a.size == 32 and b.size == 32
    return int(lib().ref_descriptor_distance(_p(a), _p(b)))


class RefFrame:
    """A YDORBSLAM::Frame of the compiled reference: either its stereo constructor (src/frame.cpp:57-110: two
    extractors on two threads, computeStereoMatches, computeImageBounds, assignKeyPointsToGrid) or a frame holding
    given key points with the grid built by assignKeyPointsToGrid."""

    def __init__(self, handle):
        self._h = C.c_void_p(handle)
        n = np.zeros(2, np.int32)
        lib().ref_frame_sizes(self._h, _p(n))
        self.n_left, self.n_right = int(n[0]), int(n[1])

    @classmethod
    def stereo(cls, left, right, fx, fy, cx, cy, bf, depth_thd=35.0, n_features=2000, scale=1.2, n_levels=8, ini_th=20, min_th=7,
               baseline_before=0.0):
        """baseline_before: value of the static Frame::m_flt_baseLine when the constructor runs -- 0 for the first frame
        of a run (maxD = +inf), bf / fx afterwards (src/frame.cpp:101 vs :108)."""
        left = np.ascontiguousarray(left, dtype=np.uint8)
        right = np.ascontiguousarray(right, dtype=np.uint8)
        h, w = left.shape
        assert right.shape == left.shape
        lib().ref_frame_stereo_create.restype = C.c_void_p
        hdl = lib().ref_frame_stereo_create(n_features, C.c_float(scale), n_levels, ini_th, min_th, _p(left), _p(right), w, h, w,
                                            C.c_float(fx), C.c_float(fy), C.c_float(cx), C.c_float(cy), C.c_float(bf),
                                            C.c_float(depth_thd), C.c_float(baseline_before))
        return cls(hdl)

    @classmethod
    def from_keypoints(cls, kps, bounds):
        kps = np.ascontiguousarray(kps, dtype=KP_DTYPE)
        lib().ref_frame_from_keypoints.restype = C.c_void_p
        hdl = lib().ref_frame_from_keypoints(_p(kps), len(kps), *(C.c_float(float(v)) for v in bounds))
        return cls(hdl)

    def get(self):
        """(kps_left, desc_left, kps_right, desc_right, right_x, depth, bounds)."""
        kl, kr = np.zeros(self.n_left, KP_DTYPE), np.zeros(self.n_right, KP_DTYPE)
        dl, dr = np.zeros((self.n_left, 32), np.uint8), np.zeros((self.n_right, 32), np.uint8)
        rx, dp = np.zeros(self.n_left, np.float32), np.zeros(self.n_left, np.float32)
        b = np.zeros(4, np.float32)
        lib().ref_frame_get(self._h, _p(kl), _p(dl), _p(kr), _p(dr), _p(rx), _p(dp), _p(b))
        return kl, dl, kr, dr, rx, dp, b

    def keypoints_in_area(self, x, y, r, min_level=-1, max_level=-1):
        out = np.zeros(max(self.n_left, 1), np.int32)
        n = lib().ref_frame_keypoints_in_area(self._h, C.c_float(x), C.c_float(y), C.c_float(r), int(min_level), int(max_level),
                                              _p(out), len(out))
        return out[:n].copy()

    def set_features(self, desc, right_x, scale_factors):
        """Descriptors, rightX and scale factors of a frame made by from_keypoints (what M1 reads from the frame)."""
        desc = np.ascontiguousarray(desc, dtype=np.uint8)
        rx = None if right_x is None else np.ascontiguousarray(right_x, dtype=np.float32)
        sf = np.ascontiguousarray(scale_factors, dtype=np.float32)
        assert desc.shape == (self.n_left, 32)
        lib().ref_frame_set_features(self._h, _p(desc), _p(rx), _p(sf), len(sf))

    def search_by_projection(self, mps, mp_desc, mp_obs_positive, kp_occupied, thd, ratio):
        """OrbMatcher(ratio).searchByProjectionInFrameAndMapPoint(frame, mapPoints, thd) of the compiled reference
        (src/orbMatcher.cpp:24-64); the MapPoint objects are built by the reference's own constructor."""
        mps = np.ascontiguousarray(mps, dtype=orc.MPPROJ_DTYPE)
        mpd = np.ascontiguousarray(mp_desc, dtype=np.uint8)
        obs = np.ascontiguousarray(mp_obs_positive, dtype=np.uint8)
        occ = None if kp_occupied is None else np.ascontiguousarray(kp_occupied, dtype=np.uint8)
        out = np.full(max(self.n_left, 1), -1, np.int32)
        n = lib().ref_search_by_projection(self._h, _p(mps), _p(mpd), _p(obs), len(mps), _p(occ), C.c_float(thd), C.c_float(ratio),
                                           _p(out))
        return n, out[: self.n_left]

    def search_last_current(self, last_kps, mp_pos, mp_desc, mp_obs_positive, outlier, last_t, cam, kp_occupied, thd, ratio,
                            check_orientation=True):
        """OrbMatcher(ratio, check).searchByProjectionInLastAndCurrentFrame(this frame [pose I], last frame, thd) of the
        compiled reference (src/orbMatcher.cpp:65-155).  mp_pos [n_last][3] (NaN x: no map point), last_t: translation of
        the last frame's pose, cam = (fx, fy, cx, cy, bf, baseline).  Returns (return value, match per current key point)."""
        lk = np.ascontiguousarray(last_kps, dtype=KP_DTYPE)
        pos = np.ascontiguousarray(mp_pos, dtype=np.float32)
        mpd = np.ascontiguousarray(mp_desc, dtype=np.uint8)
        obs = np.ascontiguousarray(mp_obs_positive, dtype=np.uint8)
        out_ = np.ascontiguousarray(outlier, dtype=np.uint8)
        lt = np.ascontiguousarray(last_t, dtype=np.float32)
        cm = np.ascontiguousarray(cam, dtype=np.float32)
        occ = None if kp_occupied is None else np.ascontiguousarray(kp_occupied, dtype=np.uint8)
        match = np.full(max(self.n_left, 1), -1, np.int32)
        n = lib().ref_search_last_current(self._h, _p(lk), len(lk), _p(pos), _p(mpd), _p(obs), _p(out_), _p(lt), _p(cm), _p(occ),
                                          C.c_float(thd), C.c_float(ratio), int(bool(check_orientation)), _p(match))
        return n, match[: self.n_left]

    def search_keyframe_current(self, kf_kps, mp_pos, mp_desc, min_dist, max_dist, is_bad, in_found, cam, kp_occupied, thd,
                                orb_dist, ratio, check_orientation=True):
        """OrbMatcher::searchByProjectionInKeyFrameAndCurrentFrame (src/orbMatcher.cpp:156-239) of the compiled reference,
        this frame (pose I) as the current frame.  Returns (return value, match per current key point, dbg [n_kf][4] =
        projX, projY, predicted level, gates passed -- evaluated by the reference's own methods)."""
        kk = np.ascontiguousarray(kf_kps, dtype=KP_DTYPE)
        arrs = [np.ascontiguousarray(a, dtype=t) for a, t in ((mp_pos, np.float32), (mp_desc, np.uint8), (min_dist, np.float32),
                                                               (max_dist, np.float32), (is_bad, np.uint8), (in_found, np.uint8),
                                                               (cam, np.float32))]
        occ = None if kp_occupied is None else np.ascontiguousarray(kp_occupied, dtype=np.uint8)
        match = np.full(max(self.n_left, 1), -1, np.int32)
        dbg = np.zeros((max(len(kk), 1), 4), np.float32)
        n = lib().ref_search_keyframe_current(self._h, _p(kk), len(kk), *[_p(a) for a in arrs], _p(occ), C.c_float(thd), int(orb_dist),
                                              C.c_float(ratio), int(bool(check_orientation)), _p(match), _p(dbg))
        return n, match[: self.n_left], dbg[: len(kk)]

    def search_in_sim(self, mp_pos, mp_desc, min_dist, max_dist, is_bad, cam, kp_occupied, thd, ratio=0.6):
        """OrbMatcher::searchByProjectionInSim (src/orbMatcher.cpp:240-302) of the compiled reference with the identity
        similarity, this frame as the KeyFrame.  Returns (return value, match per key point, dbg [n_mp][4])."""
        arrs = [np.ascontiguousarray(a, dtype=t) for a, t in ((mp_pos, np.float32), (mp_desc, np.uint8), (min_dist, np.float32),
                                                               (max_dist, np.float32), (is_bad, np.uint8), (cam, np.float32))]
        n_mp = len(arrs[0])
        occ = None if kp_occupied is None else np.ascontiguousarray(kp_occupied, dtype=np.uint8)
        match = np.full(max(self.n_left, 1), -1, np.int32)
        dbg = np.zeros((max(n_mp, 1), 4), np.float32)
        n = lib().ref_search_in_sim(self._h, n_mp, *[_p(a) for a in arrs], _p(occ), int(thd), C.c_float(ratio), _p(match), _p(dbg))
        return n, match[: self.n_left], dbg[:n_mp]

    def close(self):
        if self._h:
            lib().ref_frame_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def search_by_bow_keyframe_frame(kps1, desc1, valid1, fv1, kps2, desc2, fv2, ratio, check_orientation=True):
    """OrbMatcher::searchByBowInKeyFrameAndFrame (src/orbMatcher.cpp:303-379) of the compiled reference on a KeyFrame /
    Frame pair built from the arrays; fv = (node_ids, node_start, items).  Returns (return value, match per side-2 key point)."""
    k1, k2 = np.ascontiguousarray(kps1, dtype=KP_DTYPE), np.ascontiguousarray(kps2, dtype=KP_DTYPE)
    d1, d2 = np.ascontiguousarray(desc1, dtype=np.uint8), np.ascontiguousarray(desc2, dtype=np.uint8)
    v1 = np.ascontiguousarray(valid1, dtype=np.uint8)
    a = [np.ascontiguousarray(x, dtype=np.int32) for x in fv1]
    b = [np.ascontiguousarray(x, dtype=np.int32) for x in fv2]
    out = np.full(max(len(k2), 1), -1, np.int32)
    n = lib().ref_search_by_bow_kf_frame(_p(k1), _p(d1), _p(v1), len(k1), _p(a[0]), _p(a[1]), _p(a[2]), len(a[0]), _p(k2), _p(d2),
                                         len(k2), _p(b[0]), _p(b[1]), _p(b[2]), len(b[0]), C.c_float(ratio), int(bool(check_orientation)),
                                         _p(out))
    return n, out[: len(k2)]


def search_for_triangulation(kps1, desc1, rx1, has_mp1, fv1, kps2, desc2, rx2, has_mp2, fv2, F, origin1, cam, sf, sf2, stereo_only,
                             ratio=0.6, check_orientation=True):
    """OrbMatcher::searchForTriangulation (src/orbMatcher.cpp:463-565) of the compiled reference on two KeyFrames built from
    the arrays; origin1 = camera origin of key frame 1 in the frame of key frame 2 (identity pose).  Returns (return value,
    pairs [n][2], epipole as the reference computed it)."""
    k1, k2 = np.ascontiguousarray(kps1, dtype=KP_DTYPE), np.ascontiguousarray(kps2, dtype=KP_DTYPE)
    d1, d2 = np.ascontiguousarray(desc1, dtype=np.uint8), np.ascontiguousarray(desc2, dtype=np.uint8)
    r1, r2 = np.ascontiguousarray(rx1, dtype=np.float32), np.ascontiguousarray(rx2, dtype=np.float32)
    h1, h2 = np.ascontiguousarray(has_mp1, dtype=np.uint8), np.ascontiguousarray(has_mp2, dtype=np.uint8)
    a = [np.ascontiguousarray(x, dtype=np.int32) for x in fv1]
    b = [np.ascontiguousarray(x, dtype=np.int32) for x in fv2]
    Fm = np.ascontiguousarray(F, dtype=np.float32).reshape(9)
    o1 = np.ascontiguousarray(origin1, dtype=np.float32)
    cm = np.ascontiguousarray(cam, dtype=np.float32)
    sfa, sf2a = np.ascontiguousarray(sf, dtype=np.float32), np.ascontiguousarray(sf2, dtype=np.float32)
    pairs = np.zeros((max(len(k1), 1), 2), np.int32)
    ep = np.zeros(2, np.float32)
    n = lib().ref_search_for_triangulation(_p(k1), _p(d1), _p(r1), _p(h1), len(k1), _p(a[0]), _p(a[1]), _p(a[2]), len(a[0]),
                                           _p(k2), _p(d2), _p(r2), _p(h2), len(k2), _p(b[0]), _p(b[1]), _p(b[2]), len(b[0]), _p(Fm),
                                           _p(o1), _p(cm), _p(sfa), _p(sf2a), len(sfa), int(bool(stereo_only)), C.c_float(ratio),
                                           int(bool(check_orientation)), _p(pairs), _p(ep))
    return n, pairs[:n].copy(), (ep[0], ep[1])


def search_by_bow_two_keyframes(kps1, desc1, valid1, fv1, kps2, desc2, valid2, fv2, ratio, check_orientation=True):
    """OrbMatcher::searchByBowInTwoKeyFrames (src/orbMatcher.cpp:380-462) of the compiled reference.  Returns (return
    value, match per side-1 key point)."""
    k1, k2 = np.ascontiguousarray(kps1, dtype=KP_DTYPE), np.ascontiguousarray(kps2, dtype=KP_DTYPE)
    d1, d2 = np.ascontiguousarray(desc1, dtype=np.uint8), np.ascontiguousarray(desc2, dtype=np.uint8)
    v1, v2 = np.ascontiguousarray(valid1, dtype=np.uint8), np.ascontiguousarray(valid2, dtype=np.uint8)
    a = [np.ascontiguousarray(x, dtype=np.int32) for x in fv1]
    b = [np.ascontiguousarray(x, dtype=np.int32) for x in fv2]
    out = np.full(max(len(k1), 1), -1, np.int32)
    n = lib().ref_search_by_bow_two_keyframes(_p(k1), _p(d1), _p(v1), len(k1), _p(a[0]), _p(a[1]), _p(a[2]), len(a[0]), _p(k2), _p(d2),
                                              _p(v2), len(k2), _p(b[0]), _p(b[1]), _p(b[2]), len(b[0]), C.c_float(ratio),
                                              int(bool(check_orientation)), _p(out))
    return n, out[: len(k1)]


def frustum(R, t, cam, bounds, scale_factor, n_levels, map_points, view_cos_limit=0.5):
    """Frame::isInCameraFrustum + MapPoint::predictScaleLevel of the compiled reference (src/frame.cpp:295-326) for every map
    point (orc.MAPPOINT3D_DTYPE).  Returns orc.MPPROJ_DTYPE records; fields other than `active` are what the reference
    wrote into the MapPoint, i.e. meaningful where active."""
    mp = np.ascontiguousarray(map_points, orc.MAPPOINT3D_DTYPE)
    out = np.zeros(max(len(mp), 1), orc.MPPROJ_DTYPE)
    a = [np.ascontiguousarray(x, dtype=np.float32) for x in (np.asarray(R).reshape(9), t, cam, bounds)]
    lib().ref_frustum(*[_p(x) for x in a], C.c_float(scale_factor), int(n_levels), _p(mp), len(mp), C.c_float(view_cos_limit), _p(out))
    return out[: len(mp)]


def distinctive_descriptors(desc, start):
    """MapPoint::computeDistinctiveDescriptors (src/mapPoint.cpp:169-218) of the compiled reference for CSR observation
    sets; returns the [n_points][32] descriptors it left in the MapPoints (zeros where a point has no observation)."""
    d = np.ascontiguousarray(desc, np.uint8).reshape(-1, 32)
    s = np.ascontiguousarray(start, np.int32)
    out = np.zeros((len(s) - 1, 32), np.uint8)
    lib().ref_distinctive_descriptors(_p(d), _p(s), len(s) - 1, _p(out))
    return out


def search_by_sim3(side1, side2, cam, bounds, sf, thd, pre1=None):
    """OrbMatcher::searchBySim3 (src/orbMatcher.cpp:566-681) of the compiled reference between two KeyFrames at the identity
    pose.  side = dict(kps, desc, pos [n][3] (NaN x: no map point), mp_desc, min_d, max_d, bad).  pre1[i] >= 0: slot i is
    matched on entry to a map point key frame 2 observes at key point pre1[i].  Returns (return value, match per side-1 key
    point (-2 = pre-matched), dbg1, dbg2)."""
    def arrs(sd):
        return (np.ascontiguousarray(sd["kps"], dtype=KP_DTYPE), np.ascontiguousarray(sd["desc"], dtype=np.uint8),
                np.ascontiguousarray(sd["pos"], dtype=np.float32), np.ascontiguousarray(sd["mp_desc"], dtype=np.uint8),
                np.ascontiguousarray(sd["min_d"], dtype=np.float32), np.ascontiguousarray(sd["max_d"], dtype=np.float32),
                np.ascontiguousarray(sd["bad"], dtype=np.uint8))
    a, b = arrs(side1), arrs(side2)
    n1, n2 = len(a[0]), len(b[0])
    pre = np.full(max(n1, 1), -1, np.int32) if pre1 is None else np.ascontiguousarray(pre1, dtype=np.int32)
    cm, bd, sfa = (np.ascontiguousarray(x, dtype=np.float32) for x in (cam, bounds, sf))
    match = np.full(max(n1, 1), -1, np.int32)
    dbg1, dbg2 = np.zeros((max(n1, 1), 4), np.float32), np.zeros((max(n2, 1), 4), np.float32)
    n = lib().ref_search_by_sim3(_p(a[0]), _p(a[1]), n1, _p(a[2]), _p(a[3]), _p(a[4]), _p(a[5]), _p(a[6]), _p(pre),
                                 _p(b[0]), _p(b[1]), n2, _p(b[2]), _p(b[3]), _p(b[4]), _p(b[5]), _p(b[6]), _p(cm), _p(bd), _p(sfa),
                                 len(sfa), C.c_float(thd), _p(match), _p(dbg1), _p(dbg2))
    return n, match[:n1], dbg1[:n1], dbg2[:n2]


def fuse(mode, kps, desc, right_x, slot_kind, slot_obs, points, cam, bounds, sf, thd, sim_scale=1.0, sim_t=(0.0, 0.0, 0.0)):
    """OrbMatcher::fuseByProjection (mode 0, src/orbMatcher.cpp:682-745) / fuseBySim3 (mode 1, :746-807) of the compiled
    reference on a KeyFrame at the identity pose.  points = dict(pos [m][3] (NaN x: null entry), desc, min_d, max_d,
    normal [m][3], bad, obs).  Entities are named slot point i -> i, candidate j -> n + j, none -> -1.  Returns a dict:
    fused (return value), slot (occupant per key point), slot_repl, slot_obs, mp_repl, mp_obs, mp_bad, dbg [m][5]."""
    k = np.ascontiguousarray(kps, dtype=KP_DTYPE)
    n, m = len(k), len(points["pos"])
    d = np.ascontiguousarray(desc, dtype=np.uint8)
    rx = np.ascontiguousarray(right_x, dtype=np.float32)
    sk, so = np.ascontiguousarray(slot_kind, dtype=np.uint8), np.ascontiguousarray(slot_obs, dtype=np.int32)
    pos, mpd = np.ascontiguousarray(points["pos"], dtype=np.float32), np.ascontiguousarray(points["desc"], dtype=np.uint8)
    mind, maxd = np.ascontiguousarray(points["min_d"], dtype=np.float32), np.ascontiguousarray(points["max_d"], dtype=np.float32)
    nrm, bad = np.ascontiguousarray(points["normal"], dtype=np.float32), np.ascontiguousarray(points["bad"], dtype=np.uint8)
    obs = np.ascontiguousarray(points["obs"], dtype=np.int32)
    cm, bd, sfa, st = (np.ascontiguousarray(x, dtype=np.float32) for x in (cam, bounds, sf, sim_t))
    out = dict(slot=np.zeros(max(n, 1), np.int32), slot_repl=np.zeros(max(n, 1), np.int32), slot_obs=np.zeros(max(n, 1), np.int32),
               mp_repl=np.zeros(max(m, 1), np.int32), mp_obs=np.zeros(max(m, 1), np.int32), mp_bad=np.zeros(max(m, 1), np.uint8),
               dbg=np.zeros((max(m, 1), 5), np.float32))
    out["fused"] = lib().ref_fuse(mode, _p(k), _p(d), _p(rx), n, _p(sk), _p(so), _p(pos), _p(mpd), _p(mind), _p(maxd), _p(nrm), _p(bad),
                                  _p(obs), m, _p(cm), _p(bd), _p(sfa), len(sfa), C.c_float(thd), C.c_float(sim_scale), _p(st),
                                  _p(out["slot"]), _p(out["slot_repl"]), _p(out["slot_obs"]), _p(out["mp_repl"]), _p(out["mp_obs"]),
                                  _p(out["mp_bad"]), _p(out["dbg"]))
    for key in ("slot", "slot_repl", "slot_obs"):
        out[key] = out[key][:n]
    for key in ("mp_repl", "mp_obs", "mp_bad", "dbg"):
        out[key] = out[key][:m]
    return out


class RefVocabulary:
    """The reference's own DBoW3::Vocabulary (thirdParty/DBow3/src/Vocabulary.cpp compiled unmodified)."""

    def __init__(self, handle):
        assert handle, "the reference could not build / load the vocabulary"
        self._h = C.c_void_p(handle)

    @classmethod
    def create(cls, k, L, weighting, scoring, desc, image_start, seed=1):
        """Vocabulary(k, L, weighting, scoring).create(training features): the reference's own k-means++ / k-medians."""
        lib().ref_voc_create.restype = C.c_void_p
        d = np.ascontiguousarray(desc, dtype=np.uint8)
        st = np.ascontiguousarray(image_start, dtype=np.int32)
        return cls(lib().ref_voc_create(k, L, weighting, scoring, _p(d), _p(st), len(st) - 1, C.c_uint(seed)))

    @classmethod
    def load(cls, path):
        lib().ref_voc_load.restype = C.c_void_p
        return cls(lib().ref_voc_load(str(path).encode()))

    def save(self, path, compressed):
        assert lib().ref_voc_save(self._h, str(path).encode(), int(compressed)) == 0

    def info(self):
        out = np.zeros(5, np.int32)
        lib().ref_voc_info(self._h, _p(out))
        return dict(k=int(out[0]), L=int(out[1]), weighting=int(out[2]), scoring=int(out[3]), n_words=int(out[4]))

    def transform(self, desc, levels_up=4):
        d = np.ascontiguousarray(desc, dtype=np.uint8).reshape(-1, 32)
        n = len(d)
        m = max(n, 1)
        word, node, wgt = np.zeros(m, np.int32), np.zeros(m, np.int32), np.zeros(m, np.float64)
        bi, bv = np.zeros(m, np.int32), np.zeros(m, np.float64)
        fi, fs, ft = np.zeros(m, np.int32), np.zeros(m + 1, np.int32), np.zeros(m, np.int32)
        nb, nf = C.c_int32(0), C.c_int32(0)
        lib().ref_voc_transform(self._h, _p(d), n, levels_up, _p(word), _p(node), _p(wgt), _p(bi), _p(bv), C.byref(nb), _p(fi), _p(fs),
                                _p(ft), C.byref(nf))
        return dict(word_id=word[:n], node_id=node[:n], weight=wgt[:n], bow_ids=bi[: nb.value], bow_values=bv[: nb.value],
                    fv_node_ids=fi[: nf.value], fv_node_start=fs[: nf.value + 1], fv_items=ft[: int(fs[nf.value])])

    def close(self):
        if self._h:
            lib().ref_voc_destroy(self._h)
            self._h = None
