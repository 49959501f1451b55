"""Host logic of the projection searches: what each OrbMatcher member does between the projection of a map point and
the getKeyPointsInArea call, restated as one YdWindowQuery row per map point plus one YdWindowParams per function.

The arithmetic here is float32 with one rounding per operation, as the reference's float expressions compile on
x86-64 without FMA.  The projection itself (rotation * pos + translation, division by z) is the caller's: in a
reference build the C++ shim keeps the reference's own cv::Mat lines for it.
"""
import numpy as np

from ._lib import MAX_LEVELS, Q_ACTIVE, Q_BLOCKS, WINDOW_QUERY_DTYPE, WindowParams

NO_LO, NO_HI = -(1 << 30), 1 << 30
HIGH_THD, LOW_THD = 100, 50  # src/orbMatcher.cpp:7-8


def _f32(a):
    return np.ascontiguousarray(a, dtype=np.float32)


def _params(max_dist, strict, exclusive, check_orientation, chi2_gate=False, inv_sf2=None):
    p = WindowParams()
    p.max_dist, p.strict, p.exclusive = int(max_dist), int(strict), int(exclusive)
    p.check_orientation, p.chi2_gate = int(check_orientation), int(chi2_gate)
    for i in range(MAX_LEVELS):
        p.inv_sf2[i] = float(inv_sf2[min(i, len(inv_sf2) - 1)]) if inv_sf2 is not None else 1.0
    return p


def _queries(n):
    q = np.zeros(n, WINDOW_QUERY_DTYPE)
    q["min_level"] = -1
    q["max_level"] = -1
    q["oct_lo"] = NO_LO
    q["oct_hi"] = NO_HI
    q["right_r"] = -1.0
    return q


def last_and_current_frame(proj_x, proj_y, proj_right_x, octave, angle, active, obs_positive, scale_factors, thd, motion,
                           check_orientation=True):
    """searchByProjectionInLastAndCurrentFrame, src/orbMatcher.cpp:65-155.
    proj_right_x = projX - bf/z (:109); octave/angle of the last frame's keypoint; active = map point present, not an
    outlier, z >= 0 and isInImage (:83, :92); obs_positive = the map point's observations > 0 (what makes a claim block
    later points, :106-107); motion in {'forward', 'backward', 'none'} (:79-80, :95-101)."""
    sf = _f32(scale_factors)
    octave = np.asarray(octave, np.int32)
    q = _queries(len(octave))
    q["x"], q["y"] = _f32(proj_x), _f32(proj_y)
    q["r"] = np.float32(thd) * sf[octave]  # :93
    if motion == "forward":
        q["min_level"] = octave  # :96 (maxLevel defaulted to -1)
    elif motion == "backward":
        q["min_level"], q["max_level"] = 0, octave  # :98
    else:
        q["min_level"], q["max_level"] = octave - 1, octave + 1  # :100
    q["right_x"], q["right_r"] = _f32(proj_right_x), q["r"]  # :108-109
    q["angle"] = _f32(angle)
    q["flags"] = np.where(active, Q_ACTIVE, 0) | np.where(obs_positive, Q_BLOCKS, 0)
    return q, _params(HIGH_THD, strict=True, exclusive=True, check_orientation=check_orientation)  # :117


def keyframe_and_current_frame(proj_x, proj_y, predicted_level, angle, active, scale_factors, thd, orb_dist,
                               check_orientation=True):
    """searchByProjectionInKeyFrameAndCurrentFrame, src/orbMatcher.cpp:156-239.  active = map point present, not bad,
    not in the found set, z >= 0, isInImage, distance inside the invariance range (:170, :187-189)."""
    sf = _f32(scale_factors)
    lvl = np.asarray(predicted_level, np.int32)
    q = _queries(len(lvl))
    q["x"], q["y"] = _f32(proj_x), _f32(proj_y)
    q["r"] = np.float32(thd) * sf[lvl]  # :185
    q["min_level"], q["max_level"] = lvl - 1, lvl + 1  # :186
    q["angle"] = _f32(angle)
    q["flags"] = np.where(active, Q_ACTIVE | Q_BLOCKS, 0)  # any non-null slot blocks (:194)
    return q, _params(orb_dist, strict=False, exclusive=True, check_orientation=check_orientation)  # :202


def _sim_like(proj_x, proj_y, predicted_level, active, scale_factors, thd):
    sf = _f32(scale_factors)
    lvl = np.asarray(predicted_level, np.int32)
    q = _queries(len(lvl))
    q["x"], q["y"] = _f32(proj_x), _f32(proj_y)
    q["r"] = np.float32(thd) * sf[lvl]  # :269, :606, :702, :771
    q["oct_lo"], q["oct_hi"] = lvl - 1, lvl  # :285-286, :616, :716-717, :782
    return q


def in_sim(proj_x, proj_y, predicted_level, active, scale_factors, thd):
    """searchByProjectionInSim, src/orbMatcher.cpp:240-302 (kp_occupied = _vSptrMatchedMapPoints[idx] non-null)."""
    q = _sim_like(proj_x, proj_y, predicted_level, active, scale_factors, thd)
    q["flags"] = np.where(active, Q_ACTIVE | Q_BLOCKS, 0)  # :284
    return q, _params(LOW_THD, strict=False, exclusive=True, check_orientation=False)  # :294


def sim3_direction(proj_x, proj_y, predicted_level, active, scale_factors, thd):
    """One direction of searchBySim3, src/orbMatcher.cpp:593-630 / :633-667 (no exclusion, result per map point)."""
    q = _sim_like(proj_x, proj_y, predicted_level, active, scale_factors, thd)
    q["flags"] = np.where(active, Q_ACTIVE, 0)
    return q, _params(HIGH_THD, strict=False, exclusive=False, check_orientation=False)  # :624, :661


def sim3_agreement(first_matches, second_matches):
    """The agreement check of searchBySim3, src/orbMatcher.cpp:669-679: returns (foundNum, matched second index per
    first index or -1)."""
    first_matches = np.asarray(first_matches)
    second_matches = np.asarray(second_matches)
    out = np.full(len(first_matches), -1, np.int32)
    for i, j in enumerate(first_matches):
        if j >= 0 and second_matches[j] == i:
            out[i] = j
    return int((out >= 0).sum()), out


def fuse_by_projection(proj_x, proj_y, proj_right_x, predicted_level, active, scale_factors, inv_sf2, thd):
    """Candidate search of fuseByProjection, src/orbMatcher.cpp:682-727 (map mutation :728-738 stays on the host)."""
    q = _sim_like(proj_x, proj_y, predicted_level, active, scale_factors, thd)
    q["right_x"] = _f32(proj_right_x)  # :696, used by the chi2 gate :715
    q["flags"] = np.where(active, Q_ACTIVE, 0)
    return q, _params(LOW_THD, strict=False, exclusive=False, check_orientation=False, chi2_gate=True, inv_sf2=inv_sf2)


def fuse_by_sim3(proj_x, proj_y, predicted_level, active, scale_factors, thd):
    """Candidate search of fuseBySim3, src/orbMatcher.cpp:746-790 (map mutation :791-799 stays on the host)."""
    q = _sim_like(proj_x, proj_y, predicted_level, active, scale_factors, thd)
    q["flags"] = np.where(active, Q_ACTIVE, 0)
    return q, _params(LOW_THD, strict=False, exclusive=False, check_orientation=False)
