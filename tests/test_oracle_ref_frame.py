"""The compiled REFERENCE (oracle/_ref: src/frame.cpp, src/orbMatcher.cpp ... unmodified, against the OpenCV stand-in)
vs the oracle restatement, for the rows the extractor tests do not reach:
  M0 computeDescriptorsDistance, G1 assignKeyPointsToGrid, G2 getKeyPointsInArea, S1 computeStereoMatches
(the stereo constructor also runs both extractors on two threads, src/frame.cpp:84-87).  CPU only."""
import numpy as np
import pytest

from oracle import orc, ref
from tests import matcher_cases as mc
from ydorbslam_b200 import synth

pytestmark = pytest.mark.skipif(not ref.available(), reason="oracle/_ref not built and /root/reference absent")

KITTI = dict(fx=718.856, fy=718.856, cx=607.1928, cy=185.2157, bf=386.1448)


def test_descriptor_distance_is_popcount():
    d = synth.synth_descriptors(300, 3)
    for i in range(0, 299):
        assert ref.descriptor_distance(d[i], d[i + 1]) == orc.hamming(d[i], d[i + 1]) == int(np.unpackbits(d[i] ^ d[i + 1]).sum())
    assert ref.descriptor_distance(d[0], d[0]) == 0
    assert ref.descriptor_distance(np.zeros(32, np.uint8), np.full(32, 255, np.uint8)) == 256


def _stereo_case(w, h, nf, seed, baseline_before):
    left, right = synth.synth_stereo_pair(w, h, seed)
    f = ref.RefFrame.stereo(left, right, n_features=nf, baseline_before=baseline_before, **KITTI)
    kl, dl, kr, dr, rx, dp, bounds = f.get()
    p = orc.make_params(nf, 1.2, 8, 20)
    ok, od, _ = orc.extract(p, left, cap=nf + 64)
    rk, rd, _ = orc.extract(p, right, cap=nf + 64)
    return f, (kl, dl, kr, dr, rx, dp, bounds), p, (ok, od, rk, rd), left, right


@pytest.mark.parametrize("w,h,nf,seed,steady", [(1241, 376, 2000, 64, False), (1241, 376, 2000, 65, True), (640, 480, 1000, 66, False),
                                                (752, 480, 1200, 67, True)])
def test_stereo_frame_reference_equals_oracle(w, h, nf, seed, steady):
    """The reference's stereo Frame constructor == oracle extract x 2 + stereo_match(mode 0) -- rightX and depth bit for
    bit, the leftIdx lag included.  steady: m_flt_baseLine already set by an earlier frame (finite maxD)."""
    base = np.float32(np.float32(KITTI["bf"]) / np.float32(KITTI["fx"]))
    f, (kl, dl, kr, dr, rx, dp, bounds), p, (ok, od, rk, rd), left, right = _stereo_case(w, h, nf, seed, float(base) if steady else 0.0)
    assert np.array_equal(kl, ok) and np.array_equal(dl, od) and np.array_equal(kr, rk) and np.array_equal(dr, rd)
    assert bounds.tolist() == [0.0, 0.0, float(w), float(h)]
    (_, pl), (_, pr) = orc.pyramid(p, left), orc.pyramid(p, right)
    max_d = np.float32(np.float32(KITTI["bf"]) / base) if steady else np.inf
    orx, odp = orc.stereo_match(p, w, h, pl, pr, ok, od, rk, rd, KITTI["bf"], max_d, 0)
    assert np.array_equal(orx.view(np.uint32), rx.view(np.uint32))
    assert np.array_equal(odp.view(np.uint32), dp.view(np.uint32))
    assert (rx >= 0).sum() > 20  # matches exist
    f.close()


def test_keypoints_in_area_reference_equals_oracle():
    """G1 + G2 on the frame the stereo constructor built, and on lattice key points with shifted image bounds."""
    f, (kl, *_), p, (ok, *_), _, _ = _stereo_case(1241, 376, 2000, 64, 0.0)
    g = orc.make_geom(p, 1241, 376)
    rng = np.random.RandomState(1)
    nonempty = 0
    for _ in range(400):
        x, y, r = rng.uniform(-20, 1260), rng.uniform(-20, 400), rng.uniform(1, 90)
        mn, mx = rng.randint(-1, 8), rng.randint(-1, 8)
        a, o = f.keypoints_in_area(x, y, r, mn, mx), orc.keypoints_in_area(g, ok, x, y, r, mn, mx)
        assert np.array_equal(a, o), (x, y, r, mn, mx)
        nonempty += len(a) > 0
    assert nonempty > 100
    f.close()
    kps, _ = mc.synth_keypoints(1500, 9)
    bounds = (-12.5, -7.25, 655.0, 490.5)  # undistorted bounds need not start at 0 (src/frame.cpp:238-246)
    f = ref.RefFrame.from_keypoints(kps, bounds)
    g.min_x, g.min_y, g.max_x, g.max_y = bounds
    for _ in range(400):
        x, y, r = rng.uniform(-30, 680), rng.uniform(-30, 510), rng.uniform(1, 60)
        mn, mx = rng.randint(-1, 8), rng.randint(-1, 8)
        assert np.array_equal(f.keypoints_in_area(x, y, r, mn, mx), orc.keypoints_in_area(g, kps, x, y, r, mn, mx))
    f.close()


@pytest.mark.parametrize("nkp,nmp,th,occupied,seed", [(900, 3000, 3.0, False, 1), (900, 12000, 1.0, True, 2), (400, 500, 5.0, True, 3),
                                                      (40, 2000, 7.0, False, 4), (900, 0, 3.0, False, 5)])
def test_search_by_projection_reference_equals_oracle(nkp, nmp, th, occupied, seed):
    """M1: the reference's own OrbMatcher::searchByProjectionInFrameAndMapPoint (on MapPoint objects built by its own
    constructor) == the oracle's restatement -- assignments per key point and the return value."""
    kps, desc = mc.synth_keypoints(nkp, seed)
    p = orc.make_params(1000, 1.2, 8, 20)
    g = orc.make_geom(p, mc.W, mc.H)
    # points projected just outside the radius in x (the reference keeps |dx| > r); view_cos 0.9 -> factor 4.0
    mp = mc.window_scene(kps, desc, max(nmp, 1), 50 + seed, th * 4.0)
    mps = np.zeros(max(nmp, 1), orc.MPPROJ_DTYPE)
    for f in ("proj_x", "proj_y", "proj_right_x", "level", "active"):
        mps[f][: len(mp[f])] = mp[f]
    mps["view_cos"] = np.float32(0.9)
    mps, mpd, obs = mps[:nmp], mp["desc"][:nmp], mp["obs_positive"][:nmp]
    rng = np.random.RandomState(seed)
    right_x = mp["right_x_kp"]
    occ = (rng.rand(nkp) < 0.1).astype(np.uint8) if occupied else np.zeros(nkp, np.uint8)
    f = ref.RefFrame.from_keypoints(kps, (0, 0, mc.W, mc.H))
    f.set_features(desc, right_x, mc.SF)
    n, match = f.search_by_projection(mps, mpd, obs, occ, th, 0.8)
    on, omatch = orc.search_by_projection(g, kps, desc, right_x, occ, mps, mpd, obs, th, 0.8)
    assert n == on
    assert np.array_equal(match, omatch)
    if nmp >= 3000:
        assert n > 30
    f.close()


@pytest.mark.parametrize("motion,tz", [("forward", 1.0), ("backward", -1.0), ("none", 0.0)])
@pytest.mark.parametrize("check_ori", [True, False])
def test_search_last_and_current_frame_reference_equals_oracle(motion, tz, check_ori):
    """M2 (+ M7): the reference's own searchByProjectionInLastAndCurrentFrame -- projection by its cv::Mat arithmetic,
    level window by camera motion, stereo gate, greedy claims, rotation histogram + computeThreeMaxima -- == the
    oracle's window search on queries projected with the same float32 operations."""
    from ydorbslam_b200 import queries as Q

    kps, desc = mc.synth_keypoints(900, 21)
    s = mc.last_frame_scene(kps, desc, 3000, 31, 7.0)
    f = ref.RefFrame.from_keypoints(kps, (0, 0, mc.W, mc.H))
    f.set_features(desc, s["right_x_kp"], mc.SF)
    cam = [mc.CAM[k] for k in ("fx", "fy", "cx", "cy", "bf", "baseline")]
    n, match = f.search_last_current(s["last_kps"], s["pos"], s["desc"], s["obs_positive"], s["outlier"], (0.0, 0.0, tz), cam,
                                     s["occupied"], 7.0, 0.9, check_ori)
    q, prm = Q.last_and_current_frame(s["proj_x"], s["proj_y"], s["proj_right_x"], s["last_kps"]["octave"], s["last_kps"]["angle"],
                                      s["active"], s["obs_positive"].astype(bool), mc.SF, 7.0, motion, check_orientation=check_ori)
    g = orc.make_geom(orc.make_params(1000, 1.2, 8, 20), mc.W, mc.H)
    on, obest, omatch = orc.search_windows(g, kps, desc, s["right_x_kp"], s["occupied"], q, s["desc"], prm)
    assert n == on
    assert np.array_equal(match, np.where(omatch < 0, -1, omatch))
    assert n > 40
    f.close()


@pytest.mark.parametrize("n1,n2,seed,check_ori", [(800, 900, 3, True), (800, 900, 4, False), (300, 1500, 5, True), (50, 40, 6, True)])
def test_search_by_bow_reference_equals_oracle(n1, n2, seed, check_ori):
    """M5: the reference's own searchByBowInKeyFrameAndFrame (on a KeyFrame built by its constructor, DBoW3 FeatureVector
    std::maps filled through addFeature) == the oracle's CSR restatement."""
    s = mc.bow_scene(n1, n2, seed)
    n, match = ref.search_by_bow_keyframe_frame(s["k1"], s["d1"], s["v1"], s["fv1"], s["k2"], s["d2"], s["fv2"], 0.7, check_ori)
    on, omatch = orc.search_by_bow(0, s["k1"], s["d1"], s["v1"], s["fv1"], s["k2"], s["d2"], None, s["fv2"], 0.7, check_ori)
    assert n == on
    assert np.array_equal(match, omatch)
    if n1 >= 300:
        assert n > 30


def _keyframe_scene(kps, desc, n_kf, seed, thd):
    """Map points of a key frame for M3: world positions projecting where window_scene wants them, invariance distances
    around the actual distance (some out of range), bad / already-found flags."""
    s = mc.last_frame_scene(kps, desc, n_kf, seed, thd)
    rng = np.random.RandomState(seed + 7)
    pos = np.where(np.isnan(s["pos"]), 0.0, s["pos"]).astype(np.float64)
    dist = np.sqrt((pos ** 2).sum(1))
    lvl = s["last_kps"]["octave"].astype(np.int64)
    max_d = (dist * mc.SF[lvl] * rng.uniform(0.95, 1.05, n_kf)).astype(np.float32)
    min_d = (max_d / mc.SF[-1]).astype(np.float32)
    far = rng.rand(n_kf) < 0.05
    max_d[far] *= np.float32(0.3)  # dist3D > 1.2 * maxDistance: dropped by :188
    s.update(min_d=min_d, max_d=max_d, bad=(rng.rand(n_kf) < 0.04).astype(np.uint8), found=(rng.rand(n_kf) < 0.06).astype(np.uint8))
    return s


@pytest.mark.parametrize("check_ori,orb_dist", [(True, 100), (False, 64)])
def test_search_keyframe_and_current_frame_reference_equals_oracle(check_ori, orb_dist):
    """M3: the reference's own searchByProjectionInKeyFrameAndCurrentFrame == the oracle's window search on the queries
    its own methods produce (projection, predictScaleLevel, invariance gate); also pins the +-1 level window, the
    any-slot exclusion (:194) and the non-strict threshold (:202)."""
    from ydorbslam_b200 import queries as Q

    kps, desc = mc.synth_keypoints(900, 23)
    s = _keyframe_scene(kps, desc, 3000, 33, 7.0)
    f = ref.RefFrame.from_keypoints(kps, (0, 0, mc.W, mc.H))
    f.set_features(desc, s["right_x_kp"], mc.SF)
    cam = [mc.CAM[k] for k in ("fx", "fy", "cx", "cy", "bf", "baseline")]
    n, match, dbg = f.search_keyframe_current(s["last_kps"], s["pos"], s["desc"], s["min_d"], s["max_d"], s["bad"], s["found"], cam,
                                              s["occupied"], 7.0, orb_dist, 0.9, check_ori)
    active = dbg[:, 3] > 0
    q, prm = Q.keyframe_and_current_frame(dbg[:, 0], dbg[:, 1], dbg[:, 2].astype(np.int32), s["last_kps"]["angle"], active, mc.SF, 7.0,
                                          orb_dist, check_orientation=check_ori)
    g = orc.make_geom(orc.make_params(1000, 1.2, 8, 20), mc.W, mc.H)
    on, obest, omatch = orc.search_windows(g, kps, desc, s["right_x_kp"], s["occupied"], q, s["desc"], prm)
    assert n == on
    assert np.array_equal(match, np.where(omatch < 0, -1, omatch))
    assert n > 40 and active.sum() > 1000
    f.close()


@pytest.mark.parametrize("n1,n2,seed,stereo_only,check_ori", [(1000, 1100, 3, False, True), (1000, 1100, 4, True, True),
                                                              (600, 400, 5, False, False)])
def test_search_for_triangulation_reference_equals_oracle(n1, n2, seed, stereo_only, check_ori):
    """M6: the reference's own searchForTriangulation (epipole from its cv::Mat arithmetic, epipole-distance gate,
    isEpipolarLineDistCorrect, greedy exclusion, rotation histogram) == the oracle, on the epipole the reference computed."""
    s = mc.bow_scene(n1, n2, seed)
    F, ep = mc.fundamental(seed)
    has1, has2 = 1 - s["v1"], 1 - s["v2"]
    cam = [mc.CAM[k] for k in ("fx", "fy", "cx", "cy", "bf", "baseline")]
    origin1 = [(float(ep[0]) - float(mc.CAM["cx"])) / float(mc.CAM["fx"]), (float(ep[1]) - float(mc.CAM["cy"])) / float(mc.CAM["fy"]), 1.0]
    n, pairs, rep = ref.search_for_triangulation(s["k1"], s["d1"], s["rx1"], has1, s["fv1"], s["k2"], s["d2"], s["rx2"], has2, s["fv2"],
                                                 F, origin1, cam, mc.SF, mc.SF2, stereo_only, 0.6, check_ori)
    assert abs(rep[0] - ep[0]) < 1e-2 and abs(rep[1] - ep[1]) < 1e-2
    on, opairs = orc.search_for_triangulation(s["k1"], s["d1"], s["rx1"], has1, s["fv1"], s["k2"], s["d2"], s["rx2"], has2, s["fv2"],
                                              F, rep, mc.SF, mc.SF2, stereo_only, check_ori)
    assert n == on and np.array_equal(pairs, opairs)
    if n1 >= 1000:
        assert n > 15


@pytest.mark.parametrize("n1,n2,seed,check_ori", [(800, 900, 3, True), (300, 1500, 5, False)])
def test_search_by_bow_two_keyframes_reference_equals_oracle(n1, n2, seed, check_ori):
    """M5 second mode: the reference's own searchByBowInTwoKeyFrames == the oracle (mode 1)."""
    s = mc.bow_scene(n1, n2, seed)
    n, match = ref.search_by_bow_two_keyframes(s["k1"], s["d1"], s["v1"], s["fv1"], s["k2"], s["d2"], s["v2"], s["fv2"], 0.7, check_ori)
    on, omatch = orc.search_by_bow(1, s["k1"], s["d1"], s["v1"], s["fv1"], s["k2"], s["d2"], s["v2"], s["fv2"], 0.7, check_ori)
    assert n == on and np.array_equal(match, omatch)
    assert n > 30


def test_search_by_projection_in_sim_reference_equals_oracle():
    """M4 (first member): the reference's own searchByProjectionInSim (identity similarity) == the oracle's window search
    with the octave gate [level-1, level] (:285-286), the matched-slot exclusion (:284) and the <= lowThd rule (:294)."""
    from ydorbslam_b200 import queries as Q

    kps, desc = mc.synth_keypoints(900, 25)
    s = _keyframe_scene(kps, desc, 3000, 35, 7.0)
    f = ref.RefFrame.from_keypoints(kps, (0, 0, mc.W, mc.H))
    f.set_features(desc, s["right_x_kp"], mc.SF)
    cam = [mc.CAM[k] for k in ("fx", "fy", "cx", "cy", "bf", "baseline")]
    n, match, dbg = f.search_in_sim(s["pos"], s["desc"], s["min_d"], s["max_d"], s["bad"], cam, s["occupied"], 7, 0.6)
    q, prm = Q.in_sim(dbg[:, 0], dbg[:, 1], dbg[:, 2].astype(np.int32), dbg[:, 3] > 0, mc.SF, 7)
    g = orc.make_geom(orc.make_params(1000, 1.2, 8, 20), mc.W, mc.H)
    on, obest, omatch = orc.search_windows(g, kps, desc, s["right_x_kp"], s["occupied"], q, s["desc"], prm)
    assert n == on
    assert np.array_equal(match, np.where(omatch < 0, -1, omatch))
    assert n > 20 and (dbg[:, 3] > 0).sum() > 1000
    f.close()


def test_frustum_reference_equals_oracle():
    """N2: the reference's own Frame::isInCameraFrustum + MapPoint::predictScaleLevel == the oracle, every gate exercised on
    both sides (behind the camera, outside the image, out of the invariance range, oblique normal)."""
    from tests import map_cases

    cam, mp = map_cases.camera_and_points(4000, 5)
    R, t = np.array(cam.R[:], np.float32), np.array(cam.t[:], np.float32)
    base = np.float32(cam.bf) / np.float32(cam.fx)
    got = ref.frustum(R, t, (cam.fx, cam.fy, cam.cx, cam.cy, cam.bf, base), (cam.min_x, cam.min_y, cam.max_x, cam.max_y), 1.2,
                      cam.n_levels, mp, 0.5)
    want = orc.frustum(cam, mp, 0.5)
    assert np.array_equal(got["active"], want["active"])
    a = want["active"] != 0
    assert 200 < a.sum() < len(mp) - 200
    for f in ("proj_x", "proj_y", "proj_right_x", "view_cos"):
        assert np.array_equal(got[f][a].view(np.uint32), want[f][a].view(np.uint32)), f
    assert np.array_equal(got["level"][a], want["level"][a])


def test_distinctive_descriptors_reference_equals_oracle():
    """N3: the reference's own MapPoint::computeDistinctiveDescriptors (observations through addObservation on KeyFrames,
    the float distance matrix, std::sort median at index 0.5 * n) picks the descriptor the oracle picks."""
    from tests import map_cases

    desc, start = map_cases.observation_sets(400, 11)
    got = ref.distinctive_descriptors(desc, start)
    best = orc.distinctive_descriptors(desc, start)
    for p in range(len(best)):
        if best[p] < 0:
            assert not got[p].any(), p  # no observation: the MapPoint keeps what it had
        else:
            assert np.array_equal(got[p], desc[start[p] + best[p]]), p
    assert (best >= 0).sum() > 300


def _sim3_sides(seed):
    """Two key frames at the identity pose seeing the same scene: about 60% of frame 2's key points are frame 1's moved
    to just outside the window radius with a few descriptor bits flipped, the rest are unrelated.  Every key point owns a map point on its
    own viewing ray, so both directions find candidates and most related pairs agree (:669-679); some do not (heavy
    descriptor noise, bad points, invariance distance out of range, crowded windows)."""
    rng = np.random.RandomState(seed)
    k1, d1 = mc.synth_keypoints(700, seed)
    ke, de = mc.synth_keypoints(800, seed + 1)
    k2, d2 = ke.copy(), de.copy()
    twin = rng.permutation(800)[:420]  # frame-2 slots that copy a frame-1 key point
    src = rng.permutation(700)[:420]
    k2[twin] = k1[src]
    rad = 7.5 * mc.SF[k1["octave"][src].astype(np.int64)]  # a candidate lies OUTSIDE the radius in x, inside in y
    k2["x"][twin] += (np.where(rng.rand(420) < 0.5, 1.0, -1.0) * (rad + 0.05 + rng.uniform(0, 6, 420))).astype(np.float32)
    k2["y"][twin] += (rng.uniform(-0.9, 0.9, 420) * rad).astype(np.float32)
    k2["x"][twin] = np.clip(k2["x"][twin], 1, mc.W - 2)
    k2["y"][twin] = np.clip(k2["y"][twin], 1, mc.H - 2)
    d2[twin] = d1[src]
    flips = rng.randint(0, 256, (420, 40))
    nflip = np.where(rng.rand(420) < 0.15, 40, rng.randint(0, 14, 420))
    for r in range(420):
        for b in flips[r, : nflip[r]]:
            d2[twin[r], b >> 3] ^= np.uint8(1 << (b & 7))

    def side(k, d, sd):
        r = np.random.RandomState(sd)
        n = len(k)
        z = r.uniform(2.0, 9.0, n)
        pos = np.stack([(k["x"] - mc.CAM["cx"]) / mc.CAM["fx"] * z, (k["y"] - mc.CAM["cy"]) / mc.CAM["fy"] * z, z], 1).astype(np.float32)
        dist = np.sqrt((pos.astype(np.float64) ** 2).sum(1))
        max_d = (dist * mc.SF[k["octave"].astype(np.int64)] * r.uniform(0.93, 1.0, n)).astype(np.float32)  # predicted level == octave, mostly
        min_d = (max_d / mc.SF[-1]).astype(np.float32)
        max_d[r.rand(n) < 0.04] *= np.float32(0.3)
        pos[r.rand(n) < 0.05, 0] = np.nan  # no map point in this slot
        return dict(kps=k, desc=d, pos=pos, mp_desc=d.copy(), min_d=min_d, max_d=max_d, bad=(r.rand(n) < 0.04).astype(np.uint8))

    return side(k1, d1, seed + 2), side(k2, d2, seed + 3)


def test_search_by_sim3_reference_equals_oracle():
    """M4 (second member): the reference's own searchBySim3 -- both directions, the pre-matched exclusions (:580-588) and
    the agreement check (:669-679) -- == two oracle window searches + sim3_agreement."""
    from ydorbslam_b200 import queries as Q

    side1, side2 = _sim3_sides(41)
    n1, n2 = len(side1["kps"]), len(side2["kps"])
    rng = np.random.RandomState(3)
    pre = np.where(rng.rand(n1) < 0.05, rng.randint(0, n2, n1), -1).astype(np.int32)
    cam = [mc.CAM[k] for k in ("fx", "fy", "cx", "cy", "bf", "baseline")]
    n, match, dbg1, dbg2 = ref.search_by_sim3(side1, side2, cam, (0, 0, mc.W, mc.H), mc.SF, 7.5, pre)
    g = orc.make_geom(orc.make_params(1000, 1.2, 8, 20), mc.W, mc.H)
    act1 = (dbg1[:, 3] > 0) & (pre < 0)
    blocked2 = np.zeros(n2, bool)
    blocked2[pre[pre >= 0]] = True
    act2 = (dbg2[:, 3] > 0) & ~blocked2
    q1, p1 = Q.sim3_direction(dbg1[:, 0], dbg1[:, 1], dbg1[:, 2].astype(np.int32), act1, mc.SF, 7.5)
    q2, p2 = Q.sim3_direction(dbg2[:, 0], dbg2[:, 1], dbg2[:, 2].astype(np.int32), act2, mc.SF, 7.5)
    none1, none2 = np.full(n1, -1, np.float32), np.full(n2, -1, np.float32)
    _, best12, _ = orc.search_windows(g, side2["kps"], side2["desc"], none2, None, q1, side1["mp_desc"], p1)
    _, best21, _ = orc.search_windows(g, side1["kps"], side1["desc"], none1, None, q2, side2["mp_desc"], p2)
    found, out = Q.sim3_agreement(best12, best21)
    assert n == found
    assert np.array_equal(np.where(match == -2, -1, match), out)
    assert (match == -2).sum() == (pre >= 0).sum()
    assert found > 150 and (best12 >= 0).sum() > found + 20


def fuse_scene(mode, seed, thd, sim_t=(0.0, 0.0, 0.0)):
    """A key frame (some key points already holding good or bad map points) and candidate map points whose projections
    land where window_scene wants them; about three candidates per key point, so replace chains occur (:728-735)."""
    kps, desc = mc.synth_keypoints(900, seed)
    nmp = 3000
    ws = mc.window_scene(kps, desc, nmp, seed + 1, thd, dx_extra=1.0 if mode == 0 else 8.0)
    rng = np.random.RandomState(seed + 2)
    z = (float(mc.CAM["bf"]) / 10.0 * rng.uniform(0.98, 1.02, nmp)).astype(np.float32)  # projRightX ~ projX - 10, as the scene's rightX
    z[rng.rand(nmp) < 0.03] *= -1
    pc = np.stack([(ws["proj_x"].astype(np.float64) - float(mc.CAM["cx"])) / float(mc.CAM["fx"]) * z,
                   (ws["proj_y"].astype(np.float64) - float(mc.CAM["cy"])) / float(mc.CAM["fy"]) * z, z], 1)
    pos = (pc - np.asarray(sim_t, np.float64)).astype(np.float32)  # camera = world + t
    dist = np.sqrt((pc ** 2).sum(1))
    max_d = (dist * mc.SF[ws["level"]] * rng.uniform(0.93, 1.0, nmp)).astype(np.float32)
    min_d = (max_d / mc.SF[-1]).astype(np.float32)
    max_d[rng.rand(nmp) < 0.04] *= np.float32(0.3)
    normal = (pc / dist[:, None]).astype(np.float32)
    normal[rng.rand(nmp) < 0.05] *= -1  # seen from behind: :706 / :774 drop it
    pos[rng.rand(nmp) < 0.03, 0] = np.nan
    points = dict(pos=pos, desc=ws["desc"], min_d=min_d, max_d=max_d, normal=normal, bad=(rng.rand(nmp) < 0.04).astype(np.uint8),
                  obs=rng.randint(0, 6, nmp).astype(np.int32))
    u = rng.rand(900)
    slot_kind = np.where(u < 0.4, 0, np.where(u < 0.92, 1, 2)).astype(np.uint8)
    return dict(kps=kps, desc=desc, right_x=ws["right_x_kp"], slot_kind=slot_kind, slot_obs=rng.randint(0, 6, 900).astype(np.int32),
                points=points)


def replay_fuse(mode, best, s):
    """What the caller of the candidate search does with best[j] (src/orbMatcher.cpp:728-739 / :791-800 and
    MapPoint::beReplacedBy, src/mapPoint.cpp:128-156): entity ids as ref.fuse names them."""
    n, m = len(s["kps"]), len(best)
    step = np.where(s["right_x"] >= 0, 2, 1)  # MapPoint::addObservation, src/mapPoint.cpp:96-100
    occupant = np.where(s["slot_kind"] > 0, np.arange(n), -1)
    obs = np.concatenate([s["slot_obs"] + np.where(s["slot_kind"] > 0, step, 0), s["points"]["obs"]]).astype(np.int64)
    bad = np.concatenate([s["slot_kind"] == 2, s["points"]["bad"] > 0])
    where = np.concatenate([np.where(s["slot_kind"] > 0, np.arange(n), -1), np.full(m, -1)])  # key point the entity is observed at
    repl = np.full(n + m, -1)
    fused = 0
    for j in range(m):
        idx, me = int(best[j]), n + j
        if idx < 0:
            continue
        fused += 1
        o = int(occupant[idx])
        if o < 0:
            where[me], occupant[idx] = idx, me
            obs[me] += step[idx]
            continue
        if bad[o]:
            continue
        loser, winner = (me, o) if (mode == 1 or obs[o] > obs[me]) else (o, me)
        bad[loser], repl[loser] = True, winner
        if where[loser] >= 0:
            at = int(where[loser])
            if where[winner] < 0:
                occupant[at], where[winner] = winner, at
                obs[winner] += step[at]
            else:
                occupant[at] = -1
            where[loser] = -1
    return dict(fused=fused, slot=occupant, slot_repl=repl[:n], mp_repl=repl[n:], slot_obs=np.where(s["slot_kind"] > 0, obs[:n], 0),
                mp_obs=obs[n:], mp_bad=bad[n:])


def fuse_queries(mode, r, s, thd):
    from ydorbslam_b200 import queries as Q

    dbg = r["dbg"]
    active, level = dbg[:, 4] > 0, dbg[:, 3].astype(np.int32)
    if mode == 0:
        return Q.fuse_by_projection(dbg[:, 0], dbg[:, 1], dbg[:, 2], level, active, mc.SF, mc.INV_SF2, thd)
    return Q.fuse_by_sim3(dbg[:, 0], dbg[:, 1], level, active, mc.SF, thd)


def check_fuse(r, mine, s):
    assert r["fused"] == mine["fused"]
    for key in ("slot", "slot_repl", "mp_repl", "slot_obs"):
        assert np.array_equal(r[key], mine[key]), key
    live = ~np.isnan(s["points"]["pos"][:, 0])  # null entries of the list have no state
    assert np.array_equal(r["mp_obs"][live], mine["mp_obs"][live])
    assert np.array_equal(r["mp_bad"][live] > 0, mine["mp_bad"][live])


@pytest.mark.parametrize("mode,thd,sim_t", [(0, 1.2, (0, 0, 0)), (1, 4.0, (0, 0, 0)), (1, 4.0, (0.3, -0.2, 0.5))])
def test_fuse_reference_equals_oracle(mode, thd, sim_t):
    """M4 (third and fourth members): the reference's own fuseByProjection / fuseBySim3 -- projection, gates, window,
    chi-square test, best candidate, and the map mutation that follows -- == the oracle's candidate search + the
    caller-side replay, on every observable: occupant per key point, replacements, observation counts, bad flags."""
    s = fuse_scene(mode, 61 + mode, thd, sim_t)
    cam = [mc.CAM[k] for k in ("fx", "fy", "cx", "cy", "bf", "baseline")]
    r = ref.fuse(mode, s["kps"], s["desc"], s["right_x"], s["slot_kind"], s["slot_obs"], s["points"], cam, (0, 0, mc.W, mc.H), mc.SF, thd,
                 2.0 if mode else 1.0, sim_t)
    q, prm = fuse_queries(mode, r, s, thd)
    g = orc.make_geom(orc.make_params(1000, 1.2, 8, 20), mc.W, mc.H)
    _, best, _ = orc.search_windows(g, s["kps"], s["desc"], s["right_x"], None, q, s["points"]["desc"], prm)
    mine = replay_fuse(mode, best, s)
    check_fuse(r, mine, s)
    assert r["fused"] > (100 if mode else 30)
    assert (r["mp_repl"] >= 0).sum() > 10
    assert ((r["slot_repl"] >= 0).sum() > 5) == (mode == 0)  # only fuseByProjection ever replaces the key frame's own point
