"""GPU parity: M2/M3/M4 window searches, M5 BoW matchers, M6 searchForTriangulation, M7 computeThreeMaxima vs the CPU
oracle -- match indices and counts bit-exact.  All calls go through the C ABI (libydorb.so)."""
import numpy as np
import pytest

from oracle import orc
from tests import matcher_cases as mc
from ydorbslam_b200 import OrbExtractor, OrbMatcher, synth
from ydorbslam_b200 import queries as Q

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def extracted():
    img = synth.synth_frame(640, 480, 43)
    ex = OrbExtractor(1000, 1.2, 8, 20, 7, max_width=640, max_height=480)
    kps, desc = ex.extractAndCompute(img)
    ex.close()
    return kps, desc


def _geom():
    return orc.make_geom(orc.make_params(1000, 1.2, 8, 20), mc.W, mc.H)


@pytest.mark.parametrize("source", ["extracted", "lattice"])
@pytest.mark.parametrize("nmp,seed", [(3000, 11), (12000, 12), (1, 13), (0, 14)])
def test_window_searches(extracted, source, nmp, seed):
    kps, desc = extracted if source == "extracted" else mc.synth_keypoints(900, seed)
    g = _geom()
    m = OrbMatcher(0.8)
    total = 0
    for name, q, prm, occ, mp in mc.all_window_cases(kps, desc, nmp, seed):
        fr = m.make_frame(kps, desc, mp["right_x_kp"], bounds=(0, 0, mc.W, mc.H), scale_factors=mc.SF)
        n, best, match = m.searchWindows(fr, q, mp["desc"], prm, occ)
        rn, rbest, rmatch = orc.search_windows(g, kps, desc, mp["right_x_kp"], occ, q, mp["desc"], prm)
        assert n == rn, name
        assert np.array_equal(best, rbest), name
        assert np.array_equal(match, rmatch), name
        total += n
        fr.close()
    if nmp >= 3000:
        assert total > 500
    m.close()


def test_window_greedy_chain():
    """Many queries fighting for few key points, every claim blocking: the exclusion chain is as long as it gets."""
    kps, desc = mc.synth_keypoints(40, 5)
    g = _geom()
    m = OrbMatcher()
    mp = mc.window_scene(kps, desc, 5000, 9, 7.0)
    q, prm = Q.keyframe_and_current_frame(mp["proj_x"], mp["proj_y"], np.maximum(mp["level"] - 1, 0), mp["angle"],
                                          np.ones(5000, bool), mc.SF, 7.0, 256)
    fr = m.make_frame(kps, desc, None, bounds=(0, 0, mc.W, mc.H), scale_factors=mc.SF)
    n, best, match = m.searchWindows(fr, q, mp["desc"], prm)
    rn, rbest, rmatch = orc.search_windows(g, kps, desc, np.full(len(kps), -1, np.float32), None, q, mp["desc"], prm)
    assert n == rn and np.array_equal(best, rbest) and np.array_equal(match, rmatch)
    assert (rbest >= 0).sum() >= 10
    fr.close(); m.close()


def test_named_members(extracted):
    """The reference-named wrappers (host mirror) agree with the generic entry point + oracle."""
    kps, desc = extracted
    g = _geom()
    m = OrbMatcher(0.9, True)
    mp = mc.window_scene(kps, desc, 2000, 21, 7.0)
    a = mp["active"].astype(bool)
    fr = m.make_frame(kps, desc, mp["right_x_kp"], bounds=(0, 0, mc.W, mc.H), scale_factors=mc.SF)
    n, match = m.searchByProjectionInLastAndCurrentFrame(fr, mp["proj_x"], mp["proj_y"], mp["proj_right_x"], mp["level"],
                                                         mp["angle"], a, mp["obs_positive"].astype(bool), mp["desc"], mc.SF,
                                                         7.0, "forward", mp["occupied"])
    q, prm = Q.last_and_current_frame(mp["proj_x"], mp["proj_y"], mp["proj_right_x"], mp["level"], mp["angle"], a,
                                      mp["obs_positive"].astype(bool), mc.SF, 7.0, "forward")
    rn, _, rmatch = orc.search_windows(g, kps, desc, mp["right_x_kp"], mp["occupied"], q, mp["desc"], prm)
    assert n == rn and np.array_equal(match, rmatch) and n > 20
    # searchBySim3: two directions + agreement (:669-679)
    kps2, desc2 = mc.synth_keypoints(800, 77)
    mp2 = mc.window_scene(kps2, desc2, 1500, 22, 7.0)
    fr2 = m.make_frame(kps2, desc2, None, bounds=(0, 0, mc.W, mc.H), scale_factors=mc.SF)
    # first key frame's points projected into the second and vice versa; indices wrap so that agreement can happen
    f2s = (mp2["proj_x"], mp2["proj_y"], mp2["level"], mp2["active"].astype(bool), mp2["desc"])
    s2f = (mp["proj_x"][:800], mp["proj_y"][:800], mp["level"][:800], a[:800], mp["desc"][:800])
    found, pairs = m.searchBySim3(fr, fr2, f2s, s2f, mc.SF, 7.0)
    q12, p12 = Q.sim3_direction(*f2s[:4], mc.SF, 7.0)
    q21, p21 = Q.sim3_direction(*s2f[:4], mc.SF, 7.0)
    _, b12, _ = orc.search_windows(g, kps2, desc2, np.full(len(kps2), -1, np.float32), None, q12, f2s[4], p12)
    _, b21, _ = orc.search_windows(g, kps, desc, mp["right_x_kp"], None, q21, s2f[4], p21)
    rfound, rpairs = Q.sim3_agreement(b12, b21)
    assert found == rfound and np.array_equal(pairs, rpairs)
    fr.close(); fr2.close(); m.close()


@pytest.mark.parametrize("sizes", [[0] * 30, [100, 9, 8] + [0] * 27, [100, 10, 9] + [0] * 27, [3, 3, 3, 3] + [0] * 26,
                                   list(range(30)), list(range(30, 0, -1))])
def test_three_maxima(sizes):
    assert OrbMatcher.computeThreeMaxima(sizes) == orc.three_maxima(sizes)


@pytest.mark.parametrize("n1,n2,seed", [(1000, 1000, 5), (2000, 1700, 6), (300, 4000, 7), (1, 1, 8), (0, 50, 9)])
@pytest.mark.parametrize("mode", [0, 1])
@pytest.mark.parametrize("check_ori", [True, False])
def test_search_by_bow(n1, n2, seed, mode, check_ori):
    s = mc.bow_scene(n1, n2, seed)
    m = OrbMatcher(0.75, check_ori)
    if mode == 0:
        n, out = m.searchByBowInKeyFrameAndFrame(s["k1"], s["d1"], s["v1"], s["fv1"], s["k2"], s["d2"], s["fv2"])
    else:
        n, out = m.searchByBowInTwoKeyFrames(s["k1"], s["d1"], s["v1"], s["fv1"], s["k2"], s["d2"], s["v2"], s["fv2"])
    rn, rout = orc.search_by_bow(mode, s["k1"], s["d1"], s["v1"], s["fv1"], s["k2"], s["d2"], s["v2"], s["fv2"], 0.75, check_ori)
    assert n == rn and np.array_equal(out, rout)
    if min(n1, n2) >= 1000:
        assert n > 100
    m.close()


def test_bow_single_node_ties():
    """Everything in ONE node with many duplicate descriptors: exercises tie order (first wins) and the in-node
    exclusion chain at its longest."""
    s = mc.bow_scene(400, 400, 15)
    s["d2"][::3] = s["d2"][0]
    s["d1"][::4] = s["d2"][0]
    for side, n in (("fv1", 400), ("fv2", 400)):
        s[side] = (np.array([7], np.int32), np.array([0, n], np.int32), np.arange(n, dtype=np.int32))
    for mode in (0, 1):
        m = OrbMatcher(1.1, True)  # ratio > 1 lets exact duplicates (best == second) through? no: 0 < 1.1*0 is false
        n, out = (m.searchByBowInKeyFrameAndFrame(s["k1"], s["d1"], s["v1"], s["fv1"], s["k2"], s["d2"], s["fv2"]) if mode == 0
                  else m.searchByBowInTwoKeyFrames(s["k1"], s["d1"], s["v1"], s["fv1"], s["k2"], s["d2"], s["v2"], s["fv2"]))
        rn, rout = orc.search_by_bow(mode, s["k1"], s["d1"], s["v1"], s["fv1"], s["k2"], s["d2"], s["v2"], s["fv2"], 1.1, True)
        assert n == rn and np.array_equal(out, rout)
        m.close()


@pytest.mark.parametrize("n1,n2,seed", [(1000, 1000, 25), (2000, 1500, 26), (0, 10, 27)])
@pytest.mark.parametrize("stereo_only", [False, True])
@pytest.mark.parametrize("check_ori", [True, False])
def test_search_for_triangulation(n1, n2, seed, stereo_only, check_ori):
    s = mc.bow_scene(n1, n2, seed)
    F, ep = mc.fundamental(seed)
    has1, has2 = 1 - s["v1"], 1 - s["v2"]
    m = OrbMatcher(0.6, check_ori)
    n, pairs = m.searchForTriangulation(s["k1"], s["d1"], s["rx1"], has1, s["fv1"], s["k2"], s["d2"], s["rx2"], has2, s["fv2"],
                                        F, ep, mc.SF, mc.SF2, stereo_only)
    rn, rpairs = orc.search_for_triangulation(s["k1"], s["d1"], s["rx1"], has1, s["fv1"], s["k2"], s["d2"], s["rx2"], has2,
                                              s["fv2"], F, ep, mc.SF, mc.SF2, stereo_only, check_ori)
    assert n == rn and np.array_equal(pairs, rpairs)
    if n1 >= 1000:
        assert n > 30
    m.close()


def test_triangulation_gates_both_ways():
    """With no stereo on either side every candidate goes through the epipole-distance and epipolar-line gates: both
    outcomes must occur or the test scene is not testing them."""
    s = mc.bow_scene(1500, 1500, 31)
    F, ep = mc.fundamental(31)
    none1, none2 = np.full(1500, -1, np.float32), np.full(1500, -1, np.float32)
    zeros = np.zeros(1500, np.uint8)
    m = OrbMatcher(0.6, False)
    n, pairs = m.searchForTriangulation(s["k1"], s["d1"], none1, zeros, s["fv1"], s["k2"], s["d2"], none2, zeros, s["fv2"], F, ep,
                                        mc.SF, mc.SF2, False)
    rn, rpairs = orc.search_for_triangulation(s["k1"], s["d1"], none1, zeros, s["fv1"], s["k2"], s["d2"], none2, zeros, s["fv2"],
                                              F, ep, mc.SF, mc.SF2, False, False)
    assert n == rn and np.array_equal(pairs, rpairs)
    # same scene with a gate-free F (huge scale => lhs tiny => always inside the line gate) accepts strictly more
    n_free, _ = orc.search_for_triangulation(s["k1"], s["d1"], none1 + 1.0, zeros, s["fv1"], s["k2"], s["d2"], none2 + 1.0,
                                             zeros, s["fv2"], F * 1e4, ep, mc.SF, mc.SF2, False, False)
    assert 0 < n < n_free
    m.close()


@pytest.mark.parametrize("area_mode", [0, 1])
@pytest.mark.parametrize("clusters,noise,seed,check_ori", [(150, 200, 7, True), (600, 800, 9, True), (600, 100, 10, False),
                                                         (40, 0, 11, True), (0, 50, 12, True)])
def test_search_for_initialization(area_mode, clusters, noise, seed, check_ori):
    """EXTRA `ydorb_match_windowed_top2` (the reference has no searchForInitialization): CUDA fixed-point
    rounds == the sequential restatement in the oracle -- matches, count and updated window centres."""
    kps1, d1, kps2, d2 = mc.init_scene(clusters, noise, seed, area_mode)
    g = _geom()
    prev = np.stack([kps1["x"], kps1["y"]], 1).astype(np.float32) if len(kps1) else np.zeros((0, 2), np.float32)
    rn, rm12, rprev = orc.search_for_initialization(g, kps1, d1.reshape(-1, 32), kps2, d2, prev, 30, 0.9, 50, check_ori, area_mode)
    m = OrbMatcher(0.9, check_ori)
    fr = m.make_frame(kps2, d2, None, bounds=(0, 0, mc.W, mc.H), scale_factors=mc.SF)
    n, m12, pout = m.searchForInitialization(fr, kps1, d1.reshape(-1, 32), prev, window_size=30, area_mode=area_mode)
    assert n == rn
    assert np.array_equal(m12, rm12)
    assert np.array_equal(pout, rprev)
    if clusters >= 150:
        assert n > clusters // 3
    fr.close()
    m.close()
