"""GPU parity against the REFERENCE's own code: CUDA (through the C ABI) == oracle/_ref (orbExtractor.cpp compiled
unmodified, see tests/test_oracle_ref.py) on the BASELINE config shapes, and the device trigonometry == host libm sincosf
over the whole input domain.  /root/reference is not needed at run time: the prebuilt oracle/_ref travels with the repo."""
import ctypes as C
import os
import struct

import numpy as np
import pytest

from oracle import orc, ref
from ydorbslam_b200 import OrbExtractor, synth
from ydorbslam_b200._lib import check, lib

pytestmark = pytest.mark.gpu


def _same(kps, desc, rk, rd):
    assert len(kps) == len(rk)
    for f in orc.KP_DTYPE.names:
        assert np.array_equal(kps[f].view(np.uint32), rk[f].view(np.uint32)), f
    assert np.array_equal(desc, rd)


@pytest.mark.skipif(not ref.available(), reason="oracle/_ref not shipped")
@pytest.mark.parametrize("shape,nf,seed", [((480, 640), 1000, 31), ((376, 1241), 2000, 32), ((1080, 1920), 4000, 33),
                                           ((2160, 3840), 8000, 34), ((97, 97), 1000, 4)])
def test_cuda_matches_compiled_reference(shape, nf, seed):
    """Key points, order, octaves, angles and descriptors bit-exact against the reference build (configs 1, 2, 3, 5)."""
    h, w = shape
    img = synth.synth_frame(w, h, seed)
    rk, rd, _ = ref.extract(img, nf, cap=nf + 64)
    ex = OrbExtractor(nf, 1.2, 8, 20, 7, max_width=w, max_height=h, max_batch=1)
    kps, desc = ex.extractAndCompute(img)
    _same(kps, desc, rk, rd)
    ex.close()


@pytest.mark.skipif(not ref.available(), reason="oracle/_ref not shipped")
def test_cuda_batch_matches_compiled_reference():
    imgs = synth.synth_batch(640, 480, 6, seed0=300)
    imgs[4] = synth.synth_sparse_frame(640, 480, 12)
    ex = OrbExtractor(1000, 1.2, 8, 20, 7, max_width=640, max_height=480, max_batch=6)
    kps, desc, n = ex.extract_batch(imgs)
    for b in range(6):
        rk, rd, _ = ref.extract(imgs[b], 1000)
        _same(kps[b, : n[b]], desc[b, : n[b]], rk, rd)
    ex.close()


def test_device_sincosf_equals_libm_everywhere():
    """src/orbExtractor.cpp:426 is libm sincosf on the host.  glibc_sincosf (csrc/common.cuh) must return the same bits
    for every float an angle in radians can be: all of [0, 2*pi], 1.09e9 values."""
    hi = struct.unpack("<I", struct.pack("<f", 6.2831855))[0]
    chunk = 1 << 25
    l = orc.lib()
    l.orc_sincosf_range.argtypes = [C.c_uint32, C.c_uint32, C.c_void_p, C.c_int]
    dev = np.empty(2 * chunk, np.float32)
    host = np.empty(2 * chunk, np.float32)
    bad = 0
    for first in range(0, hi + 1, chunk):
        cnt = min(chunk, hi + 1 - first)
        check(lib().ydorb_diag_sincosf(0, C.c_uint32(first), C.c_uint32(cnt), dev.ctypes.data_as(C.c_void_p)))
        l.orc_sincosf_range(first, cnt, host.ctypes.data_as(C.c_void_p), os.cpu_count() or 8)
        bad += int((dev[: 2 * cnt].view(np.uint32) != host[: 2 * cnt].view(np.uint32)).sum())
    assert bad == 0, f"{bad} of {2 * (hi + 1)} sin/cos values differ from libm"


KITTI = dict(fx=718.856, fy=718.856, cx=607.1928, cy=185.2157, bf=386.1448)


@pytest.mark.skipif(not ref.available(), reason="oracle/_ref not shipped")
@pytest.mark.parametrize("w,h,nf,seed,steady", [(1241, 376, 2000, 64, False), (1241, 376, 2000, 65, True), (640, 480, 1000, 66, False)])
def test_cuda_stereo_frame_matches_compiled_reference(w, h, nf, seed, steady):
    """CUDA extract x 2 + computeStereoMatches + frame grid == the reference's own stereo Frame constructor
    (src/frame.cpp:57-110 compiled unmodified): key points, descriptors, rightX, depth, getKeyPointsInArea."""
    from ydorbslam_b200 import OrbMatcher

    left, right = synth.synth_stereo_pair(w, h, seed)
    base = np.float32(np.float32(KITTI["bf"]) / np.float32(KITTI["fx"]))
    f = ref.RefFrame.stereo(left, right, n_features=nf, baseline_before=float(base) if steady else 0.0, **KITTI)
    rkl, rdl, rkr, rdr, rrx, rdp, bounds = f.get()
    ex = OrbExtractor(nf, 1.2, 8, 20, 7, max_width=w, max_height=h, max_batch=2)
    kps, desc, n = ex.extract_batch(np.stack([left, right]))
    kl, dl, kr, dr = kps[0, : n[0]], desc[0, : n[0]], kps[1, : n[1]], desc[1, : n[1]]
    _same(kl, dl, rkl, rdl)
    _same(kr, dr, rkr, rdr)
    m = OrbMatcher()
    max_d = np.float32(np.float32(KITTI["bf"]) / base) if steady else np.inf
    rx, dp = m.computeStereoMatches(ex, 0, ex, 1, kl, dl, kr, dr, KITTI["bf"], max_d, 0)
    assert np.array_equal(rx.view(np.uint32), rrx.view(np.uint32))
    assert np.array_equal(dp.view(np.uint32), rdp.view(np.uint32))
    assert (rx >= 0).sum() > 20
    sf = np.array([np.float32(pow(np.float32(1.2), l)) for l in range(8)], np.float32)
    fr = m.make_frame(kl, dl, rx, bounds=tuple(float(v) for v in bounds), scale_factors=sf)
    rng = np.random.RandomState(seed)
    for _ in range(200):
        x, y, r = rng.uniform(-20, w + 20), rng.uniform(-20, h + 20), rng.uniform(1, 90)
        mn, mx = rng.randint(-1, 8), rng.randint(-1, 8)
        assert np.array_equal(fr.getKeyPointsInArea(x, y, r, mn, mx), f.keypoints_in_area(x, y, r, mn, mx))
    fr.close(); m.close(); ex.close(); f.close()


@pytest.mark.skipif(not ref.available(), reason="oracle/_ref not shipped")
def test_cuda_hamming_matches_compiled_reference():
    """M0: the device distance (through ydorb_hamming_top2 on 1-row databases) == OrbMatcher::computeDescriptorsDistance."""
    from ydorbslam_b200 import OrbMatcher

    d = synth.synth_descriptors(400, 5)
    m = OrbMatcher()
    t = m.top2(d[:200], d[200:])
    for i in range(200):
        j = int(t["best_idx"][i])
        assert int(t["best_dist"][i]) == ref.descriptor_distance(d[i], d[200 + j])
        assert m.computeDescriptorsDistance(d[i], d[200 + i]) == ref.descriptor_distance(d[i], d[200 + i])
    m.close()


@pytest.mark.skipif(not ref.available(), reason="oracle/_ref not shipped")
@pytest.mark.parametrize("nkp,nmp,th,occupied,seed", [(900, 3000, 3.0, False, 1), (900, 12000, 1.0, True, 2), (400, 500, 5.0, True, 3)])
def test_cuda_search_by_projection_matches_compiled_reference(nkp, nmp, th, occupied, seed):
    """M1: CUDA == the reference's own OrbMatcher::searchByProjectionInFrameAndMapPoint (src/orbMatcher.cpp:24-64,
    compiled unmodified, MapPoint objects from its own constructor)."""
    from tests import matcher_cases as mc
    from ydorbslam_b200 import MPPROJ_DTYPE, OrbMatcher

    kps, desc = mc.synth_keypoints(nkp, seed)
    mp = mc.window_scene(kps, desc, nmp, 50 + seed, th * 4.0)
    mps = np.zeros(nmp, MPPROJ_DTYPE)
    for f in ("proj_x", "proj_y", "proj_right_x", "level", "active"):
        mps[f] = mp[f]
    mps["view_cos"] = np.float32(0.9)
    rng = np.random.RandomState(seed)
    occ = (rng.rand(nkp) < 0.1).astype(np.uint8) if occupied else np.zeros(nkp, np.uint8)
    f = ref.RefFrame.from_keypoints(kps, (0, 0, mc.W, mc.H))
    f.set_features(desc, mp["right_x_kp"], mc.SF)
    rn, rmatch = f.search_by_projection(mps, mp["desc"], mp["obs_positive"], occ, th, 0.8)
    m = OrbMatcher(0.8)
    fr = m.make_frame(kps, desc, mp["right_x_kp"], bounds=(0, 0, mc.W, mc.H), scale_factors=mc.SF)
    n, match = m.searchByProjectionInFrameAndMapPoint(fr, mps, mp["desc"], mp["obs_positive"], th, kp_occupied=occ if occupied else None)
    assert n == rn and np.array_equal(match, rmatch)
    assert n > 30
    fr.close(); m.close(); f.close()


@pytest.mark.skipif(not ref.available(), reason="oracle/_ref not shipped")
@pytest.mark.parametrize("motion,tz", [("forward", 1.0), ("backward", -1.0), ("none", 0.0)])
def test_cuda_last_and_current_frame_matches_compiled_reference(motion, tz):
    """M2 + M7: CUDA window search == the reference's own searchByProjectionInLastAndCurrentFrame (src/orbMatcher.cpp:65-155
    compiled unmodified: its cv::Mat projection, level window by motion, stereo gate, claims, rotation histogram)."""
    from tests import matcher_cases as mc
    from ydorbslam_b200 import OrbMatcher
    from ydorbslam_b200 import queries as Q

    kps, desc = mc.synth_keypoints(900, 21)
    s = mc.last_frame_scene(kps, desc, 3000, 31, 7.0)
    f = ref.RefFrame.from_keypoints(kps, (0, 0, mc.W, mc.H))
    f.set_features(desc, s["right_x_kp"], mc.SF)
    cam = [mc.CAM[k] for k in ("fx", "fy", "cx", "cy", "bf", "baseline")]
    rn, rmatch = f.search_last_current(s["last_kps"], s["pos"], s["desc"], s["obs_positive"], s["outlier"], (0.0, 0.0, tz), cam,
                                       s["occupied"], 7.0, 0.9, True)
    q, prm = Q.last_and_current_frame(s["proj_x"], s["proj_y"], s["proj_right_x"], s["last_kps"]["octave"], s["last_kps"]["angle"],
                                      s["active"], s["obs_positive"].astype(bool), mc.SF, 7.0, motion)
    m = OrbMatcher(0.9)
    fr = m.make_frame(kps, desc, s["right_x_kp"], bounds=(0, 0, mc.W, mc.H), scale_factors=mc.SF)
    n, best, match = m.searchWindows(fr, q, s["desc"], prm, s["occupied"])
    assert n == rn and np.array_equal(np.where(match < 0, -1, match), rmatch)
    assert n > 40
    fr.close(); m.close(); f.close()


@pytest.mark.skipif(not ref.available(), reason="oracle/_ref not shipped")
@pytest.mark.parametrize("n1,n2,seed", [(800, 900, 3), (300, 1500, 5)])
def test_cuda_search_by_bow_matches_compiled_reference(n1, n2, seed):
    """M5: CUDA searchByBowInKeyFrameAndFrame == the reference's own (src/orbMatcher.cpp:303-379 compiled unmodified)."""
    from tests import matcher_cases as mc
    from ydorbslam_b200 import OrbMatcher

    s = mc.bow_scene(n1, n2, seed)
    rn, rmatch = ref.search_by_bow_keyframe_frame(s["k1"], s["d1"], s["v1"], s["fv1"], s["k2"], s["d2"], s["fv2"], 0.7, True)
    m = OrbMatcher(0.7)
    n, match = m.searchByBowInKeyFrameAndFrame(s["k1"], s["d1"], s["v1"], s["fv1"], s["k2"], s["d2"], s["fv2"])
    assert n == rn and np.array_equal(match, rmatch)
    assert n > 30
    m.close()


@pytest.mark.skipif(not ref.available(), reason="oracle/_ref not shipped")
def test_cuda_keyframe_and_current_frame_matches_compiled_reference():
    """M3: CUDA window search == the reference's own searchByProjectionInKeyFrameAndCurrentFrame (src/orbMatcher.cpp:156-239
    compiled unmodified) on the queries the reference's own methods produce."""
    from tests import matcher_cases as mc
    from tests.test_oracle_ref_frame import _keyframe_scene
    from ydorbslam_b200 import OrbMatcher
    from ydorbslam_b200 import queries as Q

    kps, desc = mc.synth_keypoints(900, 23)
    s = _keyframe_scene(kps, desc, 3000, 33, 7.0)
    f = ref.RefFrame.from_keypoints(kps, (0, 0, mc.W, mc.H))
    f.set_features(desc, s["right_x_kp"], mc.SF)
    cam = [mc.CAM[k] for k in ("fx", "fy", "cx", "cy", "bf", "baseline")]
    rn, rmatch, dbg = f.search_keyframe_current(s["last_kps"], s["pos"], s["desc"], s["min_d"], s["max_d"], s["bad"], s["found"], cam,
                                                s["occupied"], 7.0, 100, 0.9, True)
    q, prm = Q.keyframe_and_current_frame(dbg[:, 0], dbg[:, 1], dbg[:, 2].astype(np.int32), s["last_kps"]["angle"], dbg[:, 3] > 0, mc.SF,
                                          7.0, 100)
    m = OrbMatcher(0.9)
    fr = m.make_frame(kps, desc, s["right_x_kp"], bounds=(0, 0, mc.W, mc.H), scale_factors=mc.SF)
    n, best, match = m.searchWindows(fr, q, s["desc"], prm, s["occupied"])
    assert n == rn and np.array_equal(np.where(match < 0, -1, match), rmatch)
    assert n > 40
    fr.close(); m.close(); f.close()


@pytest.mark.skipif(not ref.available(), reason="oracle/_ref not shipped")
def test_cuda_search_by_sim3_matches_compiled_reference():
    """M4: CUDA searchBySim3 (two window searches + agreement) == the reference's own (src/orbMatcher.cpp:566-681 compiled
    unmodified), pre-matched slots included."""
    from tests import matcher_cases as mc
    from tests.test_oracle_ref_frame import _sim3_sides
    from ydorbslam_b200 import OrbMatcher

    side1, side2 = _sim3_sides(41)
    n1, n2 = len(side1["kps"]), len(side2["kps"])
    rng = np.random.RandomState(3)
    pre = np.where(rng.rand(n1) < 0.05, rng.randint(0, n2, n1), -1).astype(np.int32)
    cam = [mc.CAM[k] for k in ("fx", "fy", "cx", "cy", "bf", "baseline")]
    rn, rmatch, dbg1, dbg2 = ref.search_by_sim3(side1, side2, cam, (0, 0, mc.W, mc.H), mc.SF, 7.5, pre)
    blocked2 = np.zeros(n2, bool)
    blocked2[pre[pre >= 0]] = True
    m = OrbMatcher(0.9)
    none1, none2 = np.full(n1, -1, np.float32), np.full(n2, -1, np.float32)
    f1 = m.make_frame(side1["kps"], side1["desc"], none1, bounds=(0, 0, mc.W, mc.H), scale_factors=mc.SF)
    f2 = m.make_frame(side2["kps"], side2["desc"], none2, bounds=(0, 0, mc.W, mc.H), scale_factors=mc.SF)
    n, out = m.searchBySim3(f1, f2, (dbg1[:, 0], dbg1[:, 1], dbg1[:, 2].astype(np.int32), (dbg1[:, 3] > 0) & (pre < 0), side1["mp_desc"]),
                            (dbg2[:, 0], dbg2[:, 1], dbg2[:, 2].astype(np.int32), (dbg2[:, 3] > 0) & ~blocked2, side2["mp_desc"]), mc.SF, 7.5)
    assert n == rn and np.array_equal(out, np.where(rmatch == -2, -1, rmatch))
    assert n > 20
    f1.close(); f2.close(); m.close()


@pytest.mark.skipif(not ref.available(), reason="oracle/_ref not shipped")
@pytest.mark.parametrize("mode,thd,sim_t", [(0, 1.2, (0, 0, 0)), (1, 4.0, (0.3, -0.2, 0.5))])
def test_cuda_fuse_matches_compiled_reference(mode, thd, sim_t):
    """M4: CUDA candidate search of fuseByProjection / fuseBySim3 + the caller-side replay == every observable of the
    reference's own call (src/orbMatcher.cpp:682-807 compiled unmodified)."""
    from tests import matcher_cases as mc
    from tests.test_oracle_ref_frame import check_fuse, fuse_scene, replay_fuse
    from ydorbslam_b200 import OrbMatcher

    s = fuse_scene(mode, 61 + mode, thd, sim_t)
    cam = [mc.CAM[k] for k in ("fx", "fy", "cx", "cy", "bf", "baseline")]
    r = ref.fuse(mode, s["kps"], s["desc"], s["right_x"], s["slot_kind"], s["slot_obs"], s["points"], cam, (0, 0, mc.W, mc.H), mc.SF, thd,
                 2.0 if mode else 1.0, sim_t)
    dbg = r["dbg"]
    active, level = dbg[:, 4] > 0, dbg[:, 3].astype(np.int32)
    m = OrbMatcher(0.6)
    fr = m.make_frame(s["kps"], s["desc"], s["right_x"], bounds=(0, 0, mc.W, mc.H), scale_factors=mc.SF)
    if mode == 0:
        _, best = m.fuseByProjectionCandidates(fr, dbg[:, 0], dbg[:, 1], dbg[:, 2], level, active, s["points"]["desc"], mc.SF, mc.INV_SF2, thd)
    else:
        _, best = m.fuseBySim3Candidates(fr, dbg[:, 0], dbg[:, 1], level, active, s["points"]["desc"], mc.SF, thd)
    check_fuse(r, replay_fuse(mode, best, s), s)
    assert r["fused"] > 30
    fr.close(); m.close()


@pytest.mark.skipif(not ref.available(), reason="oracle/_ref not shipped")
@pytest.mark.parametrize("stereo_only", [False, True])
def test_cuda_search_for_triangulation_matches_compiled_reference(stereo_only):
    """M6: CUDA searchForTriangulation == the reference's own (src/orbMatcher.cpp:463-565, :808-819 compiled unmodified)."""
    from tests import matcher_cases as mc
    from ydorbslam_b200 import OrbMatcher

    s = mc.bow_scene(1000, 1100, 3)
    F, ep = mc.fundamental(3)
    has1, has2 = 1 - s["v1"], 1 - s["v2"]
    cam = [mc.CAM[k] for k in ("fx", "fy", "cx", "cy", "bf", "baseline")]
    origin1 = [(float(ep[0]) - float(mc.CAM["cx"])) / float(mc.CAM["fx"]), (float(ep[1]) - float(mc.CAM["cy"])) / float(mc.CAM["fy"]), 1.0]
    rn, rpairs, rep = ref.search_for_triangulation(s["k1"], s["d1"], s["rx1"], has1, s["fv1"], s["k2"], s["d2"], s["rx2"], has2,
                                                   s["fv2"], F, origin1, cam, mc.SF, mc.SF2, stereo_only, 0.6, True)
    m = OrbMatcher(0.6, True)
    n, pairs = m.searchForTriangulation(s["k1"], s["d1"], s["rx1"], has1, s["fv1"], s["k2"], s["d2"], s["rx2"], has2, s["fv2"],
                                        F, rep, mc.SF, mc.SF2, stereo_only)
    assert n == rn and np.array_equal(pairs, rpairs)
    assert n > 15
    m.close()
