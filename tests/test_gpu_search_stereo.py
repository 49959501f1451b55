"""GPU parity: Frame grid / getKeyPointsInArea, searchByProjectionInFrameAndMapPoint and computeStereoMatches
vs the CPU oracle (indices, counts and float outputs bit-exact)."""
import numpy as np
import pytest

from oracle import orc
from ydorbslam_b200 import MPPROJ_DTYPE, OrbExtractor, OrbMatcher, synth

pytestmark = pytest.mark.gpu


def _mps_array(mp):
    a = np.zeros(len(mp["proj_x"]), MPPROJ_DTYPE)
    for f in ("proj_x", "proj_y", "proj_right_x", "view_cos", "level", "active"):
        a[f] = mp[f]
    return a


@pytest.fixture(scope="module")
def frame640():
    img = synth.synth_frame(640, 480, 41)
    ex = OrbExtractor(1000, 1.2, 8, 20, 7, max_width=640, max_height=480)
    kps, desc = ex.extractAndCompute(img)
    sf = ex.getScaleFactors()
    ex.close()
    return img, kps, desc, sf


def test_keypoints_in_area(frame640):
    img, kps, desc, sf = frame640
    p = orc.make_params(1000, 1.2, 8, 20)
    g = orc.make_geom(p, 640, 480)
    m = OrbMatcher(0.8)
    fr = m.make_frame(kps, desc, None, bounds=(0, 0, 640, 480), scale_factors=sf)
    rng = np.random.RandomState(3)
    nonempty = 0
    for _ in range(200):
        x, y = rng.uniform(-20, 660), rng.uniform(-20, 500)
        r = rng.choice([4.0, 10.0, 25.0, 60.0])
        mn, mx = rng.choice([-1, 0, 1, 3]), rng.choice([-1, 0, 2, 5])
        got = fr.getKeyPointsInArea(x, y, r, int(mn), int(mx))
        ref = orc.keypoints_in_area(g, kps, x, y, r, int(mn), int(mx))
        assert np.array_equal(got, ref), (x, y, r, mn, mx)
        nonempty += len(ref) > 0
    assert nonempty > 20
    fr.close()


@pytest.mark.parametrize("nmp,th,occupied", [(3000, 3.0, False), (20000, 1.0, True), (500, 5.0, True), (0, 3.0, False)])
def test_search_by_projection(frame640, nmp, th, occupied):
    img, kps, desc, sf = frame640
    p = orc.make_params(1000, 1.2, 8, 20)
    g = orc.make_geom(p, 640, 480)
    mp = synth.synth_map_points(kps, desc, max(nmp, 1), 640, 480, 8, seed=50 + nmp)
    mps = _mps_array(mp)[:nmp]
    mpd = mp["desc"][:nmp]
    obs = mp["obs_positive"][:nmp]
    rng = np.random.RandomState(1)
    right_x = np.where(rng.rand(len(kps)) < 0.5, kps["x"] - rng.uniform(0, 30, len(kps)), -1.0).astype(np.float32)
    occ = (rng.rand(len(kps)) < 0.1).astype(np.uint8) if occupied else np.zeros(len(kps), np.uint8)
    m = OrbMatcher(0.8)
    fr = m.make_frame(kps, desc, right_x, bounds=(0, 0, 640, 480), scale_factors=sf)
    n, match = m.searchByProjectionInFrameAndMapPoint(fr, mps, mpd, obs, th, kp_occupied=occ if occupied else None)
    rn, rmatch = orc.search_by_projection(g, kps, desc, right_x, occ, mps, mpd, obs, th, 0.8)
    assert n == rn
    assert np.array_equal(match, rmatch)
    if nmp >= 3000:
        assert n > 50
    # the same search with the local map resident on the device
    import torch

    dev = torch.device("cuda:0")
    d_mps = torch.from_numpy(mps.view(np.uint8).reshape(-1).copy() if nmp else np.zeros(16, np.uint8)).to(dev)
    d_mpd = torch.from_numpy(np.ascontiguousarray(mpd).reshape(-1).copy() if nmp else np.zeros(32, np.uint8)).to(dev)
    d_obs = torch.from_numpy(np.ascontiguousarray(obs, dtype=np.uint8).copy() if nmp else np.zeros(1, np.uint8)).to(dev)
    d_occ = torch.from_numpy(occ).to(dev)
    d_out = torch.full((len(kps),), 7, dtype=torch.int32, device=dev)
    dn = m.searchByProjectionInFrameAndMapPoint_device(fr, d_mps.data_ptr(), d_mpd.data_ptr(), d_obs.data_ptr(), nmp,
                                                       d_out.data_ptr(), th, d_occ.data_ptr() if occupied else None)
    assert dn == rn and np.array_equal(d_out.cpu().numpy(), rmatch)
    fr.close()


def test_search_greedy_conflicts():
    """Many map points fighting for few keypoints: exercises the order-dependent exclusion."""
    img = synth.synth_sparse_frame(640, 480, 7)
    ex = OrbExtractor(300, 1.2, 8, 20, 7, max_width=640, max_height=480)
    kps, desc = ex.extractAndCompute(img)
    sf = ex.getScaleFactors()
    ex.close()
    assert len(kps) > 20
    p = orc.make_params(300, 1.2, 8, 20)
    g = orc.make_geom(p, 640, 480)
    mp = synth.synth_map_points(kps, desc, 6000, 640, 480, 8, seed=77)
    mps = _mps_array(mp)
    right_x = np.full(len(kps), -1.0, np.float32)
    occ = np.zeros(len(kps), np.uint8)
    m = OrbMatcher(0.9)
    fr = m.make_frame(kps, desc, right_x, bounds=(0, 0, 640, 480), scale_factors=sf)
    n, match = m.searchByProjectionInFrameAndMapPoint(fr, mps, mp["desc"], mp["obs_positive"], 5.0)
    rn, rmatch = orc.search_by_projection(g, kps, desc, right_x, occ, mps, mp["desc"], mp["obs_positive"], 5.0, 0.9)
    assert n == rn and np.array_equal(match, rmatch)
    fr.close()


@pytest.mark.parametrize("w,h,nf,seed,max_d", [(1241, 376, 2000, 61, np.inf), (640, 480, 1000, 62, 500.0), (1241, 376, 2000, 63, 718.856)])
@pytest.mark.parametrize("mode", [0, 1])
def test_stereo_match(w, h, nf, seed, max_d, mode):
    left, right = synth.synth_stereo_pair(w, h, seed)
    exl = OrbExtractor(nf, 1.2, 8, 20, 7, max_width=w, max_height=h)
    exr = OrbExtractor(nf, 1.2, 8, 20, 7, max_width=w, max_height=h)
    kl, dl = exl.extractAndCompute(left)
    kr, dr = exr.extractAndCompute(right)
    m = OrbMatcher()
    bf = 386.1448
    rx, dp = m.computeStereoMatches(exl, 0, exr, 0, kl, dl, kr, dr, bf, max_d, mode)
    p = orc.make_params(nf, 1.2, 8, 20)
    _, pl = orc.pyramid(p, left)
    _, pr = orc.pyramid(p, right)
    orx, odp = orc.stereo_match(p, w, h, pl, pr, kl, dl, kr, dr, bf, max_d, mode)
    assert np.array_equal(rx.view(np.uint32), orx.view(np.uint32))
    assert np.array_equal(dp.view(np.uint32), odp.view(np.uint32))
    assert (dp > 0).sum() > 50, "synthetic pair should produce stereo matches"
    exl.close(); exr.close()


def test_stereo_identical_images_median_zero():
    """left == right: every SAD is 0, the median rule invalidates all matches (-2)."""
    left = synth.synth_frame(640, 480, 71)
    exl = OrbExtractor(500, 1.2, 8, 20, 7, max_width=640, max_height=480)
    exr = OrbExtractor(500, 1.2, 8, 20, 7, max_width=640, max_height=480)
    kl, dl = exl.extractAndCompute(left)
    kr, dr = exr.extractAndCompute(left)
    m = OrbMatcher()
    rx, dp = m.computeStereoMatches(exl, 0, exr, 0, kl, dl, kr, dr, 40.0, np.inf, 0)
    p = orc.make_params(500, 1.2, 8, 20)
    _, pl = orc.pyramid(p, left)
    orx, odp = orc.stereo_match(p, 640, 480, pl, pl, kl, dl, kr, dr, 40.0, np.inf, 0)
    assert np.array_equal(rx, orx) and np.array_equal(dp, odp)
    assert (rx == -2).sum() > 0
    exl.close(); exr.close()


@pytest.mark.parametrize("keep_rows,expect_big_lag", [(0.5, True), (0.97, False)])
def test_stereo_lag_window(keep_rows, expect_big_lag):
    """Reference mode resolves the leftIdx lag (frame.cpp:388-403) from a 64-wide table; right keypoints confined to the
    top of the image leave many left rows empty, so the lag outgrows the window and the general chain takes over."""
    w, h, nf = 1241, 376, 2000
    left, right = synth.synth_stereo_pair(w, h, 64)
    exl = OrbExtractor(nf, 1.2, 8, 20, 7, max_width=w, max_height=h)
    exr = OrbExtractor(nf, 1.2, 8, 20, 7, max_width=w, max_height=h)
    kl, dl = exl.extractAndCompute(left)
    kr, dr = exr.extractAndCompute(right)
    keep = kr["y"] < h * keep_rows
    kr, dr = np.ascontiguousarray(kr[keep]), np.ascontiguousarray(dr[keep])
    empty_rows = int((kl["y"] > h * keep_rows + 40).sum())
    assert (empty_rows > 64) == expect_big_lag
    m = OrbMatcher()
    bf = 386.1448
    p = orc.make_params(nf, 1.2, 8, 20)
    _, pl = orc.pyramid(p, left)
    _, pr = orc.pyramid(p, right)
    for mode in (0, 1):
        rx, dp = m.computeStereoMatches(exl, 0, exr, 0, kl, dl, kr, dr, bf, np.inf, mode)
        orx, odp = orc.stereo_match(p, w, h, pl, pr, kl, dl, kr, dr, bf, np.inf, mode)
        assert np.array_equal(rx.view(np.uint32), orx.view(np.uint32))
        assert np.array_equal(dp.view(np.uint32), odp.view(np.uint32))
    exl.close(); exr.close()


@pytest.mark.parametrize("pool_cap", ["0", "7", "200"])
def test_search_by_projection_pool_fallbacks(frame640, pool_cap, monkeypatch):
    """M1 with the candidate pool forced small (lists that do not fit are walked again every round) or absent (the path
    of a frame with more than 2^19 key points), and with windows wide enough to span more than 32 grid columns and
    hold more than 32 candidates: same assignments as the oracle."""
    img, kps, desc, sf = frame640
    monkeypatch.setenv("YDORB_SEARCH_POOL_CAP", pool_cap)
    p = orc.make_params(1000, 1.2, 8, 20)
    g = orc.make_geom(p, 640, 480)
    mp = synth.synth_map_points(kps, desc, 4000, 640, 480, 8, seed=77)
    mps = _mps_array(mp)
    mps["level"] = np.minimum(mps["level"], 2)  # low levels keep most octaves, so windows are well filled
    rng = np.random.RandomState(2)
    right_x = np.where(rng.rand(len(kps)) < 0.5, kps["x"] - rng.uniform(0, 30, len(kps)), -1.0).astype(np.float32)
    occ = (rng.rand(len(kps)) < 0.1).astype(np.uint8)
    m = OrbMatcher(0.8)
    fr = m.make_frame(kps, desc, right_x, bounds=(0, 0, 640, 480), scale_factors=sf)
    for th in (3.0, 40.0):  # th 40: radius 100..230 px -> 10..23 grid columns either side, hundreds of candidates
        n, match = m.searchByProjectionInFrameAndMapPoint(fr, mps, mp["desc"], mp["obs_positive"], th, kp_occupied=occ)
        rn, rmatch = orc.search_by_projection(g, kps, desc, right_x, occ, mps, mp["desc"], mp["obs_positive"], th, 0.8)
        assert n == rn and np.array_equal(match, rmatch), (pool_cap, th)
    fr.close(); m.close()


def test_octaves_outside_the_pyramid_are_rejected(frame640):
    """Caller-supplied octaves / levels index per-level tables on the device: the host entry points refuse bad ones."""
    from ydorbslam_b200._lib import YdorbError

    img, kps, desc, sf = frame640
    m = OrbMatcher(0.8)
    bad = kps.copy()
    bad["octave"][3] = 8
    with pytest.raises(YdorbError):
        m.make_frame(bad, desc, None, bounds=(0, 0, 640, 480), scale_factors=sf)
    bad["octave"][3] = -1
    with pytest.raises(YdorbError):
        m.make_frame(bad, desc, None, bounds=(0, 0, 640, 480), scale_factors=sf)
    fr = m.make_frame(kps, desc, None, bounds=(0, 0, 640, 480), scale_factors=sf)
    mp = synth.synth_map_points(kps, desc, 10, 640, 480, 8, seed=5)
    mps = _mps_array(mp)
    mps["level"][4] = 9
    with pytest.raises(YdorbError):
        m.searchByProjectionInFrameAndMapPoint(fr, mps, mp["desc"], mp["obs_positive"], 3.0)
    fr.close(); m.close()
