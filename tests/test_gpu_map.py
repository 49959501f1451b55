"""GPU parity of the map-point rows through the C ABI: bit-exact against the CPU oracle."""
import numpy as np
import pytest

from oracle import orc
from tests import map_cases
from ydorbslam_b200 import OrbMatcher, mappoints

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("n_points,seed", [(1, 1), (300, 11), (5000, 12)])
def test_distinctive_descriptors(n_points, seed):
    desc, start = map_cases.observation_sets(n_points, seed)
    m = OrbMatcher()
    got = mappoints.computeDistinctiveDescriptors(m, desc, start)
    assert np.array_equal(got, orc.distinctive_descriptors(desc, start))
    m.close()


def test_distinctive_descriptors_empty_and_large():
    m = OrbMatcher()
    assert len(mappoints.computeDistinctiveDescriptors(m, np.zeros((0, 32), np.uint8), np.zeros(1, np.int32))) == 0
    # all points empty
    got = mappoints.computeDistinctiveDescriptors(m, np.zeros((0, 32), np.uint8), np.zeros(6, np.int32))
    assert (got == -1).all()
    # one point with 700 observations (rows loop over the warp many times), incl. identical rows
    rng = np.random.RandomState(3)
    desc = rng.randint(0, 256, (700, 32)).astype(np.uint8)
    desc[100:140] = desc[5]
    start = np.array([0, 700], np.int32)
    assert np.array_equal(mappoints.computeDistinctiveDescriptors(m, desc, start), orc.distinctive_descriptors(desc, start))
    m.close()


@pytest.mark.parametrize("n,seed,limit", [(1, 2, 0.5), (4000, 5, 0.5), (200_000, 6, 0.5), (50_000, 7, 0.0)])
def test_frustum_matches_oracle(n, seed, limit):
    cam, mp = map_cases.camera_and_points(n, seed)
    m = OrbMatcher()
    got = mappoints.isInCameraFrustum(m, cam, mp, limit)
    ref = orc.frustum(cam, mp, limit)
    for f in ("active", "level"):
        assert np.array_equal(got[f], ref[f]), f
    for f in ("proj_x", "proj_y", "proj_right_x", "view_cos"):
        assert np.array_equal(got[f].view(np.uint32), ref[f].view(np.uint32)), f
    if n > 1000:
        assert 0.05 < got["active"].mean() < 0.95
    m.close()


def test_frustum_feeds_search_by_projection():
    """The pair tracking runs every frame (src/tracking.cpp:585-606): frustum test, then searchByProjection on its output."""
    from ydorbslam_b200 import OrbExtractor, synth

    w, h, nf = 1241, 376, 2000
    img = synth.synth_frame(w, h, 41)
    ex = OrbExtractor(nf, 1.2, 8, 20, 7, max_width=w, max_height=h)
    kps, desc = ex.extractAndCompute(img)
    sf = ex.getScaleFactors()
    ex.close()
    cam, mp = map_cases.camera_and_points(30_000, 9, w, h)
    rng = np.random.RandomState(4)
    src = rng.randint(0, len(kps), len(mp))
    mp_desc = desc[src].copy()
    mp_desc[rng.rand(len(mp)) < 0.5] = rng.randint(0, 256, 32)
    obs = (rng.rand(len(mp)) < 0.9).astype(np.uint8)
    m = OrbMatcher(0.8)
    proj = mappoints.isInCameraFrustum(m, cam, mp, 0.5)
    ref_proj = orc.frustum(cam, mp, 0.5)
    assert np.array_equal(proj.view(np.uint8), ref_proj.view(np.uint8))
    fr = m.make_frame(kps, desc, None, bounds=(0, 0, w, h), scale_factors=sf)
    n, match = m.searchByProjectionInFrameAndMapPoint(fr, proj, mp_desc, obs, 1.0)
    p = orc.make_params(nf, 1.2, 8, 20)
    g = orc.make_geom(p, w, h)
    rn, rmatch = orc.search_by_projection(g, kps, desc, np.full(len(kps), -1, np.float32), np.zeros(len(kps), np.uint8), ref_proj,
                                          mp_desc, obs, 1.0, 0.8)
    assert n == rn and np.array_equal(match, rmatch)
    fr.close(); m.close()
