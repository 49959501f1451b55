"""GPU parity at BASELINE.json's full sizes (configs 2-5), through the C ABI: direct comparison with the oracle where it
finishes in seconds, plus size-independent properties (batch == frame-by-frame, sharded == single scan, idempotence)."""
import numpy as np
import pytest

from oracle import orc
from tests.test_gpu_search_stereo import _mps_array
from ydorbslam_b200 import OrbExtractor, OrbMatcher, sharding, synth

pytestmark = pytest.mark.gpu


def _same(kps, desc, ok, od):
    assert len(kps) == len(ok)
    for f in ("x", "y", "size", "angle", "response", "octave"):
        assert np.array_equal(kps[f], ok[f]), f
    assert np.array_equal(desc, od)


def test_config3_batch_1080p_4000kp():
    """Batched 1920x1080 x 4000 kp: a batch of distinct frames equals the oracle frame by frame; running the batch in
    two halves on two handles (the multi-GPU frame sharding, i mod G) gives the same bytes; a second run is identical."""
    B = 6
    imgs = np.stack([synth.synth_frame(1920, 1080, 300 + i) for i in range(B)])
    p = orc.make_params(4000, 1.2, 8, 20)
    ex = OrbExtractor(4000, 1.2, 8, 20, 7, max_width=1920, max_height=1080, max_batch=B)
    kps, desc, n = ex.extract_batch(imgs)
    for b in (0, B - 1):
        ok, od, _ = orc.extract(p, imgs[b], cap=4200)
        _same(kps[b, : n[b]], desc[b, : n[b]], ok, od)
    assert (n == 4000).all()
    kps2, desc2, n2 = ex.extract_batch(imgs)
    assert np.array_equal(n, n2) and np.array_equal(kps, kps2) and np.array_equal(desc, desc2)
    G = 2
    for g in range(G):
        mine = sharding.frames_of_rank(B, g, G)
        exg = OrbExtractor(4000, 1.2, 8, 20, 7, max_width=1920, max_height=1080, max_batch=len(mine))
        kg, dg, ng = exg.extract_batch(imgs[mine])
        assert np.array_equal(ng, n[mine]) and np.array_equal(kg, kps[mine]) and np.array_equal(dg, desc[mine])
        exg.close()
    ex.close()


def test_config4_hamming_2000_vs_1M_sharded():
    """2000 queries x 1 M database rows: single scan == oracle; 8 contiguous shards + merge == single scan."""
    import torch

    ndb, nq, G = 1_000_000, 2000, 8
    db = synth.synth_descriptors(ndb, 1)
    q, rows = synth.synth_query_descriptors(db, nq, 2)
    m = OrbMatcher(0.6)
    full = m.top2(q, db)
    ref = orc.top2(q, db)
    assert np.array_equal(full, ref)
    assert (full["best_idx"][: nq // 2] == rows).mean() > 0.99  # perturbed copies find their source row
    parts = np.stack([m.top2(q, db[b:e], idx_base=b) for b, e in (sharding.shard_range(ndb, g, G) for g in range(G))])
    dev = torch.device("cuda:0")
    t_parts = torch.from_numpy(parts.view(np.int32).reshape(G, nq, 4)).to(dev)
    t_out = torch.zeros(nq, 4, dtype=torch.int32, device=dev)
    torch.cuda.synchronize()
    m.top2_merge_device(t_parts.data_ptr(), G, nq, t_out.data_ptr())
    m.sync()
    assert np.array_equal(t_out.cpu().numpy().view(orc.TOP2_DTYPE).reshape(-1), full)
    acc = orc.ratio_accept(full, 0.6, 50)
    assert acc[: nq // 2].mean() > 0.9 and acc[nq // 2 :].sum() == 0
    m.close()


def test_config5_4k_8000kp_search_200k_map_points():
    """3840x2160, 8000 features, then searchByProjectionInFrameAndMapPoint against 200 000 projected map points."""
    w, h, nf, nmp = 3840, 2160, 8000, 200_000
    img = synth.synth_frame(w, h, 500, n_rect=200)
    p = orc.make_params(nf, 1.2, 8, 20)
    ex = OrbExtractor(nf, 1.2, 8, 20, 7, max_width=w, max_height=h)
    kps, desc = ex.extractAndCompute(img)
    ok, od, _ = orc.extract(p, img, cap=nf + 64)
    _same(kps, desc, ok, od)
    assert len(kps) == nf
    sf = ex.getScaleFactors()
    ex.close()
    g = orc.make_geom(p, w, h)
    mp = synth.synth_map_points(kps, desc, nmp, w, h, 8, seed=3)
    mps = _mps_array(mp)
    rng = np.random.RandomState(5)
    right_x = np.where(rng.rand(nf) < 0.5, kps["x"] - rng.uniform(0, 30, nf), -1.0).astype(np.float32)
    occ = (rng.rand(nf) < 0.05).astype(np.uint8)
    m = OrbMatcher(0.8)
    fr = m.make_frame(kps, desc, right_x, bounds=(0, 0, w, h), scale_factors=sf)
    n, match = m.searchByProjectionInFrameAndMapPoint(fr, mps, mp["desc"], mp["obs_positive"], 3.0, kp_occupied=occ)
    rn, rmatch = orc.search_by_projection(g, kps, desc, right_x, occ, mps, mp["desc"], mp["obs_positive"], 3.0, 0.8)
    assert n == rn and np.array_equal(match, rmatch)
    assert n > 100
    fr.close(); m.close()


def test_config2_kitti_pair_end_to_end():
    """1241x376 stereo pair, 2000 features per image, both extractors in one batch-of-1 each + computeStereoMatches;
    every stage equals the oracle (keypoints, descriptors, rightX, depth) and depth = bf / disparity where matched."""
    w, h, nf, bf = 1241, 376, 2000, 386.1448
    left, right = synth.synth_stereo_pair(w, h, 64)
    p = orc.make_params(nf, 1.2, 8, 20)
    exl = OrbExtractor(nf, 1.2, 8, 20, 7, max_width=w, max_height=h)
    exr = OrbExtractor(nf, 1.2, 8, 20, 7, max_width=w, max_height=h)
    kl, dl = exl.extractAndCompute(left)
    kr, dr = exr.extractAndCompute(right)
    for img, k, d in ((left, kl, dl), (right, kr, dr)):
        ok, od, _ = orc.extract(p, img, cap=nf + 64)
        _same(k, d, ok, od)
    m = OrbMatcher()
    rx, dp = m.computeStereoMatches(exl, 0, exr, 0, kl, dl, kr, dr, bf, np.inf, 0)
    _, pl = orc.pyramid(p, left)
    _, pr = orc.pyramid(p, right)
    orx, odp = orc.stereo_match(p, w, h, pl, pr, kl, dl, kr, dr, bf, np.inf, 0)
    assert np.array_equal(rx.view(np.uint32), orx.view(np.uint32)) and np.array_equal(dp.view(np.uint32), odp.view(np.uint32))
    ok = dp > 0
    assert ok.sum() > 100
    exl.close(); exr.close(); m.close()
