"""CPU: the oracle's map-point rows against independent restatements -- numpy for the integer
descriptor selection, cv2's own gemm / norm for the cv::Mat arithmetic of isInCameraFrustum."""
import numpy as np
import pytest

from oracle import orc
from tests import map_cases


def _distinctive_numpy(desc, start):
    best = np.full(len(start) - 1, -1, np.int32)
    for p in range(len(start) - 1):
        d = desc[start[p] : start[p + 1]]
        n = len(d)
        if n == 0:
            continue
        bits = np.unpackbits(d, axis=1).astype(np.int32)
        dist = (bits[:, None, :] != bits[None, :, :]).sum(2)
        med = np.sort(dist, axis=1)[:, int(0.5 * n)]
        best[p] = int(np.argmin(med))  # first minimum
    return best


def test_distinctive_descriptors_vs_numpy():
    desc, start = map_cases.observation_sets(300, 11)
    got = orc.distinctive_descriptors(desc, start)
    assert np.array_equal(got, _distinctive_numpy(desc, start))
    assert (got == -1).sum() > 0 and (got > 0).sum() > 50


def test_frustum_vs_cv2_primitives():
    cv2 = pytest.importorskip("cv2")
    cam, mp = map_cases.camera_and_points(4000, 5)
    out = orc.frustum(cam, mp, 0.5)
    R = np.array(cam.R[:], np.float32).reshape(3, 3)
    t = np.array(cam.t[:], np.float32).reshape(3, 1)
    origin = np.array(cam.origin[:], np.float32)
    f32 = np.float32
    n_active = 0
    for i in range(len(mp)):
        P = mp["pos"][i].reshape(3, 1).copy()
        pc = cv2.gemm(R, P, 1.0, t, 1.0).ravel()  # m_cvMat_R_c2w * pos + m_cvMat_t_c2w
        x = f32(f32(f32(cam.fx) * pc[0]) / pc[2]) + f32(cam.cx)
        y = f32(f32(f32(cam.fy) * pc[1]) / pc[2]) + f32(cam.cy)
        v = (mp["pos"][i] - origin).astype(np.float32)
        dist = f32(cv2.norm(v.reshape(3, 1)))
        view_cos = f32(float(v.astype(np.float64) @ mp["normal"][i].astype(np.float64)) / float(dist))
        ok = (pc[2] >= 0 and cam.min_x <= x < cam.max_x and cam.min_y <= y < cam.max_y
              and dist >= f32(0.8) * mp["min_distance"][i] and dist <= f32(1.2) * mp["max_distance"][i] and view_cos >= f32(0.5))
        assert bool(out["active"][i]) == bool(ok), i
        if ok:
            n_active += 1
            ratio = f32(mp["max_distance"][i] / dist)
            lvl = int(np.ceil(np.log(float(ratio)) / float(f32(cam.log_scale_factor))))
            lvl = min(max(lvl, 0), cam.n_levels - 1)
            assert out["proj_x"][i] == x and out["proj_y"][i] == y and out["level"][i] == lvl
            assert out["view_cos"][i] == view_cos
            assert out["proj_right_x"][i] == f32(x - f32(f32(cam.bf) / pc[2]))
    assert 200 < n_active < len(mp) - 200  # every gate is exercised on both sides
