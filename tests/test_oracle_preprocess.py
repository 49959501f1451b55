"""CPU: the oracle's pre-processing rows against the committed cv2-generated golden vectors, and live
against cv2 itself when it is importable (cvtColor *2GRAY, undistortPoints: bit-exact)."""
import numpy as np
import pytest

from oracle import orc
from tests import pre_cases
from tests.conftest import GOLDEN

CODES = {"rgb": 0, "bgr": 1, "rgba": 2, "bgra": 3}


def _bits(a):
    return np.ascontiguousarray(a, np.float32).view(np.uint32)


def test_cvt_gray_vs_golden():
    g = np.load(GOLDEN / "preprocess.npz")
    img3, img4 = pre_cases.color_image(61, 23, 3, 1), pre_cases.color_image(61, 23, 4, 2)
    for name, code in CODES.items():
        assert np.array_equal(orc.cvt_gray(img4 if code >= 2 else img3, code), g["gray_" + name]), name


@pytest.mark.parametrize("name", sorted(pre_cases.DISTS))
def test_undistort_vs_golden(name):
    g = np.load(GOLDEN / "preprocess.npz")
    pts = pre_cases.points(400, 3, spread=6.0 if name == "barrel" else 1.2)
    dist = pre_cases.DISTS[name]
    assert np.array_equal(_bits(orc.undistort_points(pts, pre_cases.K, dist, pre_cases.K)), _bits(g["und_" + name]))
    assert np.array_equal(_bits(orc.undistort_points(pts, pre_cases.K, dist, pre_cases.P_OTHER)), _bits(g["undP_" + name]))


def test_live_cv2():
    cv2 = pytest.importorskip("cv2")
    big = pre_cases.color_image(1023, 517, 3, 5)
    assert np.array_equal(orc.cvt_gray(big, 0), cv2.cvtColor(big, cv2.COLOR_RGB2GRAY))
    assert np.array_equal(orc.cvt_gray(big, 1), cv2.cvtColor(big, cv2.COLOR_BGR2GRAY))
    big4 = pre_cases.color_image(640, 480, 4, 6)
    assert np.array_equal(orc.cvt_gray(big4, 2), cv2.cvtColor(big4, cv2.COLOR_RGBA2GRAY))
    assert np.array_equal(orc.cvt_gray(big4, 3), cv2.cvtColor(big4, cv2.COLOR_BGRA2GRAY))
    # every (r, g, b) grey level boundary: a 256^3 sweep is 16M pixels, cheap
    r, gch, b = np.meshgrid(np.arange(256, dtype=np.uint8), np.arange(256, dtype=np.uint8), np.arange(0, 256, 5, dtype=np.uint8), indexing="ij")
    cube = np.stack([r, gch, b], -1).reshape(256, -1, 3)
    assert np.array_equal(orc.cvt_gray(cube, 0), cv2.cvtColor(cube, cv2.COLOR_RGB2GRAY))
    for name, dist in pre_cases.DISTS.items():
        pts = pre_cases.points(20000, 11, spread=8.0 if name == "barrel" else 1.3)
        ref = cv2.undistortPoints(pts.reshape(-1, 1, 2), pre_cases.K, dist, None, pre_cases.K).reshape(-1, 2)
        got = orc.undistort_points(pts, pre_cases.K, dist, pre_cases.K)
        same = (_bits(got) == _bits(ref)) | (np.isnan(got) & np.isnan(ref))
        assert same.all(), (name, int((~same).sum()))


def test_stereo_from_rgbd_vs_numpy():
    w, h = 640, 480
    d16 = pre_cases.depth_map_u16(w, h, 3)
    xy = pre_cases.points(5000, 4, spread=1.1)  # some fall outside the map
    kps = pre_cases.keypoints_at(xy)
    f, bf = np.float32(1.0 / 5000.0), np.float32(40.0)
    for dm, factor in [(d16, f), (d16.astype(np.float32), f), (d16.astype(np.float32) * f, np.float32(1.0))]:
        rx, dz = orc.stereo_from_rgbd(kps, dm, factor, bf)
        ix, iy = xy[:, 0].astype(np.int32), xy[:, 1].astype(np.int32)  # C truncation
        inside = (xy[:, 0] > -1) & (xy[:, 1] > -1) & (ix >= 0) & (iy >= 0) & (ix < w) & (iy < h)
        raw = dm[np.clip(iy, 0, h - 1), np.clip(ix, 0, w - 1)].astype(np.float32)
        d = raw * factor if (dm.dtype == np.uint16 or abs(factor - 1) > 1e-5) else raw
        ok = inside & (d > 0)
        assert np.array_equal(_bits(dz), _bits(np.where(ok, d, np.float32(-1))))
        with np.errstate(divide="ignore"):
            exp_rx = np.where(ok, xy[:, 0] - bf / np.where(ok, d, np.float32(1)), np.float32(-1)).astype(np.float32)
        assert np.array_equal(_bits(rx), _bits(exp_rx))
        assert 0 < ok.sum() < len(xy) and (~inside).sum() > 0
