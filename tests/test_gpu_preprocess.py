"""GPU parity: cvtColor to gray, undistortPoints / undistortKeyPoints / computeImageBounds, computeStereoFromRGBD
 through the C ABI vs the CPU oracle (itself pinned to cv2 4.13.0).  Bit-exact, floats compared as
their 32-bit patterns."""
import numpy as np
import pytest

from oracle import orc
from tests import pre_cases
from ydorbslam_b200 import OrbExtractor, OrbMatcher, preprocess, synth

pytestmark = pytest.mark.gpu


def _bits(a):
    return np.ascontiguousarray(a, np.float32).view(np.uint32)


@pytest.mark.parametrize("w,h", [(640, 480), (61, 23), (1, 1), (3, 2), (1241, 376), (1922, 7)])
@pytest.mark.parametrize("code", [0, 1, 2, 3])
def test_cvt_gray(w, h, code):
    img = pre_cases.color_image(w, h, 4 if code >= 2 else 3, 10 + code)
    m = OrbMatcher()
    assert np.array_equal(preprocess.cvtColor(m, img, code), orc.cvt_gray(img, code))


def test_cvt_gray_batch_device_feeds_extractor():
    """Device-resident chain: RGB frames -> k_cvt_gray -> ydorb_extract_batch_device, equal to extracting the oracle's
    gray images."""
    import torch

    w, h, B = 320, 240, 3
    gray = np.stack([synth.synth_frame(w, h, 60 + b) for b in range(B)])
    # an RGB image whose gray conversion is non-trivial: channels are shifted copies
    rgb = np.stack([gray, np.roll(gray, 3, 2), np.roll(gray, 5, 1)], -1)
    dev = torch.device("cuda:0")
    m = OrbMatcher()
    d_rgb = torch.from_numpy(rgb).to(dev)
    d_gray = torch.zeros((B, h, w), dtype=torch.uint8, device=dev)
    torch.cuda.synchronize()
    from ydorbslam_b200._lib import check, lib
    import ctypes as C

    check(lib().ydorb_cvt_gray_device(m._h, C.c_void_p(d_rgb.data_ptr()), w, h, w * 3, C.c_size_t(w * h * 3), 1,
                                      C.c_void_p(d_gray.data_ptr()), w, C.c_size_t(w * h), B))
    m.sync()
    got_gray = d_gray.cpu().numpy()
    for b in range(B):
        assert np.array_equal(got_gray[b], orc.cvt_gray(rgb[b], 1))
    ex = OrbExtractor(500, 1.2, 8, 20, 7, max_width=w, max_height=h, max_batch=B)
    kps, desc, n = ex.extract_batch(got_gray)
    p = orc.make_params(500, 1.2, 8, 20)
    for b in range(B):
        ok, od, _ = orc.extract(p, orc.cvt_gray(rgb[b], 1))
        assert n[b] == len(ok) and np.array_equal(desc[b, : n[b]], od)


@pytest.mark.parametrize("name", sorted(pre_cases.DISTS))
def test_undistort_points(name):
    dist = pre_cases.DISTS[name]
    m = OrbMatcher()
    for n, spread, P in [(20000, 8.0 if name == "barrel" else 1.3, pre_cases.K), (1, 1.0, pre_cases.P_OTHER), (257, 1.0, pre_cases.P_OTHER)]:
        pts = pre_cases.points(n, 11 + n, spread)
        got, ref = preprocess.undistortPoints(m, pts, pre_cases.K, dist, P), orc.undistort_points(pts, pre_cases.K, dist, P)
        same = (_bits(got) == _bits(ref)) | (np.isnan(got) & np.isnan(ref))
        assert same.all(), int((~same).sum())
    assert len(preprocess.undistortPoints(m, np.zeros((0, 2), np.float32), pre_cases.K, dist, pre_cases.K)) == 0


def test_undistort_keypoints_and_bounds():
    m = OrbMatcher()
    kps = pre_cases.keypoints_at(pre_cases.points(3000, 21))
    dist = pre_cases.DISTS["tum5"]
    got = preprocess.undistortKeyPoints(m, kps, pre_cases.K, dist, pre_cases.K)
    ref = orc.undistort_points(np.stack([kps["x"], kps["y"]], 1), pre_cases.K, dist, pre_cases.K)
    assert np.array_equal(_bits(got["x"]), _bits(ref[:, 0])) and np.array_equal(_bits(got["y"]), _bits(ref[:, 1]))
    # dist[0] == 0: untouched (src/frame.cpp:194)
    zero = np.zeros(5, np.float32)
    same = preprocess.undistortKeyPoints(m, kps, pre_cases.K, zero, pre_cases.K)
    assert np.array_equal(same, kps)
    assert preprocess.computeImageBounds(m, 640, 480, pre_cases.K, zero) == (0.0, 640.0, 0.0, 480.0)
    b = preprocess.computeImageBounds(m, 640, 480, pre_cases.K, dist)
    c = orc.undistort_points(np.array([[0, 0], [640, 0], [0, 480], [640, 480]], np.float32), pre_cases.K, dist, pre_cases.K)
    assert b == (float(min(c[0, 0], c[2, 0])), float(max(c[1, 0], c[3, 0])), float(min(c[0, 1], c[1, 1])), float(max(c[2, 1], c[3, 1])))


def test_undistort_keypoints_device_in_place():
    import ctypes as C

    import torch

    from ydorbslam_b200._lib import KP_DTYPE, check, lib

    m = OrbMatcher()
    B, cap = 3, 500
    ns = np.array([500, 0, 123], np.int32)
    kps = np.zeros((B, cap), KP_DTYPE)
    for b in range(B):
        kps[b] = pre_cases.keypoints_at(pre_cases.points(cap, 30 + b))
    dev = torch.device("cuda:0")
    d_k = torch.from_numpy(kps.view(np.uint8).reshape(B, cap, 28)).to(dev)
    d_n = torch.from_numpy(ns).to(dev)
    torch.cuda.synchronize()
    dist = pre_cases.DISTS["tum5"]
    Kf = np.ascontiguousarray(pre_cases.K, np.float32)
    check(lib().ydorb_undistort_keypoints_device(m._h, C.c_void_p(d_k.data_ptr()), C.c_void_p(d_n.data_ptr()), cap, B,
                                                 Kf.ctypes.data_as(C.c_void_p), dist.ctypes.data_as(C.c_void_p), len(dist),
                                                 Kf.ctypes.data_as(C.c_void_p)))
    m.sync()
    got = d_k.cpu().numpy().view(KP_DTYPE).reshape(B, cap)
    for b in range(B):
        ref = orc.undistort_points(np.stack([kps[b]["x"], kps[b]["y"]], 1), pre_cases.K, dist, pre_cases.K)
        assert np.array_equal(_bits(got[b]["x"][: ns[b]]), _bits(ref[: ns[b], 0]))
        assert np.array_equal(_bits(got[b]["y"][: ns[b]]), _bits(ref[: ns[b], 1]))
        assert np.array_equal(got[b][ns[b] :], kps[b][ns[b] :])  # rows past n untouched
        assert np.array_equal(got[b]["size"], kps[b]["size"])


def test_stereo_from_rgbd():
    w, h = 640, 480
    d16 = pre_cases.depth_map_u16(w, h, 3)
    kps = pre_cases.keypoints_at(pre_cases.points(8000, 4, spread=1.1))
    f, bf = np.float32(1.0 / 5000.0), np.float32(40.0)
    m = OrbMatcher()
    for dm, factor in [(d16, f), (d16.astype(np.float32), f), (d16.astype(np.float32) * f, np.float32(1.0)),
                       (d16[:, :333].copy(), f)]:
        rx, dz = preprocess.computeStereoFromRGBD(m, kps, dm, factor, bf)
        orx, odz = orc.stereo_from_rgbd(kps, dm, factor, bf)
        assert np.array_equal(_bits(rx), _bits(orx)) and np.array_equal(_bits(dz), _bits(odz))
        assert (dz > 0).sum() > 1000 and (dz == -1).sum() > 100
    rx, dz = preprocess.computeStereoFromRGBD(m, kps[:0], d16, f, bf)
    assert len(rx) == 0
