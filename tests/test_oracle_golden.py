"""CPU: pin the C++ oracle against the committed cv2-generated golden vectors (oracle/gen_golden.py)."""
import hashlib

import numpy as np
import pytest

from oracle import orc
from ydorbslam_b200 import synth


def sha(a):
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


def _input(name, g):
    if "input" in g.files:
        return g["input"]
    h, w = (int(v) for v in g["shape"])
    seed = {"synth640": 1, "sparse640": 2, "synth1241": 3, "tiny97": 4}[name]
    img = synth.synth_sparse_frame(w, h, seed) if name.startswith("sparse") else synth.synth_frame(w, h, seed)
    assert sha(img) == str(g["input_sha"]), "synthetic generator drifted from the golden fixtures"
    return img


CASES = ["image0", "synth640", "sparse640", "synth1241", "tiny97"]


def test_ctor_tables():
    p = orc.make_params(1000, 1.2, 8, 20, 7)
    assert list(p.quota)[:8] == [217, 181, 151, 126, 105, 88, 73, 59]
    assert list(p.umax) == [0] * 11 + [26, 25, 22, 19, 15]
    assert [orc.level_size(p, 640, 480, l) for l in range(8)] == [
        (640, 480), (533, 400), (444, 333), (370, 278), (309, 231), (257, 193), (214, 161), (179, 134)]
    assert list(orc.make_params(2000, 1.2, 8, 20).quota)[:8] == [434, 362, 302, 252, 210, 175, 146, 119]
    assert list(orc.make_params(4000, 1.2, 8, 20).quota)[:8] == [869, 724, 603, 502, 418, 348, 290, 246]
    assert list(orc.make_params(8000, 1.2, 8, 20).quota)[:8] == [1737, 1448, 1207, 1006, 838, 698, 582, 484]


def test_primitives(golden_dir):
    g = np.load(golden_dir / "primitives.npz")
    got = np.array([orc.fast_atan2(y, x) for y, x in zip(g["atan_y"], g["atan_x"])], dtype=np.float32)
    assert np.array_equal(got.view(np.uint32), g["atan_deg"].view(np.uint32))
    img = synth.synth_frame(321, 243, 11)
    r1 = orc.resize_linear(img, 268, 203)
    assert np.array_equal(r1, g["resize1"])
    assert np.array_equal(orc.resize_linear(r1, 223, 169), g["resize2"])
    assert np.array_equal(orc.gauss7(img), g["blur"])
    xs, ys, rs = orc.fast(img, 20)
    assert np.array_equal(np.stack([xs, ys, rs], 1).astype(np.float32), g["fast20"])
    xs, ys, rs = orc.fast(np.ascontiguousarray(img[40:77, 100:137]), 7)
    assert np.array_equal(np.stack([xs, ys, rs], 1).astype(np.float32), g["fast7_roi"])


@pytest.mark.parametrize("name", CASES)
def test_extract_stages(golden_dir, name):
    g = np.load(golden_dir / f"extract_{name}.npz")
    img = _input(name, g)
    nf = int(g["n_features"])
    p = orc.make_params(nf, 1.2, 8, 20)
    pyr, _ = orc.pyramid(p, img)
    for l in range(8):
        assert sha(pyr[l]) == str(g[f"pyr_sha_{l}"]), f"pyramid level {l}"
        c = orc.level_candidates(pyr[l], 20, 20)
        gc = g[f"cand_{l}"]
        assert len(c) == len(gc)
        if len(c):
            assert np.array_equal(np.stack([c["x"], c["y"], c["response"]], 1).astype(np.float32), gc), f"candidates {l}"
        lh, lw = pyr[l].shape[0] - 38, pyr[l].shape[1] - 38
        for form in ("list", "sortkey"):
            s = orc.quadtree(c, lw - 32, lh - 32, p.quota[l], form)
            assert np.array_equal(s, g[f"sel_{l}"]), f"quadtree {form} level {l}"
        if f"blur_sha_{l}" in g.files:
            assert sha(orc.gauss7(pyr[l][19:-19, 19:-19])) == str(g[f"blur_sha_{l}"]), f"blur {l}"


@pytest.mark.parametrize("name", CASES)
def test_extract_end_to_end(golden_dir, name):
    g = np.load(golden_dir / f"extract_{name}.npz")
    img = _input(name, g)
    p = orc.make_params(int(g["n_features"]), 1.2, 8, 20)
    kps, desc, _ = orc.extract(p, img, cap=int(g["n_features"]) + 64)
    gk = g["kps"]
    assert len(kps) == len(gk)
    for f in orc.KP_DTYPE.names:
        assert np.array_equal(kps[f], gk[f]), f
    assert np.array_equal(desc, g["desc"])


def test_empty_and_flat():
    p = orc.make_params(500, 1.2, 8, 20)
    kps, desc, per = orc.extract(p, np.full((120, 160), 77, np.uint8))
    assert len(kps) == 0 and desc.shape == (0, 32)
