"""GPU parity: CUDA extractor (through the C ABI) vs cv2-generated golden fixtures and vs the CPU oracle."""
import numpy as np
import pytest

from oracle import orc
from tests.util import golden_input, sha
from ydorbslam_b200 import OrbExtractor, synth

pytestmark = pytest.mark.gpu

CASES = ["image0", "synth640", "sparse640", "synth1241", "tiny97"]


@pytest.fixture(autouse=True, params=["cells", "blocks"])
def fast_kernel(request, monkeypatch):
    """Every test of this module runs under both FAST kernels: warp-per-cell (fast.cu, what small launches take) and
    CTA-per-cell-block with TMA staging (fast2.cu, what batches take).  libydorb reads the variable on every call."""
    monkeypatch.setenv("YDORB_FAST_KERNEL", request.param)
    return request.param


def _check_against(ex, img, kps, desc, gk, gd):
    assert len(kps) == len(gk)
    for f in ("x", "y", "size", "response", "octave", "class_id"):
        assert np.array_equal(kps[f], gk[f]), f
    # angles: bit-exact expected; north_star tolerance is 1e-4 rad
    dang = np.abs(kps["angle"] - gk["angle"])
    dang = np.minimum(dang, 360.0 - dang)
    assert float(dang.max(initial=0.0)) * np.pi / 180.0 <= 1e-4
    bits = np.unpackbits(desc ^ gd).sum()
    assert bits <= 0.001 * desc.size * 8, f"{bits} descriptor bits differ"
    return int((kps["angle"].view(np.uint32) != gk["angle"].view(np.uint32)).sum()), int(bits)


@pytest.mark.parametrize("name", CASES)
def test_extract_matches_golden(golden_dir, name):
    g = np.load(golden_dir / f"extract_{name}.npz")
    img = golden_input(name, g)
    nf = int(g["n_features"])
    h, w = img.shape
    ex = OrbExtractor(nf, 1.2, 8, 20, 7, max_width=w, max_height=h, max_batch=1)
    kps, desc = ex.extractAndCompute(img)
    # stages
    for l in range(8):
        assert sha(ex.pyramid_level(0, l, padded=True)) == str(g[f"pyr_sha_{l}"]), f"pyramid level {l}"
        xy, resp = ex.debug_candidates(0, l)
        gc = g[f"cand_{l}"]
        assert len(xy) == len(gc), f"candidate count level {l}: {len(xy)} vs {len(gc)}"
        if len(gc):
            assert np.array_equal(np.concatenate([xy, resp[:, None]], 1).astype(np.float32), gc), f"candidates level {l}"
    nang, nbits = _check_against(ex, img, kps, desc, g["kps"], g["desc"])
    assert nang == 0 and nbits == 0, (nang, nbits)  # bit-exact on the fixtures
    ex.close()


@pytest.mark.parametrize("shape,nf,seed", [((480, 640), 1000, 21), ((376, 1241), 2000, 22), ((1080, 1920), 4000, 23),
                                           ((333, 517), 700, 24)])
def test_extract_matches_oracle(shape, nf, seed):
    h, w = shape
    img = synth.synth_frame(w, h, seed)
    p = orc.make_params(nf, 1.2, 8, 20)
    ok, od, _ = orc.extract(p, img, cap=nf + 64)
    ex = OrbExtractor(nf, 1.2, 8, 20, 7, max_width=w, max_height=h, max_batch=1)
    kps, desc = ex.extractAndCompute(img)
    _check_against(ex, img, kps, desc, ok, od)
    ex.close()


@pytest.mark.parametrize("scale,levels,shape", [(1.5, 5, (480, 640)), (2.0, 4, (376, 1241)), (1.1, 6, (333, 517))])
def test_extract_other_scale_factors(scale, levels, shape):
    """The resize tiles size their shared buffers from the scale factor; check pyramids and results away from 1.2."""
    h, w = shape
    img = synth.synth_frame(w, h, 77)
    p = orc.make_params(800, scale, levels, 20)
    ok, od, _ = orc.extract(p, img, cap=900)
    ex = OrbExtractor(800, scale, levels, 20, 7, max_width=w, max_height=h, max_batch=1)
    kps, desc = ex.extractAndCompute(img)
    _check_against(ex, img, kps, desc, ok, od)
    ex.close()


def test_batch_and_reuse():
    """Batch of different frames == frame-by-frame; a reused handle gives fresh-extractor results."""
    imgs = synth.synth_batch(640, 480, 5, seed0=100)
    imgs[3] = synth.synth_sparse_frame(640, 480, 9)
    imgs[4] = 50  # flat: no keypoints
    ex = OrbExtractor(1000, 1.2, 8, 20, 7, max_width=640, max_height=480, max_batch=5)
    kps, desc, n = ex.extract_batch(imgs)
    p = orc.make_params(1000, 1.2, 8, 20)
    for b in range(5):
        ok, od, _ = orc.extract(p, imgs[b])
        assert n[b] == len(ok), (b, n[b], len(ok))
        for f in ("x", "y", "size", "response", "octave", "angle"):
            assert np.array_equal(kps[b, : n[b]][f], ok[f]), (b, f)
        assert np.array_equal(desc[b, : n[b]], od)
    assert n[4] == 0
    # second call on the same handle with another size
    img2 = synth.synth_frame(320, 240, 5)
    k2, d2 = ex.extractAndCompute(img2)
    ok, od, _ = orc.extract(p, img2)
    assert np.array_equal(k2["x"], ok["x"]) and np.array_equal(d2, od)
    ex.close()


def test_min_threshold_fallback():
    """True two-threshold fallback (library) vs reference quirk min:=ini."""
    img = (synth.synth_frame(640, 480, 31, contrast=1) // 2 + 64).astype(np.uint8)  # low contrast: many empty cells at th=20
    for quirk in (True, False):
        p = orc.make_params(1000, 1.2, 8, 20, 20 if quirk else 7)
        ok, od, _ = orc.extract(p, img)
        ex = OrbExtractor(1000, 1.2, 8, 20, 7, max_width=640, max_height=480, reference_min_th_quirk=quirk)
        k, d = ex.extractAndCompute(img)
        assert len(k) == len(ok)
        assert np.array_equal(k["x"], ok["x"]) and np.array_equal(k["y"], ok["y"]) and np.array_equal(d, od)
        ex.close()


def test_empty_image_and_errors():
    ex = OrbExtractor(500, 1.2, 8, 20, 7, max_width=320, max_height=240)
    k, d = ex.extractAndCompute(np.zeros((0, 0), np.uint8))
    assert len(k) == 0 and d.shape == (0, 32)
    from ydorbslam_b200 import YdorbError

    with pytest.raises(YdorbError):
        ex.extractAndCompute(np.zeros((480, 640), np.uint8))  # larger than max
    ex.close()


@pytest.mark.parametrize("patch,nf", [(48, 1000), (90, 2000), (24, 500)])
def test_clustered_corners_deep_quadtree(patch, nf):
    """All corners inside one small textured patch of a flat image: far fewer candidates than the quota, so the
    reference keeps splitting until every point is alone (more than 8 passes: the sort-based deep path of the quadtree),
    while the upper pyramid levels, where the patch shrinks, take the node-table path."""
    img = np.full((480, 640), 90, np.uint8)
    tex = synth.synth_frame(patch, patch, 123 + patch)
    img[200 : 200 + patch, 300 : 300 + patch] = tex
    img[37 : 37 + patch // 2, 41 : 41 + patch // 2] = tex[: patch // 2, : patch // 2]
    ex = OrbExtractor(nf, 1.2, 8, 20, 7, max_width=640, max_height=480)
    kps, desc = ex.extractAndCompute(img)
    ok, od, _ = orc.extract(orc.make_params(nf, 1.2, 8, 20), img)
    assert len(kps) == len(ok) and len(kps) > 20
    assert np.array_equal(kps, ok) and np.array_equal(desc, od)


@pytest.mark.gpu
def test_pinned_host_buffers_same_result():
    """ydorb_host_alloc buffers (include/ydorb.h) through the host entry point give what pageable arrays give."""
    from ydorbslam_b200 import PinnedArray

    imgs = np.stack([synth.synth_frame(640, 480, 90 + i) for i in range(8)])
    ex = OrbExtractor(1000, 1.2, 8, 20, 7, max_width=640, max_height=480, max_batch=8)
    kps, desc, n = ex.extract_batch(imgs)
    pin = PinnedArray(imgs.shape)
    pin.array[...] = imgs
    kps2, desc2, n2 = ex.extract_batch(pin.array)
    assert np.array_equal(n, n2) and np.array_equal(kps, kps2) and np.array_equal(desc, desc2)
    pin.close()
    ex.close()


_KNOB_SCRIPT = r"""
import hashlib, sys
import numpy as np
from ydorbslam_b200 import OrbExtractor, OrbMatcher, synth
left, right = synth.synth_stereo_pair(1241, 376, 64)
ex = OrbExtractor(2000, 1.2, 8, 20, 7, max_width=1241, max_height=376, max_batch=2)
kps, desc, n = ex.extract_batch(np.stack([left, right]))
m = OrbMatcher()
rx, dp = m.computeStereoMatches(ex, 0, ex, 1, kps[0, :n[0]], desc[0, :n[0]], kps[1, :n[1]], desc[1, :n[1]], 386.1448, np.inf, 0)
h = hashlib.sha256()
for a in (n, kps[0, :n[0]], kps[1, :n[1]], desc[0, :n[0]], desc[1, :n[1]], rx, dp):
    h.update(np.ascontiguousarray(a).tobytes())
print("HASH", h.hexdigest())
"""


def test_tuning_knobs_same_result():
    """The launch-shape knobs (cells / key points per warp, programmatic dependent launch, general stereo chain) change
    how the work is spread, never the result."""
    import os
    import subprocess
    import sys

    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

    def run(env):
        e = dict(os.environ, PYTHONPATH=root, **env)
        out = subprocess.run([sys.executable, "-c", _KNOB_SCRIPT], env=e, cwd=root, capture_output=True, text=True, timeout=600)
        assert out.returncode == 0, out.stderr[-2000:]
        return [l for l in out.stdout.splitlines() if l.startswith("HASH")][0]

    base = run({})
    assert run({"YDORB_PDL": "0", "YDORB_FAST_CPW": "4", "YDORB_DESC_KPW": "8", "YDORB_STEREO_CHAIN": "1"}) == base
    assert run({"YDORB_FAST_CPW": "2", "YDORB_DESC_KPW": "2", "YDORB_DEVICE_SPLIT": "1"}) == base
