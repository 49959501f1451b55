"""CPU: the parity chain's top link.  oracle/_ref is the REFERENCE's own orbExtractor.cpp compiled unmodified
(oracle/refshim/Makefile); these tests assert  _ref == oracle (restatement) == committed golden fixtures  on every
stage the reference exposes: ctor tables, pyramid, raw per-cell FAST candidates (order included), the std::list
quadtree with libstdc++ std::sort ties, angles, descriptors, key point order.
"""
import ctypes as C
import os
import struct

import numpy as np
import pytest

from oracle import orc, ref
from tests.util import golden_input, sha
from ydorbslam_b200 import synth

pytestmark = pytest.mark.skipif(not ref.available(), reason="oracle/_ref not built and /root/reference absent")

CASES = ["image0", "synth640", "sparse640", "synth1241", "tiny97"]


def _same_kps(a, b):
    assert len(a) == len(b)
    for f in orc.KP_DTYPE.names:
        assert np.array_equal(a[f].view(np.uint32), b[f].view(np.uint32)), f


@pytest.mark.parametrize("nf,scale,nl,ini,mn", [(1000, 1.2, 8, 20, 7), (2000, 1.2, 8, 20, 7), (4000, 1.2, 8, 20, 7),
                                                 (8000, 1.2, 8, 20, 7), (500, 1.5, 5, 12, 5), (800, 2.0, 4, 20, 7),
                                                 (700, 1.1, 6, 30, 10), (300, 1.3, 12, 20, 7)])
def test_ctor_tables(nf, scale, nl, ini, mn):
    """E0 src/orbExtractor.cpp:315-354, incl. F3 (min := ini) and the F4 half-width table."""
    t = ref.ctor_tables(nf, scale, nl, ini, mn)
    p = orc.make_params(nf, scale, nl, ini)
    for name in ("sf", "inv_sf", "sf2", "inv_sf2"):
        assert np.array_equal(t[name].view(np.uint32), np.array(getattr(p, name)[:nl], np.float32).view(np.uint32)), name
    assert list(t["quota"]) == list(p.quota)[:nl]
    assert list(t["umax"][:16]) == list(p.umax)
    assert len(t["umax"]) == 16 + 12  # resize(16) then 12 push_backs: the table the reference really builds (F4)
    assert t["ini_th"] == ini and t["min_th"] == ini  # F3


@pytest.mark.parametrize("name", CASES)
def test_reference_matches_golden_and_oracle(golden_dir, name):
    g = np.load(golden_dir / f"extract_{name}.npz")
    img = golden_input(name, g)
    nf = int(g["n_features"])
    h, w = img.shape
    rk, rd, per, pyr, cands = ref.extract(img, nf, cap=nf + 64, want_pyramid=True, want_candidates=True)
    p = orc.make_params(nf, 1.2, 8, 20)
    sizes = [orc.level_size(p, w, h, l) for l in range(8)]
    levels = ref.pyramid_levels(pyr, sizes)
    opyr, _ = orc.pyramid(p, img)
    for l in range(8):
        assert sha(levels[l]) == str(g[f"pyr_sha_{l}"]), f"pyramid level {l} vs golden"
        assert np.array_equal(levels[l], opyr[l]), f"pyramid level {l} vs oracle"
        gc = g[f"cand_{l}"]
        c = cands[l]
        assert len(c) == len(gc), f"candidate count level {l}"
        if len(c):
            assert np.array_equal(np.stack([c["x"], c["y"], c["response"]], 1).astype(np.float32), gc), f"candidates {l}"
        oc = orc.level_candidates(opyr[l], 20, 20)
        assert np.array_equal(c, oc), f"candidates level {l} vs oracle"
        # E4 in isolation: the reference's distributeQuadTree on the golden candidate list
        lw, lh = sizes[l]
        s = ref.quadtree(c, lw - 32, lh - 32, int(p.quota[l]))
        assert np.array_equal(s, g[f"sel_{l}"]), f"quadtree level {l}"
    _same_kps(rk, g["kps"])
    assert np.array_equal(rd, g["desc"])
    ok, od, oper = orc.extract(p, img, cap=nf + 64)
    _same_kps(rk, ok)
    assert np.array_equal(rd, od)
    assert np.array_equal(per, oper)


@pytest.mark.parametrize("shape,nf,seed", [((480, 640), 1000, 21), ((376, 1241), 2000, 22), ((1080, 1920), 4000, 23),
                                           ((2160, 3840), 8000, 6), ((333, 517), 700, 24), ((97, 97), 1000, 4)])
def test_reference_matches_oracle_on_config_shapes(shape, nf, seed):
    """BASELINE configs 1, 2, 3, 5 (+ an odd size and a tiny image): key points, order, angles, descriptors."""
    h, w = shape
    img = synth.synth_frame(w, h, seed)
    rk, rd, per = ref.extract(img, nf, cap=nf + 64)
    ok, od, oper = orc.extract(orc.make_params(nf, 1.2, 8, 20), img, cap=nf + 64)
    _same_kps(rk, ok)
    assert np.array_equal(rd, od)
    assert np.array_equal(per, oper)


@pytest.mark.parametrize("scale,levels,ini", [(1.5, 5, 20), (2.0, 4, 20), (1.1, 6, 20), (1.2, 8, 7), (1.2, 8, 40)])
def test_reference_matches_oracle_other_parameters(scale, levels, ini):
    img = synth.synth_frame(640, 480, 77)
    rk, rd, per = ref.extract(img, 800, scale, levels, ini, 7, cap=900)
    ok, od, oper = orc.extract(orc.make_params(800, scale, levels, ini), img, cap=900)
    _same_kps(rk, ok)
    assert np.array_equal(rd, od)


def test_reference_sparse_flat_and_empty():
    for img in (synth.synth_sparse_frame(640, 480, 9), np.full((120, 160), 77, np.uint8)):
        rk, rd, _ = ref.extract(img, 500)
        ok, od, _ = orc.extract(orc.make_params(500, 1.2, 8, 20), img)
        _same_kps(rk, ok)
        assert np.array_equal(rd, od)


@pytest.mark.parametrize("seed", range(6))
def test_quadtree_ties_against_reference(seed):
    """F5/F6/F7: random candidate clouds with few distinct responses (tied maxima inside > 16-point nodes decide through
    libstdc++ introsort), clustered so that some nodes stay large, duplicates in x so that F6 lineages never separate."""
    rng = np.random.default_rng(seed)
    rootw, rooth = 608, 448
    n = int(rng.integers(300, 6000))
    pts = set()
    while len(pts) < n:
        if rng.random() < 0.6:
            cx, cy = rng.integers(0, rootw), rng.integers(0, rooth)
            x = int(np.clip(cx + rng.normal(0, 6), 0, rootw - 1))
            y = int(np.clip(cy + rng.normal(0, 6), 0, rooth - 1))
        else:
            x, y = int(rng.integers(0, rootw)), int(rng.integers(0, rooth))
        pts.add((y, x))
    pts = sorted(pts)  # any fixed order is a legal candidate order for the quadtree
    cand = np.zeros(len(pts), orc.CAND_DTYPE)
    cand["y"] = [p[0] for p in pts]
    cand["x"] = [p[1] for p in pts]
    cand["response"] = rng.integers(20, 20 + int(rng.integers(2, 12)), len(pts)).astype(np.float32)
    for want in (1, 2, 59, 217, 1000, len(pts) + 5):
        r = ref.quadtree(cand, rootw, rooth, want)
        assert np.array_equal(r, orc.quadtree(cand, rootw, rooth, want, "list")), want
        assert np.array_equal(r, orc.quadtree(cand, rootw, rooth, want, "sortkey")), want


def test_sincosf_model_is_libm_sincosf():
    """src/orbExtractor.cpp:426 resolves to sincosf (float overloads).  orc_sincosf_model -- the restatement of glibc's
    algorithm that describe.cu follows -- must equal libm on EVERY float an angle can be: [0, 2*pi] (and beyond, < 120)."""
    l = orc.lib()
    l.orc_sincosf_model_mismatches.restype = C.c_int64
    l.orc_sincosf_model_mismatches.argtypes = [C.c_uint32, C.c_uint32, C.c_uint32, C.c_int]
    hi = struct.unpack("<I", struct.pack("<f", 6.2831855))[0]
    assert l.orc_sincosf_model_mismatches(0, hi, 1, os.cpu_count() or 4) == 0
    hi2 = struct.unpack("<I", struct.pack("<f", 119.99))[0]
    assert l.orc_sincosf_model_mismatches(hi, hi2, 7, os.cpu_count() or 4) == 0


def test_reference_object_calls_sincosf():
    """What the compiled reference really calls for `(float)cos(angle)`: the float routine, not double cos/sin."""
    import pathlib
    import subprocess

    obj = pathlib.Path(ref._SO).parent / "orbExtractor.o"
    if not obj.exists():
        pytest.skip("object file not shipped")
    syms = subprocess.run(["nm", "-u", str(obj)], capture_output=True, text=True).stdout.split()
    assert "sincosf" in syms and "cos" not in syms and "sin" not in syms and "sincos" not in syms
