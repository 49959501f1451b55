"""CPU: the oracle's M2-M7 restatements against an independent plain-Python reading of the same reference lines
(small cases), plus known-answer tests of computeThreeMaxima.  No GPU, no libydorb compute calls."""
import numpy as np
import pytest

from oracle import orc
from tests import matcher_cases as mc


def _popcount(a, b):
    return int(np.unpackbits(np.bitwise_xor(a, b)).sum())


def _bin(a, b, add_f32=False):
    rot = np.float32(a) - np.float32(b)
    if rot < 0:
        rot = np.float32(rot + np.float32(360.0)) if add_f32 else np.float32(np.float64(rot) + 360.0)
    v = float(np.float32(rot * np.float32(1.0 / 30)))
    b_ = int(np.floor(abs(v) + 0.5) * np.sign(v))  # C round(): half away from zero
    return 0 if b_ == 30 else b_


def _three_maxima_py(sizes):
    """src/orbMatcher.cpp:827-854, read independently of the oracle."""
    m1 = m2 = m3 = 0
    i1 = i2 = i3 = -1
    for i, n in enumerate(sizes):
        if n > m1:
            m3, m2, m1 = m2, m1, n
            i3, i2, i1 = i2, i1, i
        elif n > m2:
            m3, m2 = m2, n
            i3, i2 = i2, i
        elif n > m3:
            m3, i3 = n, i
    if m2 < int(m1 / 10):
        i2 = i3 = -1
    elif m3 < int(m1 / 10):
        i3 = -1
    return i1, i2, i3


@pytest.mark.parametrize("sizes,expect", [
    ([0] * 30, (-1, -1, -1)),
    ([5] + [0] * 29, (0, -1, -1)),                       # max2 = 0 < 5/10 = 0 is false -> idx2 stays -1 anyway
    ([100, 9, 8] + [0] * 27, (0, -1, -1)),               # second < max/10 clears both
    ([100, 10, 9] + [0] * 27, (0, 1, -1)),               # third < max/10 clears only the third
    ([3, 3, 3, 3] + [0] * 26, (0, 1, 2)),                # ties: strict '>' keeps the earliest bins
    ([1, 2, 3, 4] + [0] * 26, (3, 2, 1)),
])
def test_three_maxima_known_answers(sizes, expect):
    assert orc.three_maxima(sizes) == expect
    assert _three_maxima_py(sizes) == expect


def test_three_maxima_random():
    rng = np.random.RandomState(0)
    for _ in range(300):
        sizes = rng.randint(0, rng.choice([3, 20, 200]), 30).tolist()
        assert orc.three_maxima(sizes) == _three_maxima_py(sizes)


def _window_search_py(g, kps, desc, right_x, occupied, q, qd, prm):
    """Plain-Python reading of the loop shared by src/orbMatcher.cpp:82-153 / 169-237 / 251-300 / 593-630 / 687-743."""
    n = len(kps)
    slot = [-2 if (occupied is not None and occupied[i]) else -1 for i in range(n)]
    hist = [[] for _ in range(30)]
    best_of = np.full(len(q), -1, np.int32)
    matches = 0
    for j in range(len(q)):
        if not (q["flags"][j] & 1):
            continue
        cands = orc.keypoints_in_area(g, kps, q["x"][j], q["y"][j], q["r"][j], int(q["min_level"][j]), int(q["max_level"][j]))
        best, best_idx = 256, -1
        for idx in cands:
            if prm.exclusive and (slot[idx] == -2 or (slot[idx] >= 0 and q["flags"][slot[idx]] & 2)):
                continue
            if q["right_r"][j] >= 0 and not (right_x[idx] <= 0 or abs(np.float32(q["right_x"][j] - right_x[idx])) <= q["right_r"][j]):
                continue
            o = kps["octave"][idx]
            if o < q["oct_lo"][j] or o > q["oct_hi"][j]:
                continue
            if prm.chi2_gate:
                dx = float(np.float32(kps["x"][idx] - q["x"][j]))
                dy = float(np.float32(kps["y"][idx] - q["y"][j]))
                mono = np.float32(dx * dx + dy * dy)
                dr = float(np.float32(right_x[idx] - q["right_x"][j]))
                stereo = np.float32(float(mono) + dr * dr)
                w = np.float32(prm.inv_sf2[o])
                ok = (right_x[idx] >= 0 and float(np.float32(stereo * w)) <= 7.81) or \
                     (right_x[idx] < 0 and float(np.float32(mono * w)) <= 5.99)
                if not ok:
                    continue
            d = _popcount(qd[j], desc[idx])
            if d < best:
                best, best_idx = d, idx
        ok = best < prm.max_dist if prm.strict else best <= prm.max_dist
        if not ok or best_idx < 0:
            continue
        best_of[j] = best_idx
        slot[best_idx] = j
        matches += 1
        if prm.check_orientation:
            hist[_bin(q["angle"][j], kps["angle"][best_idx])].append(best_idx)
    if prm.check_orientation:
        keep = _three_maxima_py([len(h) for h in hist])
        for b in range(30):
            if b not in keep:
                for idx in hist[b]:
                    slot[idx] = -3
                    matches -= 1
    match = np.array([s if s >= 0 else (-2 if s == -3 else -1) for s in slot], np.int32)
    return matches, best_of, match


def test_window_search_oracle_vs_python():
    kps, desc = mc.synth_keypoints(400, 3)
    p = orc.make_params(1000, 1.2, 8, 20)
    g = orc.make_geom(p, mc.W, mc.H)
    cases = mc.all_window_cases(kps, desc, 600, 11)
    seen_matches = 0
    for name, q, prm, occ, mp in cases:
        n, best, match = orc.search_windows(g, kps, desc, mp["right_x_kp"], occ, q, mp["desc"], prm)
        pn, pbest, pmatch = _window_search_py(g, kps, desc, mp["right_x_kp"], occ, q, mp["desc"], prm)
        assert n == pn, name
        assert np.array_equal(best, pbest), name
        assert np.array_equal(match, pmatch), name
        assert n > 10, name
        seen_matches += n
    assert seen_matches > 100


def _bow_py(mode, s, ratio, check_ori):
    """Plain-Python reading of src/orbMatcher.cpp:303-379 / 380-462 using dicts for the FeatureVectors."""
    fv1 = {int(i): s["fv1"][2][s["fv1"][1][k]:s["fv1"][1][k + 1]].tolist() for k, i in enumerate(s["fv1"][0])}
    fv2 = {int(i): s["fv2"][2][s["fv2"][1][k]:s["fv2"][1][k + 1]].tolist() for k, i in enumerate(s["fv2"][0])}
    n_out = len(s["k2"]) if mode == 0 else len(s["k1"])
    out = np.full(n_out, -1, np.int32)
    matched2 = np.zeros(len(s["k2"]), bool)
    hist = [[] for _ in range(30)]
    matches = 0
    for node in sorted(set(fv1) & set(fv2)):
        for i1 in fv1[node]:
            if not s["v1"][i1]:
                continue
            best = second = 256
            bi = -1
            for i2 in fv2[node]:
                if matched2[i2] or (mode == 1 and not s["v2"][i2]):
                    continue
                d = _popcount(s["d1"][i1], s["d2"][i2])
                if d < best:
                    second, best, bi = best, d, i2
                elif d < second:
                    second = d
            if best <= 50 and np.float32(best) < np.float32(ratio) * np.float32(second):
                matched2[bi] = True
                slot = bi if mode == 0 else i1
                out[slot] = i1 if mode == 0 else bi
                if check_ori:
                    hist[_bin(s["k1"]["angle"][i1], s["k2"]["angle"][bi])].append(slot)
                matches += 1
    if check_ori:
        keep = _three_maxima_py([len(h) for h in hist])
        for b in range(30):
            if b not in keep:
                for slot in hist[b]:
                    out[slot] = -1
                    matches -= 1
    return matches, out


@pytest.mark.parametrize("mode", [0, 1])
@pytest.mark.parametrize("check_ori", [True, False])
def test_bow_oracle_vs_python(mode, check_ori):
    s = mc.bow_scene(300, 350, 5)
    n, out = orc.search_by_bow(mode, s["k1"], s["d1"], s["v1"], s["fv1"], s["k2"], s["d2"], s["v2"], s["fv2"], 0.75, check_ori)
    pn, pout = _bow_py(mode, s, 0.75, check_ori)
    assert n == pn and np.array_equal(out, pout)
    assert n > 20


def test_triangulation_oracle_properties():
    s = mc.bow_scene(300, 350, 6)
    F, ep = mc.fundamental(1)
    has1 = 1 - s["v1"]
    has2 = 1 - s["v2"]
    for stereo_only in (False, True):
        n, pairs = orc.search_for_triangulation(s["k1"], s["d1"], s["rx1"], has1, s["fv1"], s["k2"], s["d2"], s["rx2"], has2,
                                                s["fv2"], F, ep, mc.SF, mc.SF2, stereo_only, True)
        assert n == len(pairs)
        assert len(set(pairs[:, 1].tolist())) == len(pairs)      # a second-side feature is matched once (:497, :515)
        assert not has1[pairs[:, 0]].any() and not has2[pairs[:, 1]].any()
        if stereo_only:
            assert (s["rx1"][pairs[:, 0]] >= 0).all() and (s["rx2"][pairs[:, 1]] >= 0).all()
        for a, b in pairs:
            assert _popcount(s["d1"][a], s["d2"][b]) <= 50
    n_all, _ = orc.search_for_triangulation(s["k1"], s["d1"], s["rx1"], has1, s["fv1"], s["k2"], s["d2"], s["rx2"], has2,
                                            s["fv2"], F, ep, mc.SF, mc.SF2, False, False)
    assert n_all > 10


def _search_for_initialization_py(g2, kps1, desc1, kps2, desc2, prev, window, ratio, th_low, check_ori):
    """Plain-Python reading of ORB-SLAM2's ORBmatcher::SearchForInitialization on the reference's getKeyPointsInArea."""
    n1, n2 = len(kps1), len(kps2)
    m12 = np.full(n1, -1, np.int32)
    m21 = np.full(n2, -1, np.int64)
    mdist = np.full(n2, 2 ** 31 - 1, np.int64)
    hist = [[] for _ in range(30)]
    prev = prev.copy()
    nm = 0
    for i1 in range(n1):
        if kps1["octave"][i1] > 0:
            continue
        cands = orc.keypoints_in_area(g2, kps2, prev[i1, 0], prev[i1, 1], float(window), 0, 0)
        if len(cands) == 0:
            continue
        best, best2, bi = 2 ** 31 - 1, 2 ** 31 - 1, -1
        for i2 in cands:
            d = _popcount(desc1[i1], desc2[i2])
            if mdist[i2] <= d:
                continue
            if d < best:
                best2, best, bi = best, d, i2
            elif d < best2:
                best2 = d
        if best <= th_low and np.float32(best) < np.float32(np.float32(best2) * np.float32(ratio)):
            if m21[bi] >= 0:
                m12[m21[bi]] = -1
                nm -= 1
            m12[i1] = bi
            m21[bi] = i1
            mdist[bi] = best
            nm += 1
            if check_ori:
                hist[_bin(kps1["angle"][i1], kps2["angle"][bi])].append(i1)
    if check_ori:
        keep = _three_maxima_py([len(h) for h in hist])
        for b in range(30):
            if b not in keep:
                for i1 in hist[b]:
                    if m12[i1] >= 0:
                        m12[i1] = -1
                        nm -= 1
    for i1 in range(n1):
        if m12[i1] >= 0:
            prev[i1] = (kps2["x"][m12[i1]], kps2["y"][m12[i1]])
    return nm, m12, prev


@pytest.mark.parametrize("check_ori", [True, False])
def test_search_for_initialization_oracle_vs_python(check_ori):
    """EXTRA (no reference parity): the C++ restatement agrees with a plain-Python reading; contention
    (steals and blocked candidates) does occur in the scene."""
    p = orc.make_params(1000, 1.2, 8, 20)
    g = orc.make_geom(p, mc.W, mc.H)
    kps1, d1, kps2, d2 = mc.init_scene(150, 200, 7, 0)
    prev = np.stack([kps1["x"], kps1["y"]], 1).astype(np.float32)
    n, m12, pout = orc.search_for_initialization(g, kps1, d1, kps2, d2, prev, 30, 0.9, 50, check_ori, 0)
    pn, pm12, pprev = _search_for_initialization_py(g, kps1, d1, kps2, d2, prev, 30, 0.9, 50, check_ori)
    assert n == pn and np.array_equal(m12, pm12) and np.array_equal(pout, pprev)
    assert n == int((m12 >= 0).sum()) and n > 60
    # every match is one-to-one
    used = m12[m12 >= 0]
    assert len(np.unique(used)) == len(used)
    # mode 1 (ORB-SLAM2's own window predicate) runs and matches too
    kps1, d1, kps2, d2 = mc.init_scene(150, 200, 8, 1)
    prev = np.stack([kps1["x"], kps1["y"]], 1).astype(np.float32)
    n1_, m1_, _ = orc.search_for_initialization(g, kps1, d1, kps2, d2, prev, 30, 0.9, 50, check_ori, 1)
    assert n1_ > 60 and n1_ == int((m1_ >= 0).sum())
