"""ctypes binding for the CPU oracle (TEST INFRASTRUCTURE -- see oracle/ydorb_oracle.h).

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may import this.
"""
import ctypes as C
import os
import pathlib
import subprocess

import numpy as np

_HERE = pathlib.Path(__file__).resolve().parent
_SO = _HERE / "_build" / "libydorb_oracle.so"

MAX_LEVELS = 16
PAD = 19
BORDER = 16

KP_DTYPE = np.dtype(
    [("x", "<f4"), ("y", "<f4"), ("size", "<f4"), ("angle", "<f4"), ("response", "<f4"), ("octave", "<i4"), ("class_id", "<i4")]
)
CAND_DTYPE = np.dtype([("x", "<i4"), ("y", "<i4"), ("response", "<f4")])
TOP2_DTYPE = np.dtype([("best_idx", "<i4"), ("best_dist", "<i4"), ("second_dist", "<i4"), ("second_idx", "<i4")])
MPPROJ_DTYPE = np.dtype(
    [("proj_x", "<f4"), ("proj_y", "<f4"), ("proj_right_x", "<f4"), ("view_cos", "<f4"), ("level", "<i4"), ("active", "<i4")]
)


class Params(C.Structure):
    _fields_ = [
        ("n_features", C.c_int32),
        ("scale_factor", C.c_float),
        ("n_levels", C.c_int32),
        ("ini_th", C.c_int32),
        ("min_th", C.c_int32),
        ("sf", C.c_float * MAX_LEVELS),
        ("inv_sf", C.c_float * MAX_LEVELS),
        ("sf2", C.c_float * MAX_LEVELS),
        ("inv_sf2", C.c_float * MAX_LEVELS),
        ("quota", C.c_int32 * MAX_LEVELS),
        ("umax", C.c_int32 * 16),
    ]


class FrameGeom(C.Structure):
    _fields_ = [
        ("min_x", C.c_float),
        ("min_y", C.c_float),
        ("max_x", C.c_float),
        ("max_y", C.c_float),
        ("n_levels", C.c_int32),
        ("sf", C.c_float * MAX_LEVELS),
    ]


def build(force=False):
    srcs = [_HERE / f for f in ("ydorb_oracle.cpp", "ydorb_oracle_map.cpp", "ydorb_oracle_bow.cpp", "ydorb_oracle_pre.cpp", "ydorb_oracle.h")]
    if force or not _SO.exists() or _SO.stat().st_mtime < max(f.stat().st_mtime for f in srcs):
        subprocess.check_call(["make", "-C", str(_HERE)], stdout=subprocess.DEVNULL)
    return _SO


_lib = None


def lib():
    global _lib
    if _lib is None:
        build()
        _lib = C.CDLL(str(_SO))
        _lib.orc_fast_atan2.restype = C.c_float
        _lib.orc_fast_atan2.argtypes = [C.c_float, C.c_float]
        _lib.orc_ic_angle.restype = C.c_float
        _lib.orc_pyramid_bytes.restype = C.c_int64
        _lib.orc_cvround.argtypes = [C.c_float]
    return _lib


def _p(a):
    return a.ctypes.data_as(C.c_void_p) if a is not None else None


def make_params(n_features=1000, scale=1.2, n_levels=8, ini_th=20, min_th=None):
    """min_th=None reproduces the reference ctor (min := ini, orbExtractor.cpp:318)."""
    p = Params()
    if min_th is None:
        min_th = ini_th
    rc = lib().orc_params_init(C.byref(p), n_features, C.c_float(scale), n_levels, ini_th, min_th)
    assert rc == 0
    return p


def level_size(p, w, h, level):
    lw, lh = C.c_int(), C.c_int()
    lib().orc_level_size(C.byref(p), w, h, level, C.byref(lw), C.byref(lh))
    return lw.value, lh.value


def pyramid(p, img):
    """Returns list of padded level arrays (views into one buffer)."""
    img = np.ascontiguousarray(img, dtype=np.uint8)
    h, w = img.shape
    off = (C.c_int64 * MAX_LEVELS)()
    tot = lib().orc_pyramid_bytes(C.byref(p), w, h, off)
    buf = np.zeros(tot, dtype=np.uint8)
    lib().orc_pyramid(C.byref(p), _p(img), w, h, w, _p(buf))
    out = []
    for l in range(p.n_levels):
        lw, lh = level_size(p, w, h, l)
        out.append(buf[off[l] : off[l] + (lw + 2 * PAD) * (lh + 2 * PAD)].reshape(lh + 2 * PAD, lw + 2 * PAD))
    return out, buf


def resize_linear(src, dw, dh):
    src = np.ascontiguousarray(src, dtype=np.uint8)
    dst = np.zeros((dh, dw), dtype=np.uint8)
    lib().orc_resize_linear(_p(src), src.shape[1], src.shape[0], src.shape[1], _p(dst), dw, dh, dw)
    return dst


def gauss7(src):
    src = np.ascontiguousarray(src, dtype=np.uint8)
    dst = np.zeros_like(src)
    lib().orc_gauss7(_p(src), src.shape[1], src.shape[0], src.shape[1], _p(dst), src.shape[1])
    return dst


def fast(img, th):
    img = np.ascontiguousarray(img, dtype=np.uint8)
    cap = img.size
    xs = np.zeros(cap, np.int32)
    ys = np.zeros(cap, np.int32)
    rs = np.zeros(cap, np.int32)
    n = lib().orc_fast(_p(img), img.shape[1], img.shape[0], img.shape[1], th, _p(xs), _p(ys), _p(rs), cap)
    return xs[:n], ys[:n], rs[:n]


def fast_atan2(y, x):
    return float(lib().orc_fast_atan2(C.c_float(y), C.c_float(x)))


def level_candidates(padded_level, ini_th, min_th):
    """padded_level: (lh+38, lw+38) array.  Returns CAND_DTYPE array in reference order."""
    lh, lw = padded_level.shape[0] - 2 * PAD, padded_level.shape[1] - 2 * PAD
    st = padded_level.strides[0]
    base = padded_level.ctypes.data + PAD * st + PAD
    n = lib().orc_level_candidates(C.c_void_p(base), lw, lh, st, ini_th, min_th, None, 0)
    out = np.zeros(max(n, 1), CAND_DTYPE)
    lib().orc_level_candidates(C.c_void_p(base), lw, lh, st, ini_th, min_th, _p(out), n)
    return out[:n]


def quadtree(cand, rootw, rooth, want, form="list"):
    cand = np.ascontiguousarray(cand, dtype=CAND_DTYPE)
    sel = np.zeros(max(want, 1), np.int32)
    fn = lib().orc_quadtree_list if form == "list" else lib().orc_quadtree_sortkey
    m = fn(_p(cand), len(cand), rootw, rooth, want, _p(sel))
    return sel[:m]


def ic_angle(p, padded_level, x, y):
    st = padded_level.strides[0]
    base = padded_level.ctypes.data + PAD * st + PAD
    return float(lib().orc_ic_angle(C.byref(p), C.c_void_p(base), st, int(x), int(y)))


def descriptor(blurred, x, y, angle):
    blurred = np.ascontiguousarray(blurred, dtype=np.uint8)
    d = np.zeros(32, np.uint8)
    lib().orc_descriptor(_p(blurred), blurred.shape[1], int(x), int(y), C.c_float(angle), _p(d))
    return d


def extract(p, img, cap=None):
    img = np.ascontiguousarray(img, dtype=np.uint8)
    h, w = img.shape
    cap = cap or (p.n_features + 64)
    kps = np.zeros(cap, KP_DTYPE)
    desc = np.zeros((cap, 32), np.uint8)
    per = np.zeros(MAX_LEVELS, np.int32)
    n = lib().orc_extract(C.byref(p), _p(img), w, h, w, _p(kps), _p(desc), cap, _p(per))
    assert n <= cap
    return kps[:n], desc[:n], per[: p.n_levels]


def extract_batch(p, imgs, threads, cap=None):
    imgs = np.ascontiguousarray(imgs, dtype=np.uint8)
    b, h, w = imgs.shape
    cap = cap or (p.n_features + 64)
    kps = np.zeros((b, cap), KP_DTYPE)
    desc = np.zeros((b, cap, 32), np.uint8)
    n = np.zeros(b, np.int32)
    lib().orc_extract_batch(C.byref(p), _p(imgs), w, h, w, C.c_int64(h * w), b, _p(kps), _p(desc), cap, _p(n), threads)
    return kps, desc, n


def hamming(a, b):
    a = np.ascontiguousarray(a, dtype=np.uint8)
    b = np.ascontiguousarray(b, dtype=np.uint8)
    return int(lib().orc_hamming(_p(a), _p(b)))


def top2(q, db, idx_base=0, threads=None):
    q = np.ascontiguousarray(q, dtype=np.uint8)
    db = np.ascontiguousarray(db, dtype=np.uint8)
    out = np.zeros(len(q), TOP2_DTYPE)
    lib().orc_top2(_p(q), len(q), _p(db), len(db), idx_base, _p(out), threads or os.cpu_count())
    return out


def top2_merge(parts):
    """parts: (nshards, nq) TOP2 array."""
    parts = np.ascontiguousarray(parts, dtype=TOP2_DTYPE)
    ns, nq = parts.shape
    out = np.zeros(nq, TOP2_DTYPE)
    lib().orc_top2_merge(_p(parts), ns, nq, _p(out))
    return out


def ratio_accept(t, ratio, th):
    t = np.ascontiguousarray(t, dtype=TOP2_DTYPE)
    acc = np.zeros(len(t), np.uint8)
    lib().orc_ratio_accept(_p(t), len(t), C.c_float(ratio), th, _p(acc))
    return acc


def make_geom(p, w, h):
    g = FrameGeom()
    g.min_x, g.min_y, g.max_x, g.max_y = 0.0, 0.0, float(w), float(h)
    g.n_levels = p.n_levels
    for i in range(p.n_levels):
        g.sf[i] = p.sf[i]
    return g


def keypoints_in_area(g, kps, x, y, r, min_level=-1, max_level=-1):
    kps = np.ascontiguousarray(kps, dtype=KP_DTYPE)
    out = np.zeros(max(len(kps), 1), np.int32)
    n = lib().orc_keypoints_in_area(
        C.byref(g), _p(kps), len(kps), C.c_float(x), C.c_float(y), C.c_float(r), min_level, max_level, _p(out), len(out)
    )
    return out[:n]


def search_by_projection(g, kps, desc, right_x, kp_occupied, mps, mp_desc, mp_obs_positive, th, ratio):
    kps = np.ascontiguousarray(kps, dtype=KP_DTYPE)
    desc = np.ascontiguousarray(desc, dtype=np.uint8)
    right_x = np.ascontiguousarray(right_x, dtype=np.float32)
    kp_occupied = np.ascontiguousarray(kp_occupied, dtype=np.uint8)
    mps = np.ascontiguousarray(mps, dtype=MPPROJ_DTYPE)
    mp_desc = np.ascontiguousarray(mp_desc, dtype=np.uint8)
    mp_obs_positive = np.ascontiguousarray(mp_obs_positive, dtype=np.uint8)
    out = np.zeros(len(kps), np.int32)
    n = lib().orc_search_by_projection(
        C.byref(g), _p(kps), _p(desc), _p(right_x), len(kps), _p(kp_occupied), _p(mps), _p(mp_desc), _p(mp_obs_positive),
        len(mps), C.c_float(th), C.c_float(ratio), _p(out),
    )
    return n, out


def stereo_match(p, w, h, pyr_l, pyr_r, kpl, dl, kpr, dr, bf, max_d, mode=0):
    kpl = np.ascontiguousarray(kpl, dtype=KP_DTYPE)
    kpr = np.ascontiguousarray(kpr, dtype=KP_DTYPE)
    dl = np.ascontiguousarray(dl, dtype=np.uint8)
    dr = np.ascontiguousarray(dr, dtype=np.uint8)
    rx = np.zeros(len(kpl), np.float32)
    dp = np.zeros(len(kpl), np.float32)
    lib().orc_stereo_match(
        C.byref(p), w, h, _p(pyr_l), _p(pyr_r), _p(kpl), _p(dl), len(kpl), _p(kpr), _p(dr), len(kpr),
        C.c_float(bf), C.c_float(max_d), mode, _p(rx), _p(dp),
    )
    return rx, dp


# ---- M2/M3/M4/M7 window searches, M5/M6 BoW matchers -------------------------------------------------------
WINDOW_QUERY_DTYPE = np.dtype(
    [("x", "<f4"), ("y", "<f4"), ("r", "<f4"), ("min_level", "<i4"), ("max_level", "<i4"), ("oct_lo", "<i4"), ("oct_hi", "<i4"),
     ("right_x", "<f4"), ("right_r", "<f4"), ("angle", "<f4"), ("flags", "<i4")]
)


class WindowParams(C.Structure):
    _fields_ = [("max_dist", C.c_int32), ("strict", C.c_int32), ("exclusive", C.c_int32), ("check_orientation", C.c_int32),
                ("chi2_gate", C.c_int32), ("inv_sf2", C.c_float * 16)]


class FeatureVector(C.Structure):
    _fields_ = [("node_ids", C.c_void_p), ("node_start", C.c_void_p), ("items", C.c_void_p), ("n_nodes", C.c_int32)]

    @classmethod
    def from_arrays(cls, node_ids, node_start, items):
        fv = cls()
        fv._keep = (np.ascontiguousarray(node_ids, np.int32), np.ascontiguousarray(node_start, np.int32),
                    np.ascontiguousarray(items, np.int32))
        fv.node_ids, fv.node_start, fv.items = (a.ctypes.data for a in fv._keep)
        fv.n_nodes = len(fv._keep[0])
        return fv


def _window_params(prm):
    """Accepts the product's ctypes WindowParams (same layout) or an oracle one."""
    p = WindowParams()
    for name, _ in WindowParams._fields_[:5]:
        setattr(p, name, getattr(prm, name))
    for i in range(16):
        p.inv_sf2[i] = prm.inv_sf2[i]
    return p


def search_windows(g, kps, desc, right_x, kp_occupied, queries, q_desc, prm):
    kps = np.ascontiguousarray(kps, dtype=KP_DTYPE)
    desc = np.ascontiguousarray(desc, dtype=np.uint8)
    right_x = np.ascontiguousarray(right_x, dtype=np.float32)
    occ = None if kp_occupied is None else np.ascontiguousarray(kp_occupied, dtype=np.uint8)
    q = np.ascontiguousarray(queries, dtype=WINDOW_QUERY_DTYPE)
    qd = np.ascontiguousarray(q_desc, dtype=np.uint8)
    best = np.full(max(len(q), 1), -1, np.int32)
    match = np.full(max(len(kps), 1), -1, np.int32)
    p = _window_params(prm)
    n = lib().orc_search_windows(C.byref(g), _p(kps), _p(desc), _p(right_x), len(kps), _p(occ) if occ is not None else None,
                                 _p(q), _p(qd), len(q), C.byref(p), _p(best), _p(match))
    return n, best[: len(q)], match[: len(kps)]


def search_for_initialization(g2, kps1, desc1, kps2, desc2, prev_matched, window_size, nn_ratio, th_low, check_orientation,
                              area_mode=0):
    """EXTRA (not in the reference): ORB-SLAM2's SearchForInitialization, sequential restatement."""
    k1 = np.ascontiguousarray(kps1, dtype=KP_DTYPE)
    k2 = np.ascontiguousarray(kps2, dtype=KP_DTYPE)
    d1 = np.ascontiguousarray(desc1, dtype=np.uint8)
    d2 = np.ascontiguousarray(desc2, dtype=np.uint8)
    prev = np.array(prev_matched, dtype=np.float32).reshape(len(k1), 2).copy()
    m12 = np.full(max(len(k1), 1), -1, np.int32)
    lib().orc_search_for_initialization.restype = C.c_int
    n = lib().orc_search_for_initialization(C.byref(g2), _p(k1), _p(d1), len(k1), _p(k2), _p(d2), len(k2), _p(prev),
                                            int(window_size), C.c_float(nn_ratio), int(th_low), int(bool(check_orientation)),
                                            int(area_mode), _p(m12))
    return n, m12[: len(k1)], prev


def three_maxima(sizes):
    sz = np.ascontiguousarray(sizes, dtype=np.int32)
    a, b, c = C.c_int32(), C.c_int32(), C.c_int32()
    lib().orc_three_maxima(_p(sz), len(sz), C.byref(a), C.byref(b), C.byref(c))
    return a.value, b.value, c.value


def search_by_bow(mode, kps1, desc1, valid1, fv1, kps2, desc2, valid2, fv2, ratio, check_orientation):
    k1 = np.ascontiguousarray(kps1, dtype=KP_DTYPE)
    k2 = np.ascontiguousarray(kps2, dtype=KP_DTYPE)
    d1 = np.ascontiguousarray(desc1, dtype=np.uint8)
    d2 = np.ascontiguousarray(desc2, dtype=np.uint8)
    v1 = np.ascontiguousarray(valid1, dtype=np.uint8)
    v2 = np.ascontiguousarray(valid2 if valid2 is not None else np.ones(len(k2)), dtype=np.uint8)
    f1, f2 = FeatureVector.from_arrays(*fv1), FeatureVector.from_arrays(*fv2)
    n_out = len(k2) if mode == 0 else len(k1)
    out = np.full(max(n_out, 1), -1, np.int32)
    n = lib().orc_search_by_bow(_p(k1), _p(d1), _p(v1), len(k1), C.byref(f1), _p(k2), _p(d2), _p(v2), len(k2), C.byref(f2),
                                mode, C.c_float(ratio), int(check_orientation), _p(out))
    return n, out[:n_out]


def search_for_triangulation(kps1, desc1, right_x1, has_mp1, fv1, kps2, desc2, right_x2, has_mp2, fv2, F, epipole, sf, sf2,
                             stereo_only, check_orientation):
    k1 = np.ascontiguousarray(kps1, dtype=KP_DTYPE)
    k2 = np.ascontiguousarray(kps2, dtype=KP_DTYPE)
    d1 = np.ascontiguousarray(desc1, dtype=np.uint8)
    d2 = np.ascontiguousarray(desc2, dtype=np.uint8)
    r1 = np.ascontiguousarray(right_x1, dtype=np.float32)
    r2 = np.ascontiguousarray(right_x2, dtype=np.float32)
    h1 = np.ascontiguousarray(has_mp1, dtype=np.uint8)
    h2 = np.ascontiguousarray(has_mp2, dtype=np.uint8)
    f1, f2 = FeatureVector.from_arrays(*fv1), FeatureVector.from_arrays(*fv2)
    Fm = np.ascontiguousarray(F, dtype=np.float32).reshape(9)
    sfa = np.zeros(16, np.float32)
    sf2a = np.zeros(16, np.float32)
    sfa[: len(sf)] = sf
    sf2a[: len(sf2)] = sf2
    out = np.full(max(len(k1), 1), -1, np.int32)
    n = lib().orc_search_for_triangulation(_p(k1), _p(d1), _p(r1), _p(h1), len(k1), C.byref(f1), _p(k2), _p(d2), _p(r2), _p(h2),
                                           len(k2), C.byref(f2), _p(Fm), C.c_float(epipole[0]), C.c_float(epipole[1]), _p(sfa),
                                           _p(sf2a), int(stereo_only), int(check_orientation), _p(out))
    out = out[: len(k1)]
    first = np.nonzero(out >= 0)[0]
    return n, np.stack([first, out[first]], axis=1).astype(np.int32)


# ---- SURVEY 8(f) "next" rows (ydorb_oracle_map.cpp)
MAPPOINT3D_DTYPE = np.dtype([("pos", "<f4", 3), ("normal", "<f4", 3), ("max_distance", "<f4"), ("min_distance", "<f4")])


class Camera(C.Structure):
    _fields_ = [("R", C.c_float * 9), ("t", C.c_float * 3), ("origin", C.c_float * 3), ("fx", C.c_float), ("fy", C.c_float),
                ("cx", C.c_float), ("cy", C.c_float), ("min_x", C.c_float), ("max_x", C.c_float), ("min_y", C.c_float),
                ("max_y", C.c_float), ("bf", C.c_float), ("log_scale_factor", C.c_float), ("n_levels", C.c_int32)]


def distinctive_descriptors(desc, start):
    d = np.ascontiguousarray(desc, np.uint8).reshape(-1, 32)
    s = np.ascontiguousarray(start, np.int32)
    best = np.zeros(len(s) - 1, np.int32)
    lib().orc_distinctive_descriptors(_p(d), _p(s), len(best), _p(best))
    return best


def frustum(cam, map_points, view_cos_limit=0.5):
    """cam: any ctypes struct with YdCamera's layout."""
    mp = np.ascontiguousarray(map_points, MAPPOINT3D_DTYPE)
    out = np.zeros(len(mp), MPPROJ_DTYPE)
    lib().orc_frustum(C.byref(cam), _p(mp), len(mp), C.c_float(view_cos_limit), _p(out))
    return out


class Vocabulary(C.Structure):
    """OrcVocabulary: the DBoW3 tree in flat form (same layout as include/ydorb.h YdVocabularyDesc)."""

    _fields_ = [("n_nodes", C.c_int32), ("k", C.c_int32), ("L", C.c_int32), ("weighting", C.c_int32), ("scoring", C.c_int32),
                ("child_start", C.c_void_p), ("children", C.c_void_p), ("descriptors", C.c_void_p), ("weights", C.c_void_p),
                ("word_ids", C.c_void_p)]

    @classmethod
    def from_arrays(cls, k, L, weighting, scoring, child_start, children, descriptors, weights, word_ids):
        v = cls()
        v._keep = (np.ascontiguousarray(child_start, np.int32), np.ascontiguousarray(children, np.int32),
                   np.ascontiguousarray(descriptors, np.uint8), np.ascontiguousarray(weights, np.float64),
                   np.ascontiguousarray(word_ids, np.int32))
        v.n_nodes = len(v._keep[3])
        v.k, v.L, v.weighting, v.scoring = k, L, weighting, scoring
        v.child_start, v.children, v.descriptors, v.weights, v.word_ids = (a.ctypes.data for a in v._keep)
        return v


def voc_transform(voc, desc, levels_up=4):
    """DBoW3::Vocabulary::transform(features, v, fv, levelsup).  Returns a dict with the per-feature (word_id, node_id,
    weight), the BowVector (bow_ids, bow_values) and the FeatureVector in CSR form (fv_node_ids, fv_node_start, fv_items)."""
    d = np.ascontiguousarray(desc, np.uint8).reshape(-1, 32)
    n = len(d)
    word, node, wgt = np.zeros(n, np.int32), np.zeros(n, np.int32), np.zeros(n, np.float64)
    bi, bv = np.zeros(n, np.int32), np.zeros(n, np.float64)
    fi, fs, ft = np.zeros(n, np.int32), np.zeros(n + 1, np.int32), np.zeros(n, np.int32)
    nb, nf = C.c_int32(0), C.c_int32(0)
    lib().orc_voc_transform(C.byref(voc), _p(d), n, levels_up, _p(word), _p(node), _p(wgt), _p(bi), _p(bv), C.byref(nb),
                            _p(fi), _p(fs), _p(ft), C.byref(nf))
    return dict(word_id=word, node_id=node, weight=wgt, bow_ids=bi[: nb.value], bow_values=bv[: nb.value],
                fv_node_ids=fi[: nf.value], fv_node_start=fs[: nf.value + 1], fv_items=ft[: int(fs[nf.value])])


def cvt_gray(img, code):
    """cv::cvtColor(img, code) for code 0 RGB2GRAY, 1 BGR2GRAY, 2 RGBA2GRAY, 3 BGRA2GRAY on (h, w, 3|4) uint8."""
    a = np.ascontiguousarray(img, np.uint8)
    h, w, ch = a.shape
    assert ch == (4 if code >= 2 else 3)
    out = np.zeros((h, w), np.uint8)
    lib().orc_cvt_gray(_p(a), w, h, w * ch, code, _p(out), w)
    return out


def undistort_points(xy, K, dist, P):
    pts = np.ascontiguousarray(xy, np.float32).reshape(-1, 2)
    Kf, Pf, df = (np.ascontiguousarray(v, np.float32).ravel() for v in (K, P, dist))
    out = np.zeros_like(pts)
    lib().orc_undistort_points(_p(pts), len(pts), _p(Kf), _p(df), len(df), _p(Pf), _p(out))
    return out


def stereo_from_rgbd(kps, depth_map, factor, bf):
    k = np.ascontiguousarray(kps, KP_DTYPE)
    d = np.ascontiguousarray(depth_map)
    assert d.dtype in (np.uint16, np.float32)
    rx, dz = np.zeros(len(k), np.float32), np.zeros(len(k), np.float32)
    lib().orc_stereo_from_rgbd(_p(k), len(k), _p(d), 0 if d.dtype == np.uint16 else 1, d.shape[1], d.shape[0], d.strides[0],
                               C.c_float(factor), C.c_float(bf), _p(rx), _p(dz))
    return rx, dz
