"""Independent restatement of OrbExtractor::extractAndCompute that drives the *real* OpenCV primitives
through cv2 (resize / copyMakeBorder / FAST / GaussianBlur / fastAtan2) and transcribes the rest in Python.

Purpose: pin the C++ oracle.  cv2 cannot be assumed on every box, so this module is used by
oracle/gen_golden.py (committed fixtures under tests/golden/) and by tests that skip when cv2 is missing.
TEST INFRASTRUCTURE -- never imported by the product.

Reference lines cited are relative to the YDORBSLAM checkout (src/orbExtractor.cpp).
"""
import math

import numpy as np

PAD, BORDER, HALF = 19, 16, 15


def _pattern():
    import pathlib

    txt = (pathlib.Path(__file__).resolve().parents[1] / "ydorbslam_b200" / "csrc" / "brief_pattern.inc").read_text()
    vals = [int(t) for line in txt.splitlines() if not line.startswith("//") for t in line.split(",") if t.strip()]
    return np.array(vals, dtype=np.int32).reshape(512, 2)


PATTERN = _pattern()


def ctor(n_features, scale, n_levels):
    """:315-354 using numpy float32 where the reference stores floats."""
    sf32 = np.float32(scale)
    sfd = float(sf32)
    per = int(round_half_away(n_features * (1 - 1.0 / sfd) / (1.0 - math.pow(1.0 / sfd, n_levels))))
    sf = [np.float32(math.pow(sfd, i)) for i in range(n_levels)]
    inv = [np.float32(math.pow(sfd, -i)) for i in range(n_levels)]
    quota, tot = [], 0
    for i in range(n_levels):
        if i < n_levels - 1:
            quota.append(per)
            tot += per
            per = int(round_half_away(float(np.float32(per) / sf32)))
        else:
            quota.append(max(n_features - tot, 0))
    umax = [0] * 16
    vmax = int(math.floor(HALF * math.sqrt(2.0) / 2.0 + 1.0))
    vmin = int(math.ceil(HALF * math.sqrt(2.0) / 2.0))
    for v in range(vmax + 1):
        umax.append(int(round_half_away(math.sqrt(HALF * HALF + v * v))))
    i = 0
    for v in range(HALF, vmin - 1, -1):
        while umax[i] == umax[i + 1]:
            i += 1
        umax[v] = i
        i += 1
    return sf, inv, quota, umax[:16]


def round_half_away(x):
    return math.floor(x + 0.5) if x >= 0 else -math.floor(-x + 0.5)


def cv_round_f32(v):
    return int(np.rint(np.float32(v)))  # half-to-even


def pyramid(img, inv, n_levels):
    """:605-621 with real cv2.resize / cv2.copyMakeBorder.  Returns padded levels."""
    import cv2

    h, w = img.shape
    out, prev = [], None
    for l in range(n_levels):
        lw = cv_round_f32(np.float32(w) * inv[l])
        lh = cv_round_f32(np.float32(h) * inv[l])
        cur = img if l == 0 else cv2.resize(prev, (lw, lh), interpolation=cv2.INTER_LINEAR)
        padded = cv2.copyMakeBorder(cur, PAD, PAD, PAD, PAD, cv2.BORDER_REFLECT_101)
        out.append(padded)
        prev = cur
    return out


def level_candidates(level, ini_th, min_th):
    """:545-593 with real cv2 FAST per cell.  level = unpadded level image.  Returns list of (x, y, response)."""
    import cv2

    lh, lw = level.shape
    minbx = minby = BORDER
    maxbx, maxby = lw - BORDER, lh - BORDER
    width, height = np.float32(maxbx - minbx), np.float32(maxby - minby)
    cols, rows = int(width / np.float32(30.0)), int(height / np.float32(30.0))
    if cols <= 0 or rows <= 0:
        return []
    cw, ch = int(math.ceil(width / np.float32(cols))), int(math.ceil(height / np.float32(rows)))
    det_ini = cv2.FastFeatureDetector_create(ini_th, True, cv2.FAST_FEATURE_DETECTOR_TYPE_9_16)
    det_min = cv2.FastFeatureDetector_create(min_th, True, cv2.FAST_FEATURE_DETECTOR_TYPE_9_16)
    out = []
    for i in range(rows):
        iy = minby + i * ch
        my = iy + ch + 6
        if iy >= maxby - 3:
            continue
        my = min(my, maxby)
        for j in range(cols):
            ix = minbx + j * cw
            mx = ix + cw + 6
            if ix >= maxbx - 6:
                continue
            mx = min(mx, maxbx)
            roi = np.ascontiguousarray(level[iy:my, ix:mx])
            kps = det_ini.detect(roi)
            if len(kps) == 0:
                kps = det_min.detect(roi)
            for k in kps:
                out.append((int(k.pt[0]) + j * cw, int(k.pt[1]) + i * ch, float(k.response)))
    return out


# --- libstdc++ std::sort emulation (bits/stl_algo.h: __introsort_loop, __final_insertion_sort) --------------
def _std_sort(a, comp):
    """In-place emulation of libstdc++ std::sort(first,last,comp) on a Python list."""
    n = len(a)
    if n == 0:
        return

    def move_median_to_first(res, x, y, z):
        if comp(a[x], a[y]):
            if comp(a[y], a[z]):
                a[res], a[y] = a[y], a[res]
            elif comp(a[x], a[z]):
                a[res], a[z] = a[z], a[res]
            else:
                a[res], a[x] = a[x], a[res]
        elif comp(a[x], a[z]):
            a[res], a[x] = a[x], a[res]
        elif comp(a[y], a[z]):
            a[res], a[z] = a[z], a[res]
        else:
            a[res], a[y] = a[y], a[res]

    def unguarded_partition(first, last, pivot):
        while True:
            while comp(a[first], a[pivot]):
                first += 1
            last -= 1
            while comp(a[pivot], a[last]):
                last -= 1
            if not (first < last):
                return first
            a[first], a[last] = a[last], a[first]
            first += 1

    def heap_sort(first, last):  # __partial_sort(first,last,last): make_heap + sort_heap
        def adjust(hole, length, value):
            top = hole
            child = hole
            while child < (length - 1) // 2:
                child = 2 * (child + 1)
                if comp(a[first + child], a[first + child - 1]):
                    child -= 1
                a[first + hole] = a[first + child]
                hole = child
            if (length & 1) == 0 and child == (length - 2) // 2:
                child = 2 * (child + 1)
                a[first + hole] = a[first + child - 1]
                hole = child - 1
            parent = (hole - 1) // 2
            while hole > top and comp(a[first + parent], value):
                a[first + hole] = a[first + parent]
                hole = parent
                parent = (hole - 1) // 2
            a[first + hole] = value

        length = last - first
        if length >= 2:
            parent = (length - 2) // 2
            while True:
                adjust(parent, length, a[first + parent])
                if parent == 0:
                    break
                parent -= 1
        l = last
        while l - first > 1:
            l -= 1
            v = a[l]
            a[l] = a[first]
            adjust(0, l - first, v)

    def introsort(first, last, depth):
        while last - first > 16:
            if depth == 0:
                heap_sort(first, last)
                return
            depth -= 1
            mid = first + (last - first) // 2
            move_median_to_first(first, first + 1, mid, last - 1)
            cut = unguarded_partition(first + 1, last, first)
            introsort(cut, last, depth)
            last = cut

    def unguarded_linear_insert(last):
        val = a[last]
        nxt = last - 1
        while comp(val, a[nxt]):
            a[last] = a[nxt]
            last = nxt
            nxt -= 1
        a[last] = val

    def insertion_sort(first, last):
        if first == last:
            return
        for i in range(first + 1, last):
            if comp(a[i], a[first]):
                val = a[i]
                a[first + 1 : i + 1] = a[first:i]
                a[first] = val
            else:
                unguarded_linear_insert(i)

    introsort(0, n, 2 * (n.bit_length() - 1))
    if n > 16:
        insertion_sort(0, 16)
        for i in range(16, n):
            unguarded_linear_insert(i)
    else:
        insertion_sort(0, n)


def quadtree(cands, rootw, rooth, want):
    """:4-54, :455-544 list transcription.  cands = [(x,y,resp)].  Returns selected indices in output order."""

    class Node:
        __slots__ = ("kps", "tl", "tr", "bl", "br", "indiv")

        def __init__(self):
            self.kps, self.indiv = [], False
            self.tl = self.tr = self.bl = self.br = (0, 0)

    def divide(n):
        cx = int(math.ceil(float(np.float32(n.tr[0] + n.tl[0])) / 2.0))
        cy = int(math.ceil(float(np.float32(n.bl[1] + n.tl[1])) / 2.0))
        c = [Node() for _ in range(4)]
        c[0].tl, c[0].tr, c[0].bl, c[0].br = n.tl, (cx, n.tl[1]), (n.tl[0], cy), (cx, cy)
        c[1].tl, c[1].tr, c[1].bl, c[1].br = c[0].tr, n.tr, (cx, cy), (n.tr[0], cy)
        c[2].tl, c[2].tr, c[2].bl, c[2].br = c[0].bl, (cx, cy), n.bl, (cx, n.tl[1])
        c[3].tl, c[3].tr, c[3].bl, c[3].br = (cx, cy), c[1].br, c[2].br, n.br
        for k in n.kps:
            x, y = cands[k][0], cands[k][1]
            if x < cx and y < cy:
                c[0].kps.append(k)
            elif x < cx and y >= cy:
                c[2].kps.append(k)
            elif x >= cx and y < cy:
                c[1].kps.append(k)
            else:
                c[3].kps.append(k)
        for ch in c:
            ch.indiv = len(ch.kps) == 1
        return c

    root = Node()
    root.tl, root.tr, root.bl, root.br = (0, 0), (rootw, 0), (0, rooth), (rootw, rooth)
    root.kps = list(range(len(cands)))
    nodes = [root] if root.kps else []
    if nodes and len(root.kps) == 1:
        root.indiv = True
    last = 0
    while len(nodes) > last and len(nodes) < want:
        last = len(nodes)
        new_front, keep = [], []
        for n in nodes:  # front to back; children are pushed to the front, never revisited this pass
            if n.indiv:
                keep.append(n)
                continue
            for ch in divide(n):
                if ch.kps:
                    new_front.insert(0, ch)
        nodes = new_front + keep
    out = []
    for n in nodes:
        v = [(cands[k][2], k) for k in n.kps]
        _std_sort(v, lambda a, b: a[0] > b[0])
        out.append(v[0][1])
    return out[: max(want, 0)] if len(out) > want else out


def ic_angle(padded, umax, x, y):
    """:400-421 with cv2.fastAtan2.  padded = padded level; (x,y) level coords."""
    import cv2

    cy, cx = y + PAD, x + PAD
    row = padded[cy].astype(np.int64)
    m10 = int(sum(u * int(row[cx + u]) for u in range(-HALF, HALF + 1)))
    m01 = 0
    for v in range(1, HALF + 1):
        d = umax[v]
        pos = padded[cy + v, cx - d : cx + d + 1].astype(np.int64)
        neg = padded[cy - v, cx - d : cx + d + 1].astype(np.int64)
        us = np.arange(-d, d + 1)
        m01 += v * int((pos - neg).sum())
        m10 += int((us * (pos + neg)).sum())
    return float(cv2.fastAtan2(float(m01), float(m10)))


def descriptor(blurred, x, y, angle_deg):
    """:422-454"""
    a = np.float32(angle_deg) * np.float32(math.pi / 180.0)
    c, s = np.float32(math.cos(float(a))), np.float32(math.sin(float(a)))
    px, py = PATTERN[:, 0].astype(np.float32), PATTERN[:, 1].astype(np.float32)
    rr = np.rint((px * s).astype(np.float32) + (py * c).astype(np.float32)).astype(np.int64)
    cc = np.rint((px * c).astype(np.float32) - (py * s).astype(np.float32)).astype(np.int64)
    vals = blurred[y + rr, x + cc].astype(np.int32)
    bits = (vals[0::2] < vals[1::2]).astype(np.uint8)
    return np.packbits(bits.reshape(32, 8), axis=1, bitorder="little").reshape(32)


def extract(img, n_features=1000, scale=1.2, n_levels=8, ini_th=20, min_th=None, stages=None):
    """Whole extractAndCompute :355-399.  Returns (kps structured array fields as dict of lists, desc)."""
    import cv2

    if min_th is None:
        min_th = ini_th  # :318
    sf, inv, quota, umax = ctor(n_features, scale, n_levels)
    pyr = pyramid(img, inv, n_levels)
    kps, descs = [], []
    for l in range(n_levels):
        padded = pyr[l]
        level = padded[PAD:-PAD, PAD:-PAD]
        lh, lw = level.shape
        cands = level_candidates(level, ini_th, min_th)
        sel = quadtree(cands, (lw - BORDER) - BORDER, (lh - BORDER) - BORDER, quota[l])
        if stages is not None:
            stages.append(dict(level=l, padded=padded, cands=cands, sel=sel))
        if not sel:
            continue
        blurred = cv2.GaussianBlur(level.copy(), (7, 7), 2, sigmaY=2, borderType=cv2.BORDER_REFLECT_101)
        if stages is not None:
            stages[-1]["blurred"] = blurred
        size = float(int(np.float32(31) * sf[l]))
        for k in sel:
            x, y = cands[k][0] + BORDER, cands[k][1] + BORDER
            ang = ic_angle(padded, umax, x, y)
            descs.append(descriptor(blurred, x, y, ang))
            fx, fy = np.float32(x), np.float32(y)
            if l != 0:
                fx, fy = fx * sf[l], fy * sf[l]
            kps.append((float(fx), float(fy), size, ang, cands[k][2], l, -1))
    return kps, (np.stack(descs) if descs else np.zeros((0, 32), np.uint8))
