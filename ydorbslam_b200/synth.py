"""Seeded synthetic inputs for parity tests and bench.py (numpy only, reproducible on any box).

Shapes follow BASELINE.json `configs`; generators follow SURVEY.md 8(d).  All randomness is a counter-based
hash (splitmix64 finaliser of seed/stream/index), so any frame can be produced independently and identically
from Python on any machine.
"""
import numpy as np

_M64 = np.uint64(0xFFFFFFFFFFFFFFFF)


def _mix64(z):
    """splitmix64 finaliser, vectorised over uint64 arrays."""
    with np.errstate(over="ignore"):
        z = (z + np.uint64(0x9E3779B97F4A7C15)) & _M64
        z = ((z ^ (z >> np.uint64(30))) * np.uint64(0xBF58476D1CE4E5B9)) & _M64
        z = ((z ^ (z >> np.uint64(27))) * np.uint64(0x94D049BB133111EB)) & _M64
        return z ^ (z >> np.uint64(31))


def rand_u64(seed, stream, n):
    """n pseudo-random uint64 for (seed, stream)."""
    with np.errstate(over="ignore"):
        base = _mix64(np.uint64(seed) * np.uint64(0xD1B54A32D192ED03) + np.uint64(stream) * np.uint64(0x8CB92BA72F3D8DD7))
        idx = np.arange(n, dtype=np.uint64)
        return _mix64(idx * np.uint64(0x2545F4914F6CDD1D) + base)


def rand_bytes(seed, stream, n):
    nq = (n + 7) // 8
    return rand_u64(seed, stream, nq).view(np.uint8)[:n].copy()


def _binomial5(a):
    """separable [1 4 6 4 1] in exact integers, reflect-101 borders; returns uint16 sums (x256)."""
    a = a.astype(np.int32)
    p = np.pad(a, ((0, 0), (2, 2)), mode="reflect")
    hsum = p[:, 0:-4] + 4 * p[:, 1:-3] + 6 * p[:, 2:-2] + 4 * p[:, 3:-1] + p[:, 4:]
    p = np.pad(hsum, ((2, 2), (0, 0)), mode="reflect")
    return p[0:-4, :] + 4 * p[1:-3, :] + 6 * p[2:-2, :] + 4 * p[3:-1, :] + p[4:, :]


def synth_frame(w, h, seed, n_rect=64, contrast=3):
    """Textured synthetic gray frame: smoothed hash noise (contrast-stretched) + filled rectangles.

    Dense enough that every pyramid level yields more FAST candidates than its quota (truncation,
    tie and >16-point node paths all run)."""
    noise = rand_bytes(seed, 1, w * h).reshape(h, w)
    s = (_binomial5(noise) + 128) >> 8  # 0..255
    s = np.clip((s - 128) * contrast + 128, 0, 255).astype(np.uint8)
    r = rand_u64(seed, 2, n_rect * 5)
    for i in range(n_rect):
        x0 = int(r[5 * i] % np.uint64(w))
        y0 = int(r[5 * i + 1] % np.uint64(h))
        rw = 8 + int(r[5 * i + 2] % np.uint64(max(w // 6, 9)))
        rh = 8 + int(r[5 * i + 3] % np.uint64(max(h // 6, 9)))
        g = int(r[5 * i + 4] % np.uint64(256))
        s[y0 : min(y0 + rh, h), x0 : min(x0 + rw, w)] = g
    return s


def synth_sparse_frame(w, h, seed, n_rect=24):
    """Flat background with a few rectangles: few corners, many empty FAST cells (no-growth quadtree exit)."""
    s = np.full((h, w), 96, np.uint8)
    r = rand_u64(seed, 3, n_rect * 5)
    for i in range(n_rect):
        x0 = int(r[5 * i] % np.uint64(w))
        y0 = int(r[5 * i + 1] % np.uint64(h))
        rw = 6 + int(r[5 * i + 2] % np.uint64(max(w // 5, 7)))
        rh = 6 + int(r[5 * i + 3] % np.uint64(max(h // 5, 7)))
        g = int(r[5 * i + 4] % np.uint64(256))
        s[y0 : min(y0 + rh, h), x0 : min(x0 + rw, w)] = g
    return s


def synth_batch(w, h, batch, seed0=0, **kw):
    return np.stack([synth_frame(w, h, seed0 + i, **kw) for i in range(batch)])


def synth_stereo_pair(w, h, seed):
    """Right image = left shifted by a per-row-band disparity in 4..60 px, plus +-2 noise."""
    left = synth_frame(w, h, seed)
    bands = 8
    disp = 4 + (rand_u64(seed, 4, bands) % np.uint64(57)).astype(np.int64)
    right = np.empty_like(left)
    wide = synth_frame(w + 64, h, seed + 7919)  # fills the strip uncovered by the shift
    for b in range(bands):
        y0, y1 = b * h // bands, (b + 1) * h // bands
        d = int(disp[b])
        right[y0:y1, : w - d] = left[y0:y1, d:]
        right[y0:y1, w - d :] = wide[y0:y1, :d]
    n = (rand_bytes(seed, 5, w * h).reshape(h, w) % 5).astype(np.int16) - 2
    right = np.clip(right.astype(np.int16) + n, 0, 255).astype(np.uint8)
    return left, right


def synth_descriptors(n, seed):
    return rand_bytes(seed, 6, n * 32).reshape(n, 32)


def flip_bits(desc, p, seed):
    """Flip each bit with probability p (p quantised to 1/256)."""
    n = desc.shape[0]
    thr = int(round(p * 256))
    r = rand_bytes(seed, 7, n * 256).reshape(n, 256)
    flips = np.packbits((r < thr).astype(np.uint8), axis=1, bitorder="little")
    return desc ^ flips


def synth_query_descriptors(db, nq, seed, p=0.08):
    """Half: perturbed database rows (true matches); half: uniform random."""
    n_true = nq // 2
    rows = (rand_u64(seed, 8, n_true) % np.uint64(len(db))).astype(np.int64)
    a = flip_bits(db[rows], p, seed)
    b = synth_descriptors(nq - n_true, seed + 104729)
    return np.concatenate([a, b], axis=0), rows


def synth_map_points(kps, desc, n_mp, w, h, n_levels, seed):
    """Projected local-map points for searchByProjection.

    Returns dict of arrays: proj_x, proj_y, proj_right_x, view_cos, level, active, desc, obs_positive."""
    r = rand_u64(seed, 9, n_mp * 6)
    u = (r[0::6] >> np.uint64(11)).astype(np.float64) / float(1 << 53)
    v = (r[1::6] >> np.uint64(11)).astype(np.float64) / float(1 << 53)
    nk = max(len(kps), 1)
    src = (r[2::6] % np.uint64(nk)).astype(np.int64)
    near = (r[3::6] & np.uint64(1)).astype(bool)  # half project near an existing keypoint, half anywhere
    px = np.where(near & (len(kps) > 0), (kps["x"][src] if len(kps) else 0) + (u - 0.5) * 6.0, u * w).astype(np.float32)
    py = np.where(near & (len(kps) > 0), (kps["y"][src] if len(kps) else 0) + (v - 0.5) * 6.0, v * h).astype(np.float32)
    lvl_rand = (r[4::6] % np.uint64(n_levels)).astype(np.int32)
    level = np.where(near & (len(kps) > 0), kps["octave"][src] if len(kps) else 0, lvl_rand).astype(np.int32)
    vc = (0.9 + 0.1 * ((r[5::6] >> np.uint64(11)).astype(np.float64) / float(1 << 53))).astype(np.float32)
    if len(kps):
        d_near = flip_bits(desc[src], 0.08, seed + 1)
    else:
        d_near = synth_descriptors(n_mp, seed + 2)
    d_rand = synth_descriptors(n_mp, seed + 3)
    d = np.where(near[:, None], d_near, d_rand)
    r2 = rand_u64(seed, 10, n_mp * 2)
    active = ((r2[0::2] % np.uint64(16)) != 0).astype(np.int32)  # 1/16 not in view
    obs_pos = ((r2[1::2] % np.uint64(8)) != 0).astype(np.uint8)  # 1/8 have no observations yet
    return dict(
        proj_x=px, proj_y=py, proj_right_x=(px - 10.0).astype(np.float32), view_cos=vc, level=level, active=active,
        desc=np.ascontiguousarray(d), obs_positive=obs_pos, src=src,
    )


def synth_source_angles(kps, src, seed):
    """Angles of the source key points behind projected map points: mostly the matched key point's angle plus a small
    rotation (one dominant histogram bin), some with an arbitrary rotation (the bins the orientation check clears)."""
    n = len(src)
    r = rand_u64(seed, 11, n * 2)
    small = ((r[0::2] >> np.uint64(11)).astype(np.float64) / float(1 << 53) - 0.5) * 40.0
    big = (r[1::2] >> np.uint64(11)).astype(np.float64) / float(1 << 53) * 360.0
    base = kps["angle"][src].astype(np.float64) if len(kps) else np.zeros(n)
    rot = np.where((r[0::2] % np.uint64(5)) == 0, big, small)
    return np.mod(base + rot, 360.0).astype(np.float32)


def synth_feature_vector(desc, seed, bits=6, drop_every=5):
    """A DBoW3-like FeatureVector for `desc` without a vocabulary: node id = hash of the descriptor's first `bits` bits
    (so perturbed copies of a descriptor mostly land in the same node), ids sparse and ascending, features in
    ascending index inside a node (DBoW3 transform pushes them in feature order); every `drop_every`-th node of the
    id space (chosen by seed) is absent so the two sides of a match have unshared nodes (lower_bound path).

    Returns (node_ids, node_start, items) int32 arrays."""
    n = len(desc)
    word = desc[:, 0].astype(np.int64) & ((1 << bits) - 1)
    node = word * 7 + 3
    keep = (word + seed) % drop_every != 0
    ids = np.unique(node[keep]) if n else np.zeros(0, np.int64)
    start = [0]
    items = []
    for i in ids:
        members = np.nonzero(keep & (node == i))[0]
        items.extend(members.tolist())
        start.append(len(items))
    return ids.astype(np.int32), np.asarray(start, np.int32), np.asarray(items, np.int32)


def synth_vocabulary(k=10, L=6, seed=11, flip=0.18, stop_fraction=0.02, prune=0.0, shuffle_children=False):
    """A DBoW3-shaped vocabulary tree in the flat form of include/ydorb.h YdVocabularyDesc.

    Nodes are numbered breadth-first (root 0).  A child's descriptor is its parent's with every bit flipped with
    probability `flip` (what hierarchical k-medians leaves behind: siblings close to each other, far from other
    subtrees).  `prune` makes that fraction of interior nodes leaves (ragged depth); `shuffle_children` permutes each
    node's child order so that it differs from id order (a tree loaded from a stream, Vocabulary.cpp:1382-1393).
    Words are numbered in ascending node id (Vocabulary.cpp createWords); a leaf's weight is an idf-like positive double,
    0 for `stop_fraction` of them (stopped words, `if (w > 0)` at :781)."""
    parent = [0]
    level = [0]
    desc = [np.zeros(32, np.uint8)]
    desc[0][:] = rand_bytes(seed, 1, 32)
    frontier = [0]
    u = lambda stream, n: rand_u64(seed, stream, n).astype(np.float64) / 2.0**64
    for lv in range(1, L + 1):
        pr = u(100 + lv, len(frontier))
        keep = [p for j, p in enumerate(frontier) if lv == 1 or pr[j] >= prune]
        if not keep:
            break
        nk = len(keep) * k
        bits = (rand_u64(seed, 200 + lv, nk * 256).astype(np.float64) / 2.0**64 < flip).reshape(nk, 256)
        masks = np.packbits(bits, axis=1, bitorder="little")
        base = len(parent)
        pd = np.stack([desc[p] for p in keep]).repeat(k, axis=0)
        cd = pd ^ masks
        for j in range(nk):
            parent.append(keep[j // k])
            level.append(lv)
            desc.append(cd[j])
        frontier = list(range(base, base + nk))
    n = len(parent)
    parent = np.asarray(parent, np.int32)
    descriptors = np.stack(desc).astype(np.uint8)
    order = np.arange(1, n)
    if shuffle_children:
        order = order[np.argsort(rand_u64(seed, 3, n - 1), kind="stable")]
    order = order[np.argsort(parent[order], kind="stable")]  # grouped by parent, shuffled (or id) order inside
    counts = np.bincount(parent[1:], minlength=n)
    child_start = np.zeros(n + 1, np.int32)
    child_start[1:] = np.cumsum(counts)
    children = order.astype(np.int32)
    leaf = counts == 0
    word_ids = np.zeros(n, np.int32)
    word_ids[leaf] = np.arange(int(leaf.sum()), dtype=np.int32)
    weights = np.zeros(n, np.float64)
    w = np.log(1.0 + 1.0 / (u(4, n) + 1e-3))
    w[u(5, n) < stop_fraction] = 0.0
    weights[leaf] = w[leaf]
    return dict(k=k, L=L, child_start=child_start, children=children, descriptors=descriptors, weights=weights,
                word_ids=word_ids, parent=parent)


def synth_vocabulary_full(k=10, L=6, seed=11, flip=0.18, stop_fraction=0.02):
    """synth_vocabulary without pruning/shuffling, fully vectorised: the 1 111 111-node k=10, L=6 tree takes seconds."""
    starts = [0]
    for lv in range(1, L + 2):
        starts.append(starts[-1] + k ** (lv - 1))
    n = starts[L + 1]
    descriptors = np.zeros((n, 32), np.uint8)
    descriptors[0] = rand_bytes(seed, 1, 32)
    for lv in range(1, L + 1):
        a, b = starts[lv], starts[lv + 1]
        nk = b - a
        out = np.empty((nk, 32), np.uint8)
        step = 1 << 16
        for c in range(0, nk, step):
            m = min(step, nk - c)
            bits = (rand_u64(seed * 1000003 + lv, c, m * 256) >> np.uint64(40)).astype(np.float64) / 2.0**24 < flip
            out[c : c + m] = np.packbits(bits.reshape(m, 256), axis=1, bitorder="little")
        par = descriptors[starts[lv - 1] : starts[lv]].repeat(k, axis=0)
        descriptors[a:b] = par ^ out
    child_start = np.zeros(n + 1, np.int64)
    interior = starts[L]
    child_start[1 : interior + 1] = np.arange(1, interior + 1, dtype=np.int64) * k
    child_start[interior + 1 :] = n - 1
    children = np.arange(1, n, dtype=np.int32)
    word_ids = np.zeros(n, np.int32)
    word_ids[interior:] = np.arange(n - interior, dtype=np.int32)
    uu = lambda stream: rand_u64(seed, stream, n).astype(np.float64) / 2.0**64
    weights = np.zeros(n, np.float64)
    w = np.log(1.0 + 1.0 / (uu(4) + 1e-3))
    w[uu(5) < stop_fraction] = 0.0
    weights[interior:] = w[interior:]
    return dict(k=k, L=L, child_start=child_start.astype(np.int32), children=children, descriptors=descriptors,
                weights=weights, word_ids=word_ids)


def synth_vocabulary_queries(voc, n, seed=12, flip=0.06):
    """n descriptors near random words of `voc` (leaf descriptor with `flip` of its bits toggled), half of them plain
    uniform noise so that the descent also sees far-from-everything inputs and ties."""
    leaves = np.flatnonzero(np.diff(voc["child_start"]) == 0)
    pick = leaves[(rand_u64(seed, 1, n) % np.uint64(len(leaves))).astype(np.int64)]
    bits = (rand_u64(seed, 2, n * 256).astype(np.float64) / 2.0**64 < flip).reshape(n, 256)
    q = voc["descriptors"][pick] ^ np.packbits(bits, axis=1, bitorder="little")
    noise = rand_bytes(seed, 3, n * 32).reshape(n, 32)
    use_noise = (rand_u64(seed, 4, n) & np.uint64(1)).astype(bool)
    q[use_noise] = noise[use_noise]
    return np.ascontiguousarray(q, np.uint8)
