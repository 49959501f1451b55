"""Shared inputs of the vocabulary-transform tests."""
import numpy as np

from ydorbslam_b200 import synth

FIELDS = ("word_id", "node_id", "weight", "bow_ids", "bow_values", "fv_node_ids", "fv_node_start", "fv_items")

# (k, L, prune, shuffle_children): ragged trees and trees whose child order differs from id order included
TREES = [(10, 3, 0.0, False), (3, 4, 0.25, True), (9, 2, 0.0, True), (2, 7, 0.1, False), (4, 1, 0.0, False)]


def tree(k, L, prune, shuffle, seed=21):
    return synth.synth_vocabulary(k, L, seed=seed, prune=prune, shuffle_children=shuffle, stop_fraction=0.05)


def same(a, b):
    """Bit-exact comparison of two transform results (doubles compared as their 64-bit patterns)."""
    for f in FIELDS:
        x, y = np.asarray(a[f]), np.asarray(b[f])
        if x.dtype == np.float64:
            x, y = x.view(np.uint64), np.asarray(y, np.float64).view(np.uint64)
        assert x.shape == y.shape, (f, x.shape, y.shape)
        assert np.array_equal(x, y), f


def python_transform(voc, desc, levels_up, weighting, scoring):
    """Independent pure-Python restatement of Vocabulary::transform (Vocabulary.cpp:752-873) with dicts standing in
    for the std::maps (sorted on output); small inputs only."""
    cs, ch, D = voc["child_start"], voc["children"], voc["descriptors"]
    bits = np.unpackbits(D, axis=1)
    nid_level = voc["L"] - levels_up
    bow, fv = {}, {}
    word, node, wgt = [], [], []
    for i, d in enumerate(np.unpackbits(np.asarray(desc, np.uint8).reshape(-1, 32), axis=1)):
        cur, lvl, nid = 0, 0, 0
        while True:
            lvl += 1
            kids = ch[cs[cur] : cs[cur + 1]]
            dist = (bits[kids] != d).sum(1)
            cur = int(kids[int(np.argmin(dist))])  # argmin: first minimum = strict `<`
            if lvl == nid_level:
                nid = cur
            if cs[cur + 1] == cs[cur]:
                break
        w = float(voc["weights"][cur])
        wid = int(voc["word_ids"][cur])
        word.append(wid); node.append(nid); wgt.append(w)
        if w > 0:
            if weighting in (0, 1):
                bow[wid] = bow[wid] + w if wid in bow else w
            else:
                bow.setdefault(wid, w)
            fv.setdefault(nid, []).append(i)
    ids = sorted(bow)
    vals = [bow[i] for i in ids]
    must = scoring != 5
    if weighting in (0, 1) and ids and not must:
        vals = [v / float(len(ids)) for v in vals]
    if must:
        norm = 0.0
        if scoring == 1:
            for v in vals:
                norm += v * v
            norm = float(np.sqrt(np.float64(norm)))
        else:
            for v in vals:
                norm += abs(v)
        if norm > 0.0:
            vals = [v / norm for v in vals]
    nodes = sorted(fv)
    start = np.zeros(len(nodes) + 1, np.int32)
    items = []
    for s, nd in enumerate(nodes):
        start[s] = len(items)
        items += fv[nd]
    start[len(nodes)] = len(items)
    return dict(word_id=np.asarray(word, np.int32).reshape(-1), node_id=np.asarray(node, np.int32).reshape(-1),
                weight=np.asarray(wgt, np.float64).reshape(-1), bow_ids=np.asarray(ids, np.int32).reshape(-1),
                bow_values=np.asarray(vals, np.float64).reshape(-1), fv_node_ids=np.asarray(nodes, np.int32).reshape(-1),
                fv_node_start=start, fv_items=np.asarray(items, np.int32).reshape(-1))
