"""Host mirror of DBoW3::Vocabulary for the one call the front-end makes, transform(features, BowVector, FeatureVector,
levelsup) (thirdParty/DBow3/src/Vocabulary.cpp:752-816; called by Frame::computeBoW, src/frame.cpp:265-272), over the
C ABI of libydorb.so.  The tree descent and the assembly of both containers run on the GPU; there is no CPU path."""
import ctypes as C

import numpy as np

from ._lib import check, lib, ptr

TF_IDF, TF, IDF, BINARY = range(4)  # DBoW3::WeightingType (BowVector.h:38-44)
L1_NORM, L2_NORM, CHI_SQUARE, KL, BHATTACHARYYA, DOT_PRODUCT = range(6)  # DBoW3::ScoringType (BowVector.h:47-55)
MAX_FEATURES = 16384


class VocabularyDesc(C.Structure):
    _fields_ = [("n_nodes", C.c_int32), ("k", C.c_int32), ("L", C.c_int32), ("weighting", C.c_int32), ("scoring", C.c_int32),
                ("child_start", C.c_void_p), ("children", C.c_void_p), ("descriptors", C.c_void_p), ("weights", C.c_void_p),
                ("word_ids", C.c_void_p)]


class BowResult(C.Structure):
    _fields_ = [("word_id", C.c_void_p), ("node_id", C.c_void_p), ("weight", C.c_void_p), ("bow_ids", C.c_void_p),
                ("bow_values", C.c_void_p), ("n_bow", C.c_int32), ("fv_node_ids", C.c_void_p), ("fv_node_start", C.c_void_p),
                ("fv_items", C.c_void_p), ("n_fv_nodes", C.c_int32)]


class Vocabulary:
    """DBoW3::Vocabulary(k, L, weighting, scoring) holding an already-built tree."""

    def __init__(self, matcher, k, L, child_start, children, descriptors, weights, word_ids, weighting=TF_IDF,
                 scoring=L1_NORM, **_ignored):
        self._matcher = matcher  # keeps the stream / device alive
        a = (np.ascontiguousarray(child_start, np.int32), np.ascontiguousarray(children, np.int32),
             np.ascontiguousarray(descriptors, np.uint8), np.ascontiguousarray(weights, np.float64),
             np.ascontiguousarray(word_ids, np.int32))
        d = VocabularyDesc()
        d.n_nodes = len(a[3])
        d.k, d.L, d.weighting, d.scoring = int(k), int(L), int(weighting), int(scoring)
        d.child_start, d.children, d.descriptors, d.weights, d.word_ids = (x.ctypes.data for x in a)
        self._h = C.c_void_p()
        check(lib().ydorb_vocabulary_create(matcher._h, C.byref(d), C.byref(self._h)))

    @classmethod
    def load(cls, matcher, path):
        """Vocabulary::load of an uncompressed binary stream (Vocabulary.cpp:1082-1107 -> fromStream :1335-1409)."""
        self = cls.__new__(cls)
        self._matcher = matcher
        self._h = C.c_void_p()
        check(lib().ydorb_vocabulary_load(matcher._h, str(path).encode(), C.byref(self._h)))
        return self

    def close(self):
        if getattr(self, "_h", None):
            lib().ydorb_vocabulary_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def info(self):
        a = np.zeros(6, np.int32)
        check(lib().ydorb_vocabulary_info(self._h, ptr(a)))
        return dict(zip(("n_nodes", "k", "L", "weighting", "scoring", "n_words"), a.tolist()))

    def transform(self, features, levelsup=4):
        """features: (n, 32) uint8.  Returns a dict: per-feature word_id / node_id / weight, the BowVector as
        (bow_ids ascending, bow_values) and the FeatureVector in CSR form (fv_node_ids, fv_node_start, fv_items), which
        OrbMatcher.searchByBow* accept as (node_ids, node_start, items)."""
        d = np.ascontiguousarray(features, np.uint8).reshape(-1, 32)
        n = len(d)
        word, node, wgt = np.zeros(n, np.int32), np.zeros(n, np.int32), np.zeros(n, np.float64)
        bi, bv = np.zeros(n, np.int32), np.zeros(n, np.float64)
        fi, fs, ft = np.zeros(n, np.int32), np.zeros(n + 1, np.int32), np.zeros(n, np.int32)
        r = BowResult()
        r.word_id, r.node_id, r.weight, r.bow_ids, r.bow_values = (x.ctypes.data for x in (word, node, wgt, bi, bv))
        r.fv_node_ids, r.fv_node_start, r.fv_items = (x.ctypes.data for x in (fi, fs, ft))
        check(lib().ydorb_vocabulary_transform(self._h, ptr(d), n, int(levelsup), C.byref(r)))
        nb, nf = r.n_bow, r.n_fv_nodes
        return dict(word_id=word, node_id=node, weight=wgt, bow_ids=bi[:nb], bow_values=bv[:nb], fv_node_ids=fi[:nf],
                    fv_node_start=fs[: nf + 1], fv_items=ft[: int(fs[nf])])

    def transform_batch_device(self, d_desc, d_n, capacity, batch, levelsup, d_bow_ids, d_bow_values, d_fv_node_ids,
                               d_fv_node_start, d_fv_items, d_counts, d_word_id=None, d_node_id=None, d_weight=None):
        """Device pointers (ints); frame b owns rows [b*capacity, b*capacity + n[b]).  Asynchronous on the matcher's stream."""
        vp = lambda x: C.c_void_p(x) if x else None
        check(lib().ydorb_vocabulary_transform_batch_device(
            self._h, vp(d_desc), vp(d_n), int(capacity), int(batch), int(levelsup), vp(d_word_id), vp(d_node_id), vp(d_weight),
            vp(d_bow_ids), vp(d_bow_values), vp(d_fv_node_ids), vp(d_fv_node_start), vp(d_fv_items), vp(d_counts)))


def write_dbow3_binary(path, voc, weighting=TF_IDF, scoring=L1_NORM):
    """Writes `voc` (a synth.synth_vocabulary dict) in the layout of Vocabulary::toStream with compressed = false
    (Vocabulary.cpp:1180-1226): a test/tool helper so that ydorb_vocabulary_load can be exercised without a .dbow3 blob
    (the reference's own vocabulary file is not in the checkout)."""
    import struct

    cs, ch = voc["child_start"], voc["children"]
    out = [struct.pack("<QBI", 88877711233, 0, len(voc["weights"])), struct.pack("<iiii", voc["k"], voc["L"], scoring, weighting)]
    parents = [0]
    while parents:  # the same stack walk toStream does, so children reach fromStream in Node::children order
        pid = parents.pop()
        for j in range(cs[pid], cs[pid + 1]):
            c = int(ch[j])
            out.append(struct.pack("<IId", c, pid, float(voc["weights"][c])))
            out.append(struct.pack("<iii", 32, 1, 0))  # cols, rows, type = CV_8U (DescManip.cpp:249-258)
            out.append(voc["descriptors"][c].tobytes())
            if cs[c + 1] > cs[c]:
                parents.append(c)
    leaves = [i for i in range(len(voc["weights"])) if cs[i + 1] == cs[i]]
    out.append(struct.pack("<I", len(leaves)))
    for nid in sorted(leaves, key=lambda i: int(voc["word_ids"][i])):
        out.append(struct.pack("<II", int(voc["word_ids"][nid]), nid))
    with open(path, "wb") as f:
        f.write(b"".join(out))


def inflate_stream(data):
    """ydorb_vocabulary_stream_inflate: the plain form of a binary vocabulary stream (bytes in, bytes out).  Host only."""
    src = np.frombuffer(bytes(data), np.uint8)
    n = C.c_size_t(0)
    check(lib().ydorb_vocabulary_stream_inflate(ptr(src), C.c_size_t(len(src)), None, C.c_size_t(0), C.byref(n)))
    out = np.zeros(n.value, np.uint8)
    check(lib().ydorb_vocabulary_stream_inflate(ptr(src), C.c_size_t(len(src)), ptr(out), C.c_size_t(len(out)), C.byref(n)))
    return out[: n.value].tobytes()


def read_dbow3_binary(data):
    """Parses a PLAIN binary vocabulary stream (bytes) into the dict form synth.synth_vocabulary uses, the way
    Vocabulary::fromStream does (Vocabulary.cpp:1371-1405): a test/tool helper, the product loader is ydorb_vocabulary_load."""
    import struct

    sig, comp, n = struct.unpack_from("<QBI", data, 0)
    assert sig == 88877711233 and comp == 0
    k, L, scoring, weighting = struct.unpack_from("<iiii", data, 13)
    at = 29
    parent, weights, desc = np.full(n, -1, np.int64), np.zeros(n, np.float64), np.zeros((n, 32), np.uint8)
    order = []
    for _ in range(n - 1):
        nid, pid, w, cols, rows, typ = struct.unpack_from("<IIdiii", data, at)
        assert (cols, rows, typ) == (32, 1, 0)
        desc[nid] = np.frombuffer(data, np.uint8, 32, at + 28)
        parent[nid], weights[nid] = pid, w
        order.append(nid)
        at += 60
    (nw,) = struct.unpack_from("<I", data, at)
    words = np.frombuffer(data, np.uint32, 2 * nw, at + 4).reshape(nw, 2)
    word_ids = np.zeros(n, np.int32)
    word_ids[words[:, 1]] = words[:, 0]
    kids = [[] for _ in range(n)]
    for c in order:
        kids[parent[c]].append(c)
    child_start = np.zeros(n + 1, np.int32)
    child_start[1:] = np.cumsum([len(x) for x in kids])
    children = np.asarray([c for x in kids for c in x], np.int32)
    return dict(k=k, L=L, scoring=scoring, weighting=weighting, child_start=child_start, children=children, descriptors=desc,
                weights=weights, word_ids=word_ids)
