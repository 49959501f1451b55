"""N1 against the REFERENCE's own DBoW3 (oracle/_ref: thirdParty/DBow3/src/Vocabulary.cpp, DescManip.cpp, quicklz.c
compiled unmodified): Vocabulary::transform == the oracle, on trees the reference's own k-means builds and on trees the
reference loads from streams this repo writes; and the binary stream format both ways, the QuickLZ-compressed form
included (inflated by ydorb_vocabulary_stream_inflate, host only).  CPU only."""
import numpy as np
import pytest

from oracle import orc, ref
from tests import bow_cases
from ydorbslam_b200 import synth
from ydorbslam_b200.vocabulary import inflate_stream, read_dbow3_binary, write_dbow3_binary

pytestmark = pytest.mark.skipif(not ref.available(), reason="oracle/_ref not built and /root/reference absent")


def _orc_voc(v, weighting, scoring):
    return orc.Vocabulary.from_arrays(v["k"], v["L"], weighting, scoring, v["child_start"], v["children"], v["descriptors"],
                                      v["weights"], v["word_ids"])


def trained(k, L, weighting, scoring, seed, tmp_path):
    """A vocabulary trained by the reference on clustered descriptors, and its plain stream parsed back."""
    rng = np.random.RandomState(seed)
    centres = synth.synth_descriptors(40, seed)
    n_images, per = 12, 250
    desc = np.concatenate([synth.flip_bits(centres[rng.randint(0, 40, per)], 0.06, seed + 1 + i) for i in range(n_images)])
    rv = ref.RefVocabulary.create(k, L, weighting, scoring, desc, np.arange(n_images + 1) * per, seed)
    rv.save(tmp_path / "plain.dbow3", False)
    rv.save(tmp_path / "packed.dbow3", True)
    return rv, desc, read_dbow3_binary((tmp_path / "plain.dbow3").read_bytes())


@pytest.mark.parametrize("k,L,weighting,scoring", [(10, 3, 0, 0), (4, 5, 1, 1), (7, 2, 2, 5), (3, 6, 3, 0), (10, 4, 0, 5)])
def test_reference_trained_vocabulary_transform_equals_oracle(k, L, weighting, scoring, tmp_path):
    rv, desc, v = trained(k, L, weighting, scoring, 5 + k, tmp_path)
    info = rv.info()
    assert (v["k"], v["L"], v["weighting"], v["scoring"]) == (k, L, weighting, scoring) == (info["k"], info["L"], info["weighting"], info["scoring"])
    assert int((np.diff(v["child_start"]) == 0).sum()) - (1 if v["child_start"][1] == 0 else 0) == info["n_words"]
    ov = _orc_voc(v, weighting, scoring)
    rng = np.random.RandomState(1)
    for n, levels_up in [(0, 4), (1, 0), (500, 1), (2000, L - 1), (1500, L), (800, L + 2), (1000, 4)]:
        q = np.concatenate([synth.flip_bits(desc[rng.randint(0, len(desc), max(n, 1))], 0.1, n + 3), synth.synth_descriptors(8, n)])[:n]
        same_where_defined(rv.transform(q, levels_up), orc.voc_transform(ov, q, levels_up), v, levels_up)
    rv.close()


def node_depth(v):
    depth = np.zeros(len(v["weights"]), np.int32)
    cs, ch = v["child_start"], v["children"]
    stack = [0]
    while stack:
        p = stack.pop()
        for c in ch[cs[p] : cs[p + 1]]:
            depth[c] = depth[p] + 1
            stack.append(int(c))
    return depth


def same_where_defined(got, want, v, levels_up):
    """Bit-exact on everything the reference defines.  Vocabulary::transform leaves the caller's `NodeId nid`
    UNWRITTEN when the word sits above level L - levelsup (Vocabulary.cpp:846-867 never reaches nid_level; the variable
    at :774 / :804 has no initialiser), so the FeatureVector group of such a feature is whatever the stack held; the
    oracle and the CUDA path return node 0 there.  k-means trees have such shallow words, balanced ones do not."""
    nid_level = v["L"] - levels_up
    word_node = np.zeros(int(v["word_ids"].max()) + 1, np.int64)
    leaves = np.flatnonzero(np.diff(v["child_start"]) == 0)
    word_node[v["word_ids"][leaves]] = leaves
    defined = (nid_level <= 0) | (node_depth(v)[word_node[want["word_id"]]] >= nid_level)
    for f in ("word_id", "weight", "bow_ids", "bow_values"):
        x, y = np.asarray(got[f]), np.asarray(want[f])
        assert x.shape == y.shape and np.array_equal(x.view(np.uint8), y.view(np.uint8)), f
    assert np.array_equal(got["node_id"][defined], want["node_id"][defined])
    assert np.all(want["node_id"][~defined] == 0)

    def groups(r):
        node = np.full(len(r["word_id"]), -1, np.int64)
        for g, nid in enumerate(r["fv_node_ids"]):
            items = r["fv_items"][r["fv_node_start"][g] : r["fv_node_start"][g + 1]]
            assert np.all(np.diff(items) > 0)  # ascending feature index inside a group
            node[items] = nid
        return node

    a, b = groups(got), groups(want)
    assert np.array_equal(a >= 0, b >= 0) and np.array_equal(a[defined], b[defined])
    return int((~defined).sum())


@pytest.mark.parametrize("k,L,prune,shuffle", bow_cases.TREES)
def test_reference_loads_this_repos_stream(k, L, prune, shuffle, tmp_path):
    """write_dbow3_binary -> Vocabulary::load of the reference -> its transform == the oracle on the same tree."""
    v = bow_cases.tree(k, L, prune, shuffle)
    write_dbow3_binary(tmp_path / "v.dbow3", v, weighting=1, scoring=1)
    rv = ref.RefVocabulary.load(tmp_path / "v.dbow3")
    q = synth.synth_vocabulary_queries(v, 1200, seed=4)
    same_where_defined(rv.transform(q, 2), orc.voc_transform(_orc_voc(v, 1, 1), q, 2), v, 2)
    rv.save(tmp_path / "again.dbow3", False)  # and the reference writes back the bytes it read
    assert (tmp_path / "again.dbow3").read_bytes() == (tmp_path / "v.dbow3").read_bytes()
    rv.close()


@pytest.mark.parametrize("k,L", [(10, 3), (4, 5), (10, 5)])
def test_compressed_stream_inflates_to_the_plain_one(k, L, tmp_path):
    """Vocabulary::save(compressed = true) -- the default, QuickLZ level 1 in 10 000-byte chunks -- inflated by the
    library == the reference's own plain stream of the same vocabulary, byte for byte."""
    rv, _, _ = trained(k, L, 0, 0, 21 + L, tmp_path)
    plain, packed = (tmp_path / "plain.dbow3").read_bytes(), (tmp_path / "packed.dbow3").read_bytes()
    assert packed[8] == 1 and len(packed) < len(plain)
    assert inflate_stream(packed) == plain
    assert inflate_stream(plain) == plain
    rv.close()


def test_compressed_stream_of_a_large_synthetic_tree(tmp_path):
    """111 111 nodes (k = 10, L = 5): 660+ chunks, random descriptors (poorly compressible: stored chunks and long
    literal runs) next to highly repetitive word tables (long matches)."""
    v = synth.synth_vocabulary_full(10, 5, seed=3)
    write_dbow3_binary(tmp_path / "v.dbow3", v)
    rv = ref.RefVocabulary.load(tmp_path / "v.dbow3")
    rv.save(tmp_path / "packed.dbow3", True)
    assert inflate_stream((tmp_path / "packed.dbow3").read_bytes()) == (tmp_path / "v.dbow3").read_bytes()
    rv.close()


def test_corrupt_compressed_streams_are_refused(tmp_path):
    rv, _, _ = trained(4, 3, 0, 0, 9, tmp_path)
    packed = bytearray((tmp_path / "packed.dbow3").read_bytes())
    rv.close()
    for cut in (12, 15, 40, len(packed) - 7):
        with pytest.raises(Exception):
            inflate_stream(bytes(packed[:cut]))
    bad = bytearray(packed)
    bad[0] ^= 1
    with pytest.raises(Exception, match="signature"):
        inflate_stream(bytes(bad))
    bad = bytearray(packed)
    bad[17] = (bad[17] & ~0x0C) | (3 << 2)  # claims compression level 3
    with pytest.raises(Exception, match="level"):
        inflate_stream(bytes(bad))
