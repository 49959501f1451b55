"""GPU parity: DBoW3::Vocabulary::transform through the C ABI vs the CPU oracle.  Bit-exact on every
field, the double-precision word values included (compared as 64-bit patterns)."""
import numpy as np
import pytest

from oracle import orc
from tests import bow_cases, matcher_cases
from ydorbslam_b200 import OrbMatcher, synth
from ydorbslam_b200.vocabulary import Vocabulary, write_dbow3_binary

pytestmark = pytest.mark.gpu


def _orc_voc(v, weighting, scoring):
    return orc.Vocabulary.from_arrays(v["k"], v["L"], weighting, scoring, v["child_start"], v["children"], v["descriptors"],
                                      v["weights"], v["word_ids"])


@pytest.mark.parametrize("k,L,prune,shuffle", bow_cases.TREES)
@pytest.mark.parametrize("weighting,scoring", [(0, 0), (1, 1), (2, 5), (3, 0), (0, 5), (1, 3)])
def test_transform_vs_oracle(k, L, prune, shuffle, weighting, scoring):
    v = bow_cases.tree(k, L, prune, shuffle)
    m = OrbMatcher()
    voc = Vocabulary(m, weighting=weighting, scoring=scoring, **v)
    ov = _orc_voc(v, weighting, scoring)
    for n, levels_up in [(0, 4), (1, 0), (31, 1), (1000, L - 1), (1025, L), (3000, L + 2), (2048, 1)]:
        q = synth.synth_vocabulary_queries(v, max(n, 1), seed=n + 1)[:n]
        bow_cases.same(voc.transform(q, levels_up), orc.voc_transform(ov, q, levels_up))
    assert voc.info()["n_words"] == int((np.diff(v["child_start"]) == 0).sum())


def test_all_features_in_one_word_and_all_stopped():
    """Degenerate accumulations: 5000 identical descriptors (one word, 5000 sequential += in double), and a vocabulary
    whose every word is stopped (w = 0: both containers stay empty)."""
    v = bow_cases.tree(10, 3, 0.0, False)
    m = OrbMatcher()
    q = np.tile(synth.synth_vocabulary_queries(v, 1, seed=3), (5000, 1))
    for weighting, scoring in [(0, 0), (1, 5), (0, 1)]:
        got = Vocabulary(m, weighting=weighting, scoring=scoring, **v).transform(q, 2)
        bow_cases.same(got, orc.voc_transform(_orc_voc(v, weighting, scoring), q, 2))
    if got["weight"][0] > 0:
        assert len(got["bow_ids"]) == 1 and len(got["fv_items"]) == 5000
    z = dict(v, weights=np.zeros_like(v["weights"]))
    got = Vocabulary(m, **z).transform(q[:100], 2)
    bow_cases.same(got, orc.voc_transform(_orc_voc(z, 0, 0), q[:100], 2))
    assert len(got["bow_ids"]) == 0 and len(got["fv_node_ids"]) == 0


def test_full_size_vocabulary_k10_L6():
    """The shape the reference loads (k = 10, L = 6: 1 111 111 nodes, 10^6 words) with 8000 descriptors (config 5's
    keypoint count) and levelsup = 4, the call Frame::computeBoW makes (src/frame.cpp:270)."""
    v = synth.synth_vocabulary_full(10, 6, seed=31)
    q = synth.synth_vocabulary_queries(v, 8000, seed=32)
    m = OrbMatcher()
    voc = Vocabulary(m, **v)
    got = voc.transform(q, 4)
    ref = orc.voc_transform(_orc_voc(v, 0, 0), q, 4)
    bow_cases.same(got, ref)
    assert len(got["fv_node_ids"]) <= 100 and len(got["bow_ids"]) > 3000
    with pytest.raises(Exception):
        voc.transform(np.zeros((16385, 32), np.uint8), 4)


def test_load_binary_stream(tmp_path):
    """ydorb_vocabulary_load parses what Vocabulary::toStream(compressed = false) writes; the loaded tree transforms
    exactly like the one created from arrays.  Compressed / foreign files are refused."""
    v = bow_cases.tree(3, 4, 0.25, True)
    p = tmp_path / "voc.dbow3"
    write_dbow3_binary(p, v, weighting=1, scoring=1)
    m = OrbMatcher()
    voc = Vocabulary.load(m, p)
    info = voc.info()
    assert (info["k"], info["L"], info["weighting"], info["scoring"], info["n_nodes"]) == (3, 4, 1, 1, len(v["weights"]))
    q = synth.synth_vocabulary_queries(v, 777, seed=8)
    bow_cases.same(voc.transform(q, 2), orc.voc_transform(_orc_voc(v, 1, 1), q, 2))
    raw = bytearray(p.read_bytes())
    raw[8] = 1  # claims to be compressed, is not
    (tmp_path / "c.dbow3").write_bytes(bytes(raw))
    with pytest.raises(Exception, match="compressed vocabulary stream"):
        Vocabulary.load(m, tmp_path / "c.dbow3")
    (tmp_path / "t.dbow3").write_bytes(bytes(raw[: len(raw) // 2]))
    with pytest.raises(Exception):
        Vocabulary.load(m, tmp_path / "t.dbow3")
    with pytest.raises(Exception):
        Vocabulary.load(m, tmp_path / "missing.dbow3")


def test_batch_device_matches_per_frame():
    import torch

    v = bow_cases.tree(10, 3, 0.0, False)
    m = OrbMatcher()
    voc = Vocabulary(m, **v)
    ov = _orc_voc(v, 0, 0)
    B, cap = 5, 1200
    ns = np.array([1200, 0, 7, 1199, 600], np.int32)
    desc = np.zeros((B, cap, 32), np.uint8)
    for b in range(B):
        desc[b, : ns[b]] = synth.synth_vocabulary_queries(v, cap, seed=40 + b)[: ns[b]]
    dev = torch.device("cuda:0")
    t = lambda a: torch.from_numpy(a).to(dev)
    d_desc, d_n = t(desc), t(ns)
    bi = torch.zeros((B, cap), dtype=torch.int32, device=dev)
    bv = torch.zeros((B, cap), dtype=torch.float64, device=dev)
    fi = torch.zeros((B, cap), dtype=torch.int32, device=dev)
    fs = torch.zeros((B, cap + 1), dtype=torch.int32, device=dev)
    ft = torch.zeros((B, cap), dtype=torch.int32, device=dev)
    cnt = torch.zeros((B, 2), dtype=torch.int32, device=dev)
    torch.cuda.synchronize()
    voc.transform_batch_device(d_desc.data_ptr(), d_n.data_ptr(), cap, B, 2, bi.data_ptr(), bv.data_ptr(), fi.data_ptr(),
                               fs.data_ptr(), ft.data_ptr(), cnt.data_ptr())
    m.sync()
    cnt = cnt.cpu().numpy()
    for b in range(B):
        ref = orc.voc_transform(ov, desc[b, : ns[b]], 2)
        nb, nf = cnt[b]
        assert nb == len(ref["bow_ids"]) and nf == len(ref["fv_node_ids"])
        assert np.array_equal(bi[b, :nb].cpu().numpy(), ref["bow_ids"])
        assert np.array_equal(bv[b, :nb].cpu().numpy().view(np.uint64), ref["bow_values"].view(np.uint64))
        assert np.array_equal(fi[b, :nf].cpu().numpy(), ref["fv_node_ids"])
        assert np.array_equal(fs[b, : nf + 1].cpu().numpy(), ref["fv_node_start"])
        assert np.array_equal(ft[b, : len(ref["fv_items"])].cpu().numpy(), ref["fv_items"])


def test_feature_vector_feeds_search_by_bow():
    """The chain the reference runs: computeBoW on two key frames, then searchByBowInTwoKeyFrames on their
    FeatureVectors (src/loopClosing.cpp:119-206).  GPU transform -> GPU matcher equals oracle transform -> oracle matcher."""
    v = bow_cases.tree(10, 3, 0.0, False)
    m = OrbMatcher(0.75, True)
    voc = Vocabulary(m, **v)
    ov = _orc_voc(v, 0, 0)
    n = 900
    d1 = synth.synth_vocabulary_queries(v, n, seed=51)
    d2 = synth.flip_bits(d1, 0.03, 52)[::-1].copy()
    kps1, kps2 = matcher_cases.synth_keypoints(n, 53)[0], matcher_cases.synth_keypoints(n, 54)[0]
    valid = np.ones(n, np.uint8)
    g1, g2 = voc.transform(d1, 2), voc.transform(d2, 2)
    o1, o2 = orc.voc_transform(ov, d1, 2), orc.voc_transform(ov, d2, 2)
    fv = lambda r: (r["fv_node_ids"], r["fv_node_start"], r["fv_items"])
    n_got, got = m.searchByBowInTwoKeyFrames(kps1, d1, valid, fv(g1), kps2, d2, valid, fv(g2))
    n_ref, ref = orc.search_by_bow(1, kps1, d1, valid, fv(o1), kps2, d2, valid, fv(o2), 0.75, 1)
    assert n_got == n_ref and np.array_equal(got, ref)
    assert (got >= 0).sum() > 100



def test_load_what_the_reference_saves(tmp_path):
    """Vocabulary::save of the compiled reference (oracle/_ref), in its default QuickLZ-compressed form and plain, of a
    vocabulary its own k-means trained -> ydorb_vocabulary_load -> transform == the reference's own transform."""
    from oracle import ref
    from tests.test_oracle_ref_vocabulary import same_where_defined, trained
    from ydorbslam_b200.vocabulary import read_dbow3_binary

    if not ref.available():
        pytest.skip("oracle/_ref not shipped")
    rv, desc, v = trained(10, 4, 0, 0, 17, tmp_path)
    m = OrbMatcher()
    undefined = 0
    for name in ("packed.dbow3", "plain.dbow3"):
        voc = Vocabulary.load(m, tmp_path / name)
        assert voc.info()["n_words"] == rv.info()["n_words"]
        for n, levels_up in [(2000, 2), (1, 4), (3000, 0), (777, 3)]:
            q = synth.flip_bits(desc[:n], 0.1, n)
            undefined += same_where_defined(rv.transform(q, levels_up), voc.transform(q, levels_up), v, levels_up)
    rv.close()
