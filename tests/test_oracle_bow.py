"""CPU: the oracle's DBoW3 transform against an independent pure-Python restatement, plus the
properties the reference's containers guarantee (sorted keys, each kept feature in exactly one node, unit L1 norm)."""
import numpy as np
import pytest

from oracle import orc
from tests import bow_cases
from ydorbslam_b200 import synth


def _orc_voc(v, weighting, scoring):
    return orc.Vocabulary.from_arrays(v["k"], v["L"], weighting, scoring, v["child_start"], v["children"], v["descriptors"],
                                      v["weights"], v["word_ids"])


@pytest.mark.parametrize("k,L,prune,shuffle", bow_cases.TREES)
@pytest.mark.parametrize("weighting,scoring", [(0, 0), (1, 1), (2, 5), (3, 0), (0, 5), (1, 3)])
def test_oracle_vs_python(k, L, prune, shuffle, weighting, scoring):
    v = bow_cases.tree(k, L, prune, shuffle)
    q = synth.synth_vocabulary_queries(v, 300, seed=k * 10 + L)
    for levels_up in (0, 1, L - 1, L, L + 2):
        got = orc.voc_transform(_orc_voc(v, weighting, scoring), q, levels_up)
        bow_cases.same(got, bow_cases.python_transform(v, q, levels_up, weighting, scoring))


def test_container_properties():
    v = bow_cases.tree(10, 3, 0.0, False)
    q = synth.synth_vocabulary_queries(v, 2000, seed=5)
    r = orc.voc_transform(_orc_voc(v, 0, 0), q, 1)
    assert (np.diff(r["bow_ids"]) > 0).all() and (np.diff(r["fv_node_ids"]) > 0).all()
    kept = np.flatnonzero(r["weight"] > 0)
    assert np.array_equal(np.sort(r["fv_items"]), kept)  # every kept feature in exactly one node
    for s in range(len(r["fv_node_ids"])):
        items = r["fv_items"][r["fv_node_start"][s] : r["fv_node_start"][s + 1]]
        assert (np.diff(items) > 0).all() and (r["node_id"][items] == r["fv_node_ids"][s]).all()
    assert abs(r["bow_values"].sum() - 1.0) < 1e-12  # L1-normalised
    assert 0 < len(kept) < len(q)  # stopped words present
    # ties happen (two children at the same distance): the first child must win
    assert len(np.unique(r["word_id"])) > 100


def test_empty_input():
    v = bow_cases.tree(4, 1, 0.0, False)
    r = orc.voc_transform(_orc_voc(v, 0, 0), np.zeros((0, 32), np.uint8), 4)
    assert len(r["bow_ids"]) == 0 and len(r["fv_node_ids"]) == 0 and r["fv_node_start"].tolist() == [0]
