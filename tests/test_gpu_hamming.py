"""GPU parity: Hamming distance, dense top-2, shard merge, ratio test vs the CPU oracle (bit-exact)."""
import numpy as np
import pytest

from oracle import orc
from ydorbslam_b200 import OrbMatcher, synth

pytestmark = pytest.mark.gpu


def test_pairs():
    a = synth.synth_descriptors(5000, 1)
    b = synth.flip_bits(a, 0.2, 2)
    m = OrbMatcher()
    d = m.computeDescriptorsDistance(a, b)
    ref = np.array([orc.hamming(x, y) for x, y in zip(a, b)], np.int32)
    assert np.array_equal(d, ref)
    assert m.computeDescriptorsDistance(a[0], a[0]) == 0
    assert m.computeDescriptorsDistance(np.zeros(32, np.uint8), np.full(32, 255, np.uint8)) == 256


@pytest.mark.parametrize("nq,ndb", [(1, 1), (7, 3), (513, 1000), (2000, 50000), (100, 0), (33, 257)])
def test_top2_vs_oracle(nq, ndb):
    db = synth.synth_descriptors(max(ndb, 1), 3)[:ndb]
    q, _ = synth.synth_query_descriptors(db, nq, 4) if ndb else (synth.synth_descriptors(nq, 5), None)
    q = q[:nq]
    m = OrbMatcher()
    got = m.top2(q, db, idx_base=17)
    ref = orc.top2(q, db, idx_base=17)
    for f in ("best_idx", "best_dist", "second_dist", "second_idx"):
        assert np.array_equal(got[f], ref[f]), f


def test_ties_first_wins():
    """All database rows identical: best = first row, second = second row, equal distances."""
    db = np.tile(synth.synth_descriptors(1, 9), (5000, 1))
    q = synth.synth_descriptors(40, 10)
    m = OrbMatcher()
    got = m.top2(q, db)
    assert (got["best_idx"] == 0).all() and (got["second_idx"] == 1).all()
    assert np.array_equal(got["best_dist"], got["second_dist"])
    ref = orc.top2(q, db)
    assert np.array_equal(got, ref)


def test_shard_merge_equals_single_scan():
    import torch

    db = synth.synth_descriptors(40000, 11)
    q, _ = synth.synth_query_descriptors(db, 700, 12)
    m = OrbMatcher()
    full = m.top2(q, db)
    G = 8
    bounds = [len(db) * g // G for g in range(G + 1)]
    parts = np.stack([m.top2(q, db[bounds[g]:bounds[g + 1]], idx_base=bounds[g]) for g in range(G)])
    # oracle merge
    assert np.array_equal(orc.top2_merge(parts), orc.top2(q, db))
    # device merge
    dev = torch.device("cuda:0")
    t_parts = torch.from_numpy(parts.view(np.int32).reshape(G, len(q), 4)).to(dev)
    t_out = torch.zeros(len(q), 4, dtype=torch.int32, device=dev)
    t_acc = torch.zeros(len(q), dtype=torch.uint8, device=dev)
    torch.cuda.synchronize()
    m.top2_merge_device(t_parts.data_ptr(), G, len(q), t_out.data_ptr())
    m.ratio_accept_device(t_out.data_ptr(), len(q), t_acc.data_ptr(), ratio=0.6, th=50)
    m.sync()
    merged = t_out.cpu().numpy().view(orc.TOP2_DTYPE).reshape(-1)
    assert np.array_equal(merged, full)
    assert np.array_equal(t_acc.cpu().numpy(), orc.ratio_accept(full, 0.6, 50))
    assert 0 < t_acc.sum().item() < len(q)
