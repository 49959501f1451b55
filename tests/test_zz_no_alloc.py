"""ydorb_hamming_top2_device_noalloc (csrc/hamming_static.cu): the Hamming top-2 scan with its fold state in a static
device array instead of cudaMalloc'ed scratch.  Through the C ABI, against ydorb_hamming_top2_device on the same
inputs: bit-exact (integer work), on two calls in a row whose sizes differ, so the second call meets the first call's
leftovers in the shared array.  Collected last on purpose (file name)."""
import ctypes as C

import numpy as np
import pytest

from ydorbslam_b200 import OrbMatcher, synth
from ydorbslam_b200._lib import TOP2_DTYPE, YdorbError, check, lib

pytestmark = pytest.mark.gpu

MAX_Q = 16384  # YDORB_NOALLOC_MAX_QUERIES, include/ydorb.h


def _both(m, torch, q, db, idx_base):
    """(existing entry point's result, no-alloc entry point's result, free device memory before / after the latter)"""
    dev = torch.device("cuda:0")
    d_q = torch.from_numpy(np.ascontiguousarray(q)).to(dev)
    d_db = torch.from_numpy(np.ascontiguousarray(db)).to(dev)
    nq, ndb = len(q), len(db)
    d_ref = torch.full((max(nq, 1), 4), -7, dtype=torch.int32, device=dev)
    d_out = torch.full((max(nq, 1), 4), -7, dtype=torch.int32, device=dev)
    args = (C.c_void_p(d_q.data_ptr()), nq, C.c_void_p(d_db.data_ptr() if ndb else None), ndb, idx_base)
    check(lib().ydorb_hamming_top2_device(m._h, *args, C.c_void_p(d_ref.data_ptr())))
    m.sync()
    torch.cuda.synchronize()
    free0 = torch.cuda.mem_get_info(0)[0]
    check(lib().ydorb_hamming_top2_device_noalloc(m._h, *args, C.c_void_p(d_out.data_ptr())))
    m.sync()
    free1 = torch.cuda.mem_get_info(0)[0]
    return d_ref.cpu().numpy()[:nq], d_out.cpu().numpy()[:nq], free0, free1


def test_noalloc_top2_equals_the_allocating_entry_point_twice_in_a_row():
    torch = pytest.importorskip("torch")
    m = OrbMatcher()
    first = True
    # the second call is smaller than the first (stale state beyond and inside its range), the third one larger again
    for nq, ndb, base, seed in [(5000, 60000, 0, 1), (777, 3001, 1000, 2), (MAX_Q, 9000, 5, 3), (1, 1, 0, 4), (300, 0, 0, 5)]:
        db = synth.synth_descriptors(max(ndb, 1), seed)[:ndb]
        q = synth.flip_bits(synth.synth_descriptors(nq, seed + 10), 0.0, seed) if ndb == 0 else \
            synth.flip_bits(db[np.random.RandomState(seed).randint(0, ndb, nq)], 0.15, seed + 20)
        want, got, free0, free1 = _both(m, torch, q, db, base)
        assert np.array_equal(got, want), (nq, ndb)
        if ndb:
            t = want.view(TOP2_DTYPE).reshape(-1)
            assert (t["best_idx"] >= base).all() and (t["best_dist"] < 256).all()
        if not first:  # the very first launch may load the module (and with it the static array) lazily
            assert free1 >= free0, "the no-alloc entry point allocated device memory"
        first = False
    m.close()


def test_noalloc_top2_rejects_what_the_static_array_cannot_hold():
    torch = pytest.importorskip("torch")
    m = OrbMatcher()
    d = torch.zeros(32 * 64, dtype=torch.uint8, device="cuda:0")
    with pytest.raises(YdorbError, match="YDORB_NOALLOC_MAX_QUERIES"):
        check(lib().ydorb_hamming_top2_device_noalloc(m._h, C.c_void_p(d.data_ptr()), MAX_Q + 1, C.c_void_p(d.data_ptr()), 64, 0,
                                                      C.c_void_p(d.data_ptr())))
    m.close()
