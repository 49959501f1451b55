"""GPU: the sharded Hamming top-2 behind the C ABI (csrc/shard.cu).  One rank per GPU, merged result checked against ONE
sequential scan of the oracle over the whole database (rule of src/orbMatcher.cpp:327-336).  With a single GPU only the
world-size-1 path runs here; the 2/4/8-rank runs need `gpurun --gpus N` (their logs are kept under profiles/) and
bench.py repeats the same check on every rank before it times the N-rank path."""
import os
import subprocess
import sys
import tempfile

import numpy as np
import pytest

from oracle import orc
from ydorbslam_b200 import OrbMatcher, sharding, synth
from ydorbslam_b200._lib import lib

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _expected(db, q):
    t = orc.top2(q, db)
    return t, orc.ratio_accept(t, 0.6, 50)


def _same(t, acc, et, eacc):
    for f in ("best_idx", "best_dist", "second_dist"):
        assert np.array_equal(t[f], et[f]), f
    assert np.array_equal(acc, eacc)


def test_sharded_world1_equals_single_scan():
    m = OrbMatcher(0.6)
    db = synth.synth_descriptors(150_000, 1234)
    comm = sharding.Communicator(m, 0, 1, 2000)
    shard = sharding.Shard(m, db, 0)
    assert comm.exchange == "single"
    for c, nq in enumerate([2000, 1, 513, 1999]):
        q, _ = synth.synth_query_descriptors(db, nq, 90 + c)
        t, acc = sharding.sharded_top2_host(comm, shard, q, 0.6, 50)
        _same(t, acc, *_expected(db, q))
    shard.close()
    comm.close()
    m.close()


def _run_ranks(world, mode, rows=200_000, nq=2000, calls=5):
    with tempfile.TemporaryDirectory() as d:
        env = dict(os.environ, YDORB_SHARD_EXCHANGE=mode, YDORB_SHARD_TIMEOUT_MS="20000")
        procs = [subprocess.Popen([sys.executable, "-m", "tests.sharded_worker", "--rank", str(r), "--world", str(world), "--dir", d,
                                   "--rows", str(rows), "--nq", str(nq), "--calls", str(calls)], cwd=ROOT, env=env)
                 for r in range(world)]
        for p in procs:
            assert p.wait(timeout=300) == 0
        return [dict(np.load(os.path.join(d, f"rank{r}.npz"))) for r in range(world)]


@pytest.mark.parametrize("mode", ["p2p", "nccl"])
@pytest.mark.parametrize("world", [2, 4, 8])
def test_sharded_ranks_equal_single_scan(world, mode):
    if lib().ydorb_device_count() < world:
        pytest.skip(f"needs {world} GPUs")
    rows, nq, calls = 200_000, 2000, 5
    res = _run_ranks(world, mode, rows, nq, calls)
    db = synth.synth_descriptors(rows, 1234)
    for r in range(world):
        assert str(res[r]["exchange"]) == mode, (str(res[r]["exchange"]), mode)
    for c in range(calls):
        n = nq if c % 2 == 0 else max(1, nq - 37 * c)
        q, _ = synth.synth_query_descriptors(db, n, 50 + c)
        et, eacc = _expected(db, q)
        for r in range(world):
            _same(res[r][f"top2_{c}"], res[r][f"acc_{c}"], et, eacc)
