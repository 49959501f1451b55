"""One rank of the multi-process sharded Hamming test (tests/test_gpu_sharded.py): python -m tests.sharded_worker ...
Rank r drives GPU r through the C ABI only (no torch): communicator from a file-shipped ncclUniqueId, resident shard,
several collective calls (both call parities, varying nq); results go to <dir>/rank<r>.npz for the parent to check."""
import argparse
import os
import pathlib
import sys
import time

import numpy as np

ROOT = pathlib.Path(__file__).resolve().parents[1]
sys.path.insert(0, str(ROOT))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--rank", type=int, required=True)
    ap.add_argument("--world", type=int, required=True)
    ap.add_argument("--dir", required=True)
    ap.add_argument("--rows", type=int, default=200_000)
    ap.add_argument("--nq", type=int, default=2000)
    ap.add_argument("--calls", type=int, default=5)
    args = ap.parse_args()
    from ydorbslam_b200 import OrbMatcher, sharding, synth

    d = pathlib.Path(args.dir)
    idf = d / "nccl_id.bin"
    if args.rank == 0:
        uid = sharding.comm_unique_id()
        tmp = d / "nccl_id.tmp"
        tmp.write_bytes(uid)
        os.replace(tmp, idf)
    else:
        t0 = time.time()
        while not idf.exists():
            if time.time() - t0 > 120:
                raise SystemExit("timed out waiting for rank 0's id")
            time.sleep(0.05)
        uid = idf.read_bytes()
    m = OrbMatcher(0.6, device=args.rank)
    comm = sharding.Communicator(m, args.rank, args.world, args.nq, uid)
    db = synth.synth_descriptors(args.rows, 1234)
    lo, hi = sharding.shard_range(args.rows, args.rank, args.world)
    shard = sharding.Shard(m, db[lo:hi], lo)
    out = {"exchange": comm.exchange}
    for c in range(args.calls):
        nq = args.nq if c % 2 == 0 else max(1, args.nq - 37 * c)  # the same on every rank
        q, _ = synth.synth_query_descriptors(db, nq, 50 + c)
        t, acc = sharding.sharded_top2_host(comm, shard, q, 0.6, 50)
        out[f"top2_{c}"] = t
        out[f"acc_{c}"] = acc
    np.savez(d / f"rank{args.rank}.npz", **out)
    shard.close()
    comm.close()
    m.close()


if __name__ == "__main__":
    main()
