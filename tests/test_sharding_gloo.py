"""CPU, world_size 2 over gloo: the multi-GPU host logic (shard ranges, one all-gather, order-free merge) gives the
result of ONE sequential scan over the whole database.  The per-shard scan and the merge are the oracle's here (the
GPU kernels take their place under NCCL, tests/test_gpu_hamming.py checks those against the same oracle)."""
import os
import socket

import numpy as np
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from ydorbslam_b200 import sharding, synth


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, nq, ndb, out_dir):
    from oracle import orc

    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    db = synth.synth_descriptors(ndb, 1)
    db[ndb // 2 + 3] = db[5]          # exact duplicate rows in different shards: lowest global index must win
    db[ndb - 1] = db[7]
    q, _ = synth.synth_query_descriptors(db, nq, 2)
    q[0] = db[5]
    q[1] = db[7]
    b, e = sharding.shard_range(ndb, rank, world)
    local = torch.zeros(nq, 4, dtype=torch.int32)
    gathered = torch.zeros(world, nq, 4, dtype=torch.int32)

    def scan():
        t = orc.top2(q, db[b:e], idx_base=b)
        local.copy_(torch.from_numpy(sharding.top2_records(t["best_idx"], t["best_dist"], t["second_dist"], t["second_idx"])))

    def merge(allp):
        recs = np.ascontiguousarray(allp.numpy()).view(orc.TOP2_DTYPE).reshape(world, nq)
        return orc.top2_merge(recs)

    merged = sharding.sharded_top2(scan, merge, local, gathered)
    np.save(os.path.join(out_dir, f"merged_{rank}.npy"), merged)
    # frame sharding: every frame belongs to exactly one rank
    mine = torch.zeros(37, dtype=torch.int32)
    mine[sharding.frames_of_rank(37, rank, world)] = 1
    dist.all_reduce(mine)
    assert bool((mine == 1).all())
    dist.destroy_process_group()


def test_sharded_top2_equals_single_scan(tmp_path):
    from oracle import orc

    world, nq, ndb = 2, 96, 5001
    mp.spawn(_worker, args=(world, _free_port(), nq, ndb, str(tmp_path)), nprocs=world, join=True)
    db = synth.synth_descriptors(ndb, 1)
    db[ndb // 2 + 3] = db[5]
    db[ndb - 1] = db[7]
    q, _ = synth.synth_query_descriptors(db, nq, 2)
    q[0] = db[5]
    q[1] = db[7]
    ref = orc.top2(q, db)
    for r in range(world):
        got = np.load(tmp_path / f"merged_{r}.npy")
        for f in ("best_idx", "best_dist", "second_dist"):
            assert np.array_equal(got[f], ref[f]), (r, f)
    assert ref["best_idx"][0] == 5 and ref["best_dist"][0] == 0 and ref["second_dist"][0] == 0
    assert ref["best_idx"][1] == 7


def test_shard_ranges_cover():
    for n in (0, 1, 7, 1000, 1_000_000):
        for world in (1, 2, 3, 4, 8):
            edges = [sharding.shard_range(n, r, world) for r in range(world)]
            assert edges[0][0] == 0 and edges[-1][1] == n
            assert all(edges[i][1] == edges[i + 1][0] for i in range(world - 1))
            sizes = [e - b for b, e in edges]
            assert max(sizes) - min(sizes) <= 1
