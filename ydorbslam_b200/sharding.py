"""Multi-GPU host logic (one process per GPU, torch.distributed for the plumbing) -- SURVEY 8e.

* Extraction / stereo pairs: frames are independent; frame i belongs to rank i mod G.  No collective.
* Hamming top-2 against a large map-descriptor database: rows split contiguously, shard g owns
  [g*n/G, (g+1)*n/G); queries are replicated; every rank scans its shard, the per-shard YdTop2 records (16 B per
  query) travel through ONE all-gather, and an order-free merge reproduces the result of a single sequential scan
  (`if d < best ... else if d < second`, src/orbMatcher.cpp:327-333): best = minimum (distance, global index),
  second = second order statistic of the multiset {best_g, second_g}.

The compute steps are injected (`scan_fn`, `merge_fn`) so the same code runs on NCCL with the CUDA kernels
(`GpuShardedMatcher`) and, in the CPU tests, on gloo with stand-ins.
"""
import numpy as np


def shard_range(n_rows, rank, world):
    """Contiguous row range [begin, end) of shard `rank` out of `world`."""
    return (rank * n_rows) // world, ((rank + 1) * n_rows) // world


def frames_of_rank(n_frames, rank, world):
    """Frame indices extracted by `rank` (frame i -> GPU i mod G)."""
    return list(range(rank, n_frames, world))


def sharded_top2(scan_fn, merge_fn, local_out, gathered, group=None):
    """scan_fn() fills `local_out` (torch int32 [nq, 4] = YdTop2 records with GLOBAL indices) for this rank's shard;
    `gathered` is a torch int32 [world, nq, 4] buffer on the same device; merge_fn(gathered) returns the merged
    [nq, 4] result, identical on every rank."""
    import torch.distributed as dist

    scan_fn()
    world = gathered.shape[0]
    if world == 1:
        gathered[0].copy_(local_out)
    else:
        # gloo only takes the concatenated [world*nq, 4] form of the output; NCCL takes either
        dist.all_gather_into_tensor(gathered.view(world * local_out.shape[0], -1), local_out, group=group)
    return merge_fn(gathered)


class GpuShardedMatcher:
    """Config 4: 2000 queries vs a 1 M-row database sharded over the ranks (NCCL all-gather of 32 KB per rank)."""

    def __init__(self, matcher, d_db, rows, idx_base, nq, world, device, group=None):
        import torch

        self.m, self.d_db, self.rows, self.base, self.nq, self.world, self.group = matcher, d_db, rows, idx_base, nq, world, group
        self.d_part = torch.zeros(nq, 4, dtype=torch.int32, device=device)
        self.d_all = torch.zeros(world, nq, 4, dtype=torch.int32, device=device)
        self.d_out = torch.zeros(nq, 4, dtype=torch.int32, device=device)

    def top2(self, d_queries):
        """d_queries: torch uint8 [nq, 32] on the device.  Returns torch int32 [nq, 4] (YdTop2), same on all ranks."""

        def scan():
            self.m.top2_device(d_queries.data_ptr(), self.nq, self.d_db.data_ptr(), self.rows, self.base, self.d_part.data_ptr())

        def merge(allp):
            if self.world == 1:
                return self.d_part
            self.m.top2_merge_device(allp.data_ptr(), self.world, self.nq, self.d_out.data_ptr())
            return self.d_out

        return sharded_top2(scan, merge, self.d_part, self.d_all, self.group)


def top2_records(best_idx, best_dist, second_dist, second_idx):
    """Pack four int arrays into YdTop2 records [n, 4] int32 (field order of include/ydorb.h)."""
    return np.stack([best_idx, best_dist, second_dist, second_idx], axis=1).astype(np.int32)


# ---- the library's own multi-GPU path (csrc/shard.cu): communicator, resident shard, sharded top-2 behind the C ABI ----
def comm_unique_id():
    """rank 0: the ncclUniqueId (bytes) every rank must pass to Communicator; ship it over any host channel."""
    import ctypes as C

    from ._lib import check, lib

    buf = (C.c_uint8 * 128)()
    check(lib().ydorb_comm_unique_id(buf, 128))
    return bytes(buf)


class Communicator:
    """ydorb_comm_create: collective over `world` processes, one GPU each (the matcher's device and stream)."""

    EXCHANGE = {0: "single", 1: "nccl", 2: "p2p"}

    def __init__(self, matcher, rank, world, max_queries, unique_id=None):
        import ctypes as C

        from ._lib import check, lib

        self.matcher, self.rank, self.world, self.max_queries = matcher, rank, world, max_queries
        self._h = C.c_void_p()
        idbuf = None if unique_id is None else (C.c_uint8 * 128).from_buffer_copy(unique_id.ljust(128, b"\0"))
        check(lib().ydorb_comm_create(matcher._h, idbuf, rank, world, max_queries, C.byref(self._h)))

    @property
    def exchange(self):
        from ._lib import lib

        return self.EXCHANGE[lib().ydorb_comm_exchange_mode(self._h)]

    def check(self):
        from ._lib import check, lib

        check(lib().ydorb_comm_check(self._h))

    def close(self):
        from ._lib import lib

        if self._h:
            lib().ydorb_comm_destroy(self._h)
            self._h = None


class Shard:
    """This rank's rows [idx_base, idx_base + len(rows)) of the database, resident on the GPU (ydorb_shard_create)."""

    def __init__(self, matcher, rows, idx_base):
        import ctypes as C

        from ._lib import check, lib, ptr

        rows = np.ascontiguousarray(rows, dtype=np.uint8)
        assert rows.ndim == 2 and rows.shape[1] == 32
        self.n_rows, self.idx_base = len(rows), idx_base
        self._h = C.c_void_p()
        check(lib().ydorb_shard_create(matcher._h, ptr(rows), len(rows), idx_base, C.byref(self._h)))

    def close(self):
        from ._lib import lib

        if self._h:
            lib().ydorb_shard_destroy(self._h)
            self._h = None


def sharded_top2_host(comm, shard, queries, ratio, th):
    """ydorb_hamming_top2_sharded: merged YdTop2 records (global rows) + ratio-test flags, identical on every rank."""
    import ctypes as C

    from ._lib import TOP2_DTYPE, check, lib, ptr

    queries = np.ascontiguousarray(queries, dtype=np.uint8)
    out = np.zeros(len(queries), TOP2_DTYPE)
    acc = np.zeros(len(queries), np.uint8)
    check(lib().ydorb_hamming_top2_sharded(comm._h, shard._h, ptr(queries), len(queries), C.c_float(ratio), th, ptr(out), ptr(acc)))
    return out, acc


def sharded_top2_device(comm, d_queries, nq, d_db, n_rows, idx_base, ratio, th, d_out, d_accept):
    """ydorb_hamming_top2_sharded_device: raw device pointers, asynchronous on the matcher's stream."""
    import ctypes as C

    from ._lib import check, lib, ptr

    check(lib().ydorb_hamming_top2_sharded_device(comm._h, ptr(d_queries), nq, ptr(d_db), n_rows, idx_base, C.c_float(ratio), th,
                                                  ptr(d_out), ptr(d_accept)))
