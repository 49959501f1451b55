"""Host-side timing of the host-pointer entry points (pinned buffers): extract_batch alone, match_frames alone,
and the raw H2D / D2H copies of the same bytes, to see what bounds the e2e figure."""
import ctypes as C
import os
import sys
import time

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench
from ydorbslam_b200 import OrbExtractor, OrbMatcher
from ydorbslam_b200._lib import check, lib, ptr

B, W, H = 256, 640, 480
frames = bench.make_frames(B, 0)
ex = OrbExtractor(*bench.EXTRACTOR_ARGS, max_width=W, max_height=H, max_batch=B, device=0)
m = OrbMatcher(0.6, device=0)
cap = ex.capacity
h_img = torch.from_numpy(frames).pin_memory()
h_kps = torch.zeros(B, cap, 28, dtype=torch.uint8).pin_memory()
h_desc = torch.zeros(B, cap, 32, dtype=torch.uint8).pin_memory()
h_n = torch.zeros(B, dtype=torch.int32).pin_memory()
h_top2 = torch.zeros(B, cap, 4, dtype=torch.int32).pin_memory()
h_acc = torch.zeros(B, cap, dtype=torch.uint8).pin_memory()
qf = np.arange(B, dtype=np.int32)
df = ((np.arange(B) - 1) % B).astype(np.int32)


def ext():
    check(lib().ydorb_extract_batch(ex._h, C.c_void_p(h_img.data_ptr()), W, H, W, C.c_size_t(W * H), B, C.c_void_p(h_kps.data_ptr()),
                                    C.c_void_p(h_desc.data_ptr()), cap, C.c_void_p(h_n.data_ptr())))


def mat():
    check(lib().ydorb_match_frames(m._h, ex._h, ptr(qf), ptr(df), B, C.c_float(0.6), 50, C.c_void_p(h_top2.data_ptr()),
                                   C.c_void_p(h_acc.data_ptr())))


def timeit(f, n=10):
    for _ in range(3):
        f()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(n):
        f()
    torch.cuda.synchronize()
    return (time.perf_counter() - t0) / n * 1e3


d_img = torch.empty_like(h_img, device="cuda")
d_out = torch.zeros(B, cap, 60, dtype=torch.uint8, device="cuda")
h_out = torch.zeros(B, cap, 60, dtype=torch.uint8).pin_memory()
print("extract_batch  ms", timeit(ext))
print("match_frames   ms", timeit(mat))
print("both           ms", timeit(lambda: (ext(), mat())))
print("H2D 78.6MB     ms", timeit(lambda: d_img.copy_(h_img, non_blocking=True)))
print("D2H 15.4MB     ms", timeit(lambda: h_out.copy_(d_out, non_blocking=True)))

# ---- two handles driven from two host threads (the reference runs its left and right extractor concurrently on two
# instances, src/frame.cpp:84-87): each thread owns half of the step's frames; one pipeline's fill/drain overlaps the
# other's kernels
from concurrent.futures import ThreadPoolExecutor

for T in (2, 4):
    Bh = B // T
    exs = [OrbExtractor(*bench.EXTRACTOR_ARGS, max_width=W, max_height=H, max_batch=Bh, device=0) for _ in range(T)]
    ms = [OrbMatcher(0.6, device=0) for _ in range(T)]
    qh = np.arange(Bh, dtype=np.int32)
    dh = ((np.arange(Bh) - 1) % Bh).astype(np.int32)

    def half(t):
        o = t * Bh
        check(lib().ydorb_extract_batch(exs[t]._h, C.c_void_p(h_img[o:].data_ptr()), W, H, W, C.c_size_t(W * H), Bh,
                                        C.c_void_p(h_kps[o:].data_ptr()), C.c_void_p(h_desc[o:].data_ptr()), cap,
                                        C.c_void_p(h_n[o:].data_ptr())))
        check(lib().ydorb_match_frames(ms[t]._h, exs[t]._h, ptr(qh), ptr(dh), Bh, C.c_float(0.6), 50,
                                       C.c_void_p(h_top2[o:].data_ptr()), C.c_void_p(h_acc[o:].data_ptr())))

    pool = ThreadPoolExecutor(T)
    print(f"{T} handles / {T} threads, both  ms", timeit(lambda: list(pool.map(half, range(T)))))
    pool.shutdown()
    del exs, ms
