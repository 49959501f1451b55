"""e2e with T host threads, each driving its own extractor + matcher handle over ITS OWN batch of Bh frames per call
(not a 1/T share of a 256-frame step): how the host-call throughput depends on frames per call and on the thread count."""
import ctypes as C
import os
import sys
import threading
import time

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench
from ydorbslam_b200 import OrbExtractor, OrbMatcher
from ydorbslam_b200._lib import check, lib, ptr

W, H, cap = 640, 480, 1000
BMAX = 1024
frames = bench.make_frames(256, 0)
frames = np.concatenate([frames] * (BMAX // 256))
h_img = torch.from_numpy(frames).pin_memory()
h_kps = torch.zeros(BMAX, cap, 28, dtype=torch.uint8).pin_memory()
h_desc = torch.zeros(BMAX, cap, 32, dtype=torch.uint8).pin_memory()
h_n = torch.zeros(BMAX, dtype=torch.int32).pin_memory()
h_top2 = torch.zeros(BMAX, cap, 4, dtype=torch.int32).pin_memory()
h_acc = torch.zeros(BMAX, cap, dtype=torch.uint8).pin_memory()
for T, Bh in [(4, 64), (2, 128), (4, 128), (3, 128), (2, 256), (3, 256), (4, 256), (6, 128), (8, 64), (8, 128)]:
    exs = [OrbExtractor(*bench.EXTRACTOR_ARGS, max_width=W, max_height=H, max_batch=Bh, device=0) for _ in range(T)]
    ms = [OrbMatcher(0.6, device=0) for _ in range(T)]
    qh = np.arange(Bh, dtype=np.int32)
    dh = ((np.arange(Bh) - 1) % Bh).astype(np.int32)
    K = max(4, 20 * 256 // (T * Bh))

    def work(t, steps):
        o = t * Bh
        for _ in range(steps):
            check(lib().ydorb_extract_batch(exs[t]._h, C.c_void_p(h_img[o:].data_ptr()), W, H, W, C.c_size_t(W * H), Bh,
                                            C.c_void_p(h_kps[o:].data_ptr()), C.c_void_p(h_desc[o:].data_ptr()), cap,
                                            C.c_void_p(h_n[o:].data_ptr())))
            check(lib().ydorb_match_frames(ms[t]._h, exs[t]._h, ptr(qh), ptr(dh), Bh, C.c_float(0.6), 50,
                                           C.c_void_p(h_top2[o:].data_ptr()), C.c_void_p(h_acc[o:].data_ptr())))

    def run(steps):
        th = [threading.Thread(target=work, args=(t, steps)) for t in range(T)]
        t0 = time.perf_counter()
        for x in th:
            x.start()
        for x in th:
            x.join()
        return time.perf_counter() - t0

    run(3)
    dt = run(K)
    print(f"T={T} Bh={Bh}: {Bh * T * K / dt:.0f} frames/s", flush=True)
    del exs, ms
