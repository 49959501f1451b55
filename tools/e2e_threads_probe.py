"""e2e with T independent host threads, each driving its own extractor + matcher handle over its share of every step's
frames, K steps back to back without a per-step join (how two camera streams would drive two extractor instances)."""
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

B, W, H = 256, 640, 480
frames = bench.make_frames(B, 0)
cap = 1000
h_img = torch.from_numpy(frames).pin_memory()
h_kps = torch.zeros(B, cap, 28, dtype=torch.uint8).pin_memory()
h_desc = torch.zeros(B, cap, 32, dtype=torch.uint8).pin_memory()
h_n = torch.zeros(B, dtype=torch.int32).pin_memory()
h_top2 = torch.zeros(B, cap, 4, dtype=torch.int32).pin_memory()
h_acc = torch.zeros(B, cap, dtype=torch.uint8).pin_memory()
K = 20
for T in (1, 2, 3, 4):
    for chunks in ("1", "2", "4", ""):
        if chunks:
            os.environ["YDORB_HOST_CHUNKS"] = chunks
        else:
            os.environ.pop("YDORB_HOST_CHUNKS", None)
        Bh = B // T
        exs = [OrbExtractor(*bench.EXTRACTOR_ARGS, max_width=W, max_height=H, max_batch=Bh, device=0) for _ in range(T)]
        ms = [OrbMatcher(0.6, device=0) for _ in range(T)]
        qh = np.arange(Bh, dtype=np.int32)
        dh = ((np.arange(Bh) - 1) % Bh).astype(np.int32)

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
        print(f"T={T} chunks={chunks or 'plan'}: {dt / K * 1e3:.3f} ms per {Bh * T} frames -> {Bh * T * K / dt:.0f} frames/s")
        del exs, ms
