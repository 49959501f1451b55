"""Config 1 breakdown: one 640x480 frame.  Device-resident stage times (events), the device entry timed from the host,
and the host-pointer call with pageable and with pinned buffers."""
import os
import sys
import time

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from ydorbslam_b200 import OrbExtractor, synth

w, h, nf = (int(a) for a in sys.argv[1:4]) if len(sys.argv) > 3 else (640, 480, 1000)
img = synth.synth_frame(w, h, 21)
ex = OrbExtractor(nf, 1.2, 8, 20, 7, max_width=w, max_height=h, max_batch=1)
cap = ex.capacity
dev = torch.device("cuda:0")
d_img = torch.from_numpy(img).to(dev)
d_kps = torch.zeros(1, cap, 28, dtype=torch.uint8, device=dev)
d_desc = torch.zeros(1, cap, 32, dtype=torch.uint8, device=dev)
d_n = torch.zeros(1, dtype=torch.int32, device=dev)


def dev_call():
    ex.extract_batch_device(d_img.data_ptr(), w, h, 1, d_kps.data_ptr(), d_desc.data_ptr(), cap, d_n.data_ptr())
    ex.sync()


def best(f, n=50, warm=5):
    for _ in range(warm):
        f()
    ts = []
    for _ in range(n):
        t0 = time.perf_counter()
        f()
        ts.append(time.perf_counter() - t0)
    return round(min(ts) * 1e6, 1), round(sorted(ts)[n // 2] * 1e6, 1)


print("device entry + sync, us (best, median):", best(dev_call))
ex.set_profiling(True)
for _ in range(3):
    dev_call()
print("stage_ms (serialised):", [round(v, 4) for v in ex.stage_ms()])
ex.set_profiling(False)
print("host call, pageable numpy, us:", best(lambda: ex.extractAndCompute(img)))
print("host call (batch API), us:", best(lambda: ex.extract_batch(img[None])))
