"""Runs a few device-resident extract_batch calls on synthetic frames of a given shape (ncu target).
usage: python tools/run_shape.py W H NFEAT BATCH [reps]"""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from ydorbslam_b200 import OrbExtractor, synth

w, h, nf, B = (int(a) for a in sys.argv[1:5])
reps = int(sys.argv[5]) if len(sys.argv) > 5 else 3
base = np.stack([synth.synth_frame(w, h, 300 + i) for i in range(min(B, 4))])
imgs = np.concatenate([base] * ((B + len(base) - 1) // len(base)))[:B]
ex = OrbExtractor(nf, 1.2, 8, 20, 7, max_width=w, max_height=h, max_batch=B)
cap = ex.capacity
dev = torch.device("cuda:0")
d_img = torch.from_numpy(imgs).to(dev)
d_kps = torch.zeros(B, cap, 28, dtype=torch.uint8, device=dev)
d_desc = torch.zeros(B, cap, 32, dtype=torch.uint8, device=dev)
d_n = torch.zeros(B, dtype=torch.int32, device=dev)
ex.set_profiling(True)
for _ in range(reps):
    ex.extract_batch_device(d_img.data_ptr(), w, h, B, d_kps.data_ptr(), d_desc.data_ptr(), cap, d_n.data_ptr())
ex.sync()
print("stage_ms", ex.stage_ms(), "kp/frame", float(d_n.float().mean().item()))
