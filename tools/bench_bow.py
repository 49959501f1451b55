"""SURVEY 8f row N1 measurement: DBoW3 transform of a batch of frames on one B200 (device-resident descriptors, CUDA
events on the matcher's stream), with the CPU oracle (port of DBoW3's transform) timed beside it on a bounded sample.

    python tools/bench_bow.py [--frames 256] [--features 1000] [--no-cpu]
"""
import argparse
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import torch  # noqa: E402

from ydorbslam_b200 import OrbMatcher, synth  # noqa: E402
from ydorbslam_b200.vocabulary import Vocabulary  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--frames", type=int, default=256)
ap.add_argument("--features", type=int, default=1000)
ap.add_argument("--steps", type=int, default=20)
ap.add_argument("--no-cpu", action="store_true")
a = ap.parse_args()

v = synth.synth_vocabulary_full(10, 6, seed=31)
m = OrbMatcher()
voc = Vocabulary(m, **v)
B, cap = a.frames, a.features
q = synth.synth_vocabulary_queries(v, B * cap, seed=33).reshape(B, cap, 32)
dev = torch.device("cuda:0")
d_desc = torch.from_numpy(q).to(dev)
d_n = torch.full((B,), cap, dtype=torch.int32, device=dev)
bi = torch.zeros((B, cap), dtype=torch.int32, device=dev)
bv = torch.zeros((B, cap), dtype=torch.float64, device=dev)
fi = torch.zeros((B, cap), dtype=torch.int32, device=dev)
fs = torch.zeros((B, cap + 1), dtype=torch.int32, device=dev)
ft = torch.zeros((B, cap), dtype=torch.int32, device=dev)
cnt = torch.zeros((B, 2), dtype=torch.int32, device=dev)
stream = torch.cuda.ExternalStream(m.stream)
torch.cuda.synchronize()


def step():
    voc.transform_batch_device(d_desc.data_ptr(), d_n.data_ptr(), cap, B, 4, bi.data_ptr(), bv.data_ptr(), fi.data_ptr(),
                               fs.data_ptr(), ft.data_ptr(), cnt.data_ptr())


for _ in range(3):
    step()
m.sync()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
with torch.cuda.stream(stream):
    e0.record()
    for _ in range(a.steps):
        step()
    e1.record()
m.sync()
ms = e0.elapsed_time(e1) / a.steps
out = {"row": "N1 DBoW3 transform", "workload": f"{B} frames x {cap} descriptors, k=10 L=6 vocabulary (1111111 nodes), levelsup 4",
       "device_ms_per_batch": ms, "descriptors_per_s": B * cap / ms * 1e3, "frames_per_s": B / ms * 1e3,
       "hamming_distances_per_descriptor": 60,
       "l2_gather_bytes_per_batch": B * cap * 60 * 32}
if not a.no_cpu:
    from oracle import orc

    ov = orc.Vocabulary.from_arrays(10, 6, 0, 0, v["child_start"], v["children"], v["descriptors"], v["weights"], v["word_ids"])
    nf = min(B, 64)
    t0 = time.perf_counter()
    for b in range(nf):
        orc.voc_transform(ov, q[b], 4)
    dt = time.perf_counter() - t0
    out["cpu_port_frames_per_s_1thread"] = nf / dt
    ref = orc.voc_transform(ov, q[0], 4)
    nb = int(cnt[0, 0])
    assert np.array_equal(bi[0, :nb].cpu().numpy(), ref["bow_ids"])
    assert np.array_equal(bv[0, :nb].cpu().numpy().view(np.uint64), ref["bow_values"].view(np.uint64))
    out["checked_vs_oracle"] = True
print(json.dumps(out))
