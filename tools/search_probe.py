"""Config 5 breakdown: searchByProjectionInFrameAndMapPoint against 200k map points on a 4K frame; run with
YDORB_SEARCH_TRACE=1 for the per-phase wall times of one call."""
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
from test_gpu_search_stereo import _mps_array
from ydorbslam_b200 import OrbExtractor, OrbMatcher, synth

w, h, nf, nmp = 3840, 2160, 8000, 200_000
img = synth.synth_frame(w, h, 500, n_rect=200)
ex = OrbExtractor(nf, 1.2, 8, 20, 7, max_width=w, max_height=h)
kps, desc = ex.extractAndCompute(img)
sf = ex.getScaleFactors()
mp = synth.synth_map_points(kps, desc, nmp, w, h, 8, seed=3)
mps = _mps_array(mp)
rng = np.random.RandomState(5)
right_x = np.where(rng.rand(nf) < 0.5, kps["x"] - rng.uniform(0, 30, nf), -1.0).astype(np.float32)
occ = (rng.rand(nf) < 0.05).astype(np.uint8)
m = OrbMatcher(0.8)
fr = m.make_frame(kps, desc, right_x, bounds=(0, 0, w, h), scale_factors=sf)
f = lambda: m.searchByProjectionInFrameAndMapPoint(fr, mps, mp["desc"], mp["obs_positive"], 3.0, kp_occupied=occ)
for _ in range(3):
    f()
t0 = time.perf_counter()
for _ in range(10):
    n, match = f()
print("search ms", (time.perf_counter() - t0) / 10 * 1e3, "matches", n, "active map points", int(mps["active"].sum()), file=sys.stderr)

import torch

dev = torch.device("cuda:0")
d_mps = torch.from_numpy(mps.view(np.uint8).reshape(-1).copy()).to(dev)
d_mpd = torch.from_numpy(np.ascontiguousarray(mp["desc"]).reshape(-1).copy()).to(dev)
d_obs = torch.from_numpy(np.ascontiguousarray(mp["obs_positive"], dtype=np.uint8)).to(dev)
d_occ = torch.from_numpy(occ).to(dev)
d_out = torch.zeros(nf, dtype=torch.int32, device=dev)
g = lambda: m.searchByProjectionInFrameAndMapPoint_device(fr, d_mps.data_ptr(), d_mpd.data_ptr(), d_obs.data_ptr(), nmp,
                                                          d_out.data_ptr(), 3.0, d_occ.data_ptr())
for _ in range(3):
    g()
t0 = time.perf_counter()
for _ in range(20):
    nd = g()
print("device-resident search ms", (time.perf_counter() - t0) / 20 * 1e3, "matches", nd,
      "same", bool(np.array_equal(d_out.cpu().numpy(), match)), file=sys.stderr)
