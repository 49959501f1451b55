"""Config 2 breakdown: the two extractAndCompute host calls and computeStereoMatches (mode 0 = reference quirks, mode 1 =
corrected) timed separately on a KITTI-shape synthetic stereo pair."""
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from ydorbslam_b200 import OrbExtractor, OrbMatcher, synth

w, h, nf = 1241, 376, 2000
left, right = synth.synth_stereo_pair(w, h, 5)
exl = OrbExtractor(nf, 1.2, 8, 20, 7, max_width=w, max_height=h)
exr = OrbExtractor(nf, 1.2, 8, 20, 7, max_width=w, max_height=h)
m = OrbMatcher()


def t(f, n=20):
    for _ in range(3):
        f()
    t0 = time.perf_counter()
    for _ in range(n):
        f()
    return (time.perf_counter() - t0) / n * 1e3


kl, dl = exl.extractAndCompute(left)
kr, dr = exr.extractAndCompute(right)
bf = 386.1448
print("extract left  ms", round(t(lambda: exl.extractAndCompute(left)), 3))
for mode in (0, 1):
    for max_d in (np.inf, bf / 0.5372):
        print("stereo mode", mode, "max_d", max_d, "ms",
              round(t(lambda: m.computeStereoMatches(exl, 0, exr, 0, kl, dl, kr, dr, bf, max_d, mode)), 3))
