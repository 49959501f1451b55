"""Diagnostic: run the CTA-per-cell-block FAST kernel on a few launch shapes in fresh processes."""
import os, subprocess, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SCRIPT = r"""
import sys, numpy as np
from ydorbslam_b200 import OrbExtractor, synth
w, h, b, nf, pre = (int(v) for v in sys.argv[1:6])
ex = OrbExtractor(nf, 1.2, 8, 20, 7, max_width=w, max_height=h, max_batch=b)
imgs = np.stack([synth.synth_frame(w, h, 90 + i) for i in range(b)])
if pre:
    ex.extractAndCompute(imgs[0])
kps, desc, n = ex.extract_batch(imgs)
print("ok", n.tolist()[:4])
"""
for args in [(640, 480, 5, 1000, 0), (640, 480, 5, 1000, 1), (1241, 376, 2, 2000, 0), (1241, 376, 2, 2000, 1), (640, 480, 2, 1000, 0),
             (640, 480, 16, 1000, 0), (640, 480, 64, 1000, 0)]:
    e = dict(os.environ, PYTHONPATH=ROOT, YDORB_FAST_KERNEL="blocks", CUDA_LAUNCH_BLOCKING="1")
    out = subprocess.run([sys.executable, "-c", SCRIPT, *map(str, args)], env=e, cwd=ROOT, capture_output=True, text=True)
    print(args, out.returncode, out.stdout.strip()[-200:], out.stderr.strip()[-300:].replace("\n", " | "))
