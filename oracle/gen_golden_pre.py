"""Writes tests/golden/preprocess.npz: cv2-generated vectors for the SURVEY 8f N4 rows (cvtColor *2GRAY and
undistortPoints).  Run in a container that has cv2 (4.13.0 here):  python oracle/gen_golden_pre.py"""
import pathlib
import sys

import cv2
import numpy as np

ROOT = pathlib.Path(__file__).resolve().parents[1]
sys.path.insert(0, str(ROOT))
from tests import pre_cases  # noqa: E402

out = {"cv2_version": cv2.__version__}
img3 = pre_cases.color_image(61, 23, 3, 1)
img4 = pre_cases.color_image(61, 23, 4, 2)
out["gray_rgb"] = cv2.cvtColor(img3, cv2.COLOR_RGB2GRAY)
out["gray_bgr"] = cv2.cvtColor(img3, cv2.COLOR_BGR2GRAY)
out["gray_rgba"] = cv2.cvtColor(img4, cv2.COLOR_RGBA2GRAY)
out["gray_bgra"] = cv2.cvtColor(img4, cv2.COLOR_BGRA2GRAY)
for name, dist in pre_cases.DISTS.items():
    pts = pre_cases.points(400, 3, spread=6.0 if name == "barrel" else 1.2)
    out["und_" + name] = cv2.undistortPoints(pts.reshape(-1, 1, 2), pre_cases.K, dist, None, pre_cases.K).reshape(-1, 2)
    out["undP_" + name] = cv2.undistortPoints(pts.reshape(-1, 1, 2), pre_cases.K, dist, None, pre_cases.P_OTHER).reshape(-1, 2)
np.savez_compressed(ROOT / "tests" / "golden" / "preprocess.npz", **out)
print("wrote", ROOT / "tests" / "golden" / "preprocess.npz")
