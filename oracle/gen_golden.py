#!/usr/bin/env python3
"""Generate tests/golden/*.npz from REAL OpenCV primitives (cv2) driven by oracle/cv2_flow.py.

Runs only in the build container (needs cv2 and, for the one real image, /root/reference).  The fixtures it
writes are committed; tests compare the C++ oracle and the CUDA path against them on any box.

    python oracle/gen_golden.py
"""
import hashlib
import pathlib
import sys

import numpy as np

ROOT = pathlib.Path(__file__).resolve().parents[1]
sys.path.insert(0, str(ROOT))

from oracle import cv2_flow  # noqa: E402
from ydorbslam_b200 import synth  # noqa: E402

OUT = ROOT / "tests" / "golden"
KP_DTYPE = np.dtype(
    [("x", "<f4"), ("y", "<f4"), ("size", "<f4"), ("angle", "<f4"), ("response", "<f4"), ("octave", "<i4"), ("class_id", "<i4")]
)


def sha(a):
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


def case(name, img, nf, store_input):
    import cv2

    stages = []
    kps, desc = cv2_flow.extract(img, nf, 1.2, 8, 20, stages=stages)
    d = dict(
        n_features=np.int32(nf),
        input_sha=np.array(sha(img)),
        shape=np.array(img.shape, np.int32),
        kps=np.array(kps, dtype=KP_DTYPE),
        desc=desc,
        cv2_version=np.array(cv2.__version__),
    )
    if store_input:
        d["input"] = img
    for st in stages:
        l = st["level"]
        d[f"pyr_sha_{l}"] = np.array(sha(st["padded"]))
        d[f"cand_{l}"] = np.array(st["cands"], dtype=np.float32).reshape(-1, 3)
        d[f"sel_{l}"] = np.array(st["sel"], dtype=np.int32)
        if "blurred" in st:
            d[f"blur_sha_{l}"] = np.array(sha(st["blurred"]))
    np.savez_compressed(OUT / f"extract_{name}.npz", **d)
    print(name, "kps", len(kps), "cands", [len(s["cands"]) for s in stages])


def primitives():
    import cv2

    rng = np.random.RandomState(7)
    ys = np.concatenate([rng.randint(-200000, 200001, 20000), np.repeat(np.arange(-40, 41), 81)]).astype(np.float32)
    xs = np.concatenate([rng.randint(-200000, 200001, 20000), np.tile(np.arange(-40, 41), 81)]).astype(np.float32)
    ang = np.array([cv2.fastAtan2(float(y), float(x)) for y, x in zip(ys, xs)], dtype=np.float32)
    img = synth.synth_frame(321, 243, 11)
    rs = cv2.resize(img, (268, 203), interpolation=cv2.INTER_LINEAR)
    rs2 = cv2.resize(rs, (223, 169), interpolation=cv2.INTER_LINEAR)
    bl = cv2.GaussianBlur(img.copy(), (7, 7), 2, sigmaY=2, borderType=cv2.BORDER_REFLECT_101)
    det = cv2.FastFeatureDetector_create(20, True, cv2.FAST_FEATURE_DETECTOR_TYPE_9_16)
    fk = det.detect(img)
    fast = np.array([(k.pt[0], k.pt[1], k.response) for k in fk], dtype=np.float32).reshape(-1, 3)
    det7 = cv2.FastFeatureDetector_create(7, True, cv2.FAST_FEATURE_DETECTOR_TYPE_9_16)
    roi = np.ascontiguousarray(img[40:77, 100:137])
    fk7 = det7.detect(roi)
    fast7 = np.array([(k.pt[0], k.pt[1], k.response) for k in fk7], dtype=np.float32).reshape(-1, 3)
    np.savez_compressed(
        OUT / "primitives.npz", atan_y=ys, atan_x=xs, atan_deg=ang, resize1=rs, resize2=rs2, blur=bl, fast20=fast,
        fast7_roi=fast7, cv2_version=np.array(cv2.__version__),
    )
    print("primitives: atan", len(ang), "fast", len(fast), len(fast7))


def main():
    import cv2

    OUT.mkdir(parents=True, exist_ok=True)
    primitives()
    ref_img = pathlib.Path("/root/reference/thirdParty/DBow3/utils/images/image0.png")
    if ref_img.exists():
        case("image0", cv2.imread(str(ref_img), cv2.IMREAD_GRAYSCALE), 1000, True)
    case("synth640", synth.synth_frame(640, 480, 1), 1000, False)
    case("sparse640", synth.synth_sparse_frame(640, 480, 2), 1000, False)
    case("synth1241", synth.synth_frame(1241, 376, 3), 2000, False)
    case("tiny97", synth.synth_frame(97, 83, 4), 200, False)


if __name__ == "__main__":
    main()
