import hashlib

import numpy as np

from ydorbslam_b200 import synth

SEEDS = {"synth640": 1, "sparse640": 2, "synth1241": 3, "tiny97": 4}


def sha(a):
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


def golden_input(name, g):
    if "input" in g.files:
        return g["input"]
    h, w = (int(v) for v in g["shape"])
    seed = SEEDS[name]
    img = synth.synth_sparse_frame(w, h, seed) if name.startswith("sparse") else synth.synth_frame(w, h, seed)
    assert sha(img) == str(g["input_sha"]), "synthetic generator drifted from the golden fixtures"
    return img


def has_cuda():
    try:
        import ctypes

        from ydorbslam_b200._lib import lib

        return lib().ydorb_device_count() > 0
    except Exception:
        return False
