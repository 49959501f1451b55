"""The OpenCV-typed drop-in (include/ydorb_cv.hpp: the reference's own signatures over libydorb.so) driven next to the
REFERENCE's own classes on the same inputs: oracle/_ref/cv_shim_check is built where /root/reference exists
(oracle/refshim/Makefile: reference translation units unmodified + the OpenCV stand-in) and travels to the GPU box."""
import os
import pathlib
import subprocess

import numpy as np
import pytest

from ydorbslam_b200 import synth

pytestmark = pytest.mark.gpu
BIN = pathlib.Path(__file__).resolve().parents[1] / "oracle" / "_ref" / "cv_shim_check"


@pytest.mark.skipif(not BIN.exists(), reason="oracle/_ref/cv_shim_check not shipped")
@pytest.mark.parametrize("w,h,nf,seed", [(1241, 376, 2000, 64), (640, 480, 1000, 66)])
def test_cv_typed_shim_equals_reference_classes(tmp_path, w, h, nf, seed):
    left, right = synth.synth_stereo_pair(w, h, seed)
    (tmp_path / "l.raw").write_bytes(np.ascontiguousarray(left).tobytes())
    (tmp_path / "r.raw").write_bytes(np.ascontiguousarray(right).tobytes())
    out = subprocess.run([str(BIN), str(tmp_path / "l.raw"), str(tmp_path / "r.raw"), str(w), str(h), str(nf)],
                         capture_output=True, text=True, timeout=600)
    assert out.returncode == 0, out.stdout[-2000:] + out.stderr[-2000:]
    for tag in ("OK E1", "OK E2", "OK S1", "OK G2", "OK M0", "OK M1", "ALL OK"):
        assert tag in out.stdout, out.stdout
