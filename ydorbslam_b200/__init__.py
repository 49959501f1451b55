"""ydorbslam_b200: B200-native ORB front-end (extract / Hamming match / stereo match) behind a C ABI.

The compute path is libydorb.so (hand-written CUDA for sm_100a).  This package is the thin host-side mirror of
the reference's OrbExtractor / OrbMatcher / Frame hooks used by the tests and bench.py.
"""
from ._lib import KP_DTYPE, MPPROJ_DTYPE, TOP2_DTYPE, PinnedArray, YdorbError  # noqa: F401
from .extractor import OrbExtractor  # noqa: F401
from .matcher import OrbMatcher  # noqa: F401
