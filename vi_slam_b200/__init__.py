"""vi_slam_b200 — B200-native ORB front-end behind vi_slam's extractor / matcher interface.

The compute lives in libviorb.so (hand-written sm_100a CUDA behind the C ABI of include/viorb.h);
this package is the thin host-side mirror of the reference's interface for that path.
"""
from .extractor import KP_DTYPE, MATCH_DTYPE, FExtractor, ViorbError, pinned_empty  # noqa: F401
from . import bow, frame, matcher, synthetic  # noqa: F401
