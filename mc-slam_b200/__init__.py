"""mcslam_b200 — B200-native MC-SLAM front-end hot path.  Python side = ctypes binding of the C ABI (include/mcslam_b200.h)
plus the synthetic workload generator.  The compute lives in libmcslam_b200.so (hand-written sm_100a kernels)."""
from .binding import (CALIB_DTYPE, KP_DTYPE, Extractor, Matcher, McsError, RigFrames, Vocabulary, lib, library_path)  # noqa: F401
from . import synth  # noqa: F401
