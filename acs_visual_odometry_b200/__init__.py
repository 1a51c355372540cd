"""acs_visual_odometry_b200 — B200-native feature front-end (blur, Shi-Tomasi corners, FREAK, Hamming matching).

The product is `libvofront.so` (hand-written sm_100a CUDA kernels behind the C ABI in include/vofront.h).
This package is only the thin host-side binding used by the tests and the benchmark:

    from acs_visual_odometry_b200 import VoFront, Params
    vo = VoFront(max_rows=376, max_cols=1241, max_keypoints=2000, n_slots=4)
    vo.extract(gray, Params.cpu_ref(), slot=0)

There is no CPU fallback: if the library is missing it is an error, and creating a context without a
B200 raises VoError.
"""
from .api import (MATCH_FULL512, MATCH_REF32, Params, VoError, VoFront, device_count, lib_path, load_library,
                  feature_extraction_manager_with_points, matchCustomBinaryDescriptorsThreadPool)

__all__ = ["VoFront", "Params", "VoError", "MATCH_REF32", "MATCH_FULL512", "device_count", "lib_path", "load_library",
           "feature_extraction_manager_with_points", "matchCustomBinaryDescriptorsThreadPool"]
