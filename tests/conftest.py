import sys
from pathlib import Path

import numpy as np
import pytest

ROOT = Path(__file__).resolve().parents[1]
if str(ROOT) not in sys.path:
    sys.path.insert(0, str(ROOT))
GOLDEN = ROOT / "tests" / "golden"


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a B200 (run with -m gpu on the GPU box)")


@pytest.fixture(scope="session")
def oracle():
    from oracle.pyoracle import Oracle
    return Oracle()


@pytest.fixture(scope="session")
def reference():
    """The reference's own code (oracle/_ref/libvoref.so); skipped when it was never built."""
    from oracle import pyoracle
    if not pyoracle.reference_available():
        pytest.skip("oracle/_ref/libvoref.so not available")
    return pyoracle.Reference()


@pytest.fixture(scope="session")
def reference_serial():
    """The reference's own full-length matcher (oracle/_ref/libvoref_serial.so); skipped when it was never built."""
    from oracle import pyoracle
    if not pyoracle.reference_serial_available():
        pytest.skip("oracle/_ref/libvoref_serial.so not available")
    return pyoracle.ReferenceSerial()


@pytest.fixture(scope="session")
def factory():
    import cv2
    f1 = cv2.imread(str(GOLDEN / "factory1.png"), cv2.IMREAD_GRAYSCALE)
    f2 = cv2.imread(str(GOLDEN / "factory2.png"), cv2.IMREAD_GRAYSCALE)
    assert f1 is not None and f1.shape == (480, 752)
    return f1, f2, dict(np.load(GOLDEN / "factory.npz"))


@pytest.fixture(scope="session")
def kitti_pair():
    """Two KITTI-shaped synthetic frames (synth.make_sequence, seed 7) and the golden vectors the reference's own code
    produced on them.  The frames are the committed PNGs (tests/golden/kitti{0,1}.png), so the golden comparison never
    depends on numpy's random stream; a changed generator stream or a corrupt PNG FAILS loudly instead of dropping it."""
    import hashlib
    import cv2
    from acs_visual_odometry_b200.synth import make_sequence
    g = dict(np.load(GOLDEN / "kitti_synth.npz"))
    fr = np.stack([cv2.imread(str(GOLDEN / f"kitti{k}.png"), cv2.IMREAD_GRAYSCALE) for k in range(2)])
    assert fr.shape == (2, 376, 1241) and fr.dtype == np.uint8
    assert hashlib.sha256(fr.tobytes()).digest() == g["frames_sha256"].tobytes(), "tests/golden/kitti*.png do not match kitti_synth.npz"
    assert np.array_equal(make_sequence(2, seed=7), fr), "synth.make_sequence(seed=7) no longer reproduces the committed frames"
    return fr, g


@pytest.fixture(scope="session")
def matcher_kat():
    return dict(np.load(GOLDEN / "matcher_kat.npz"))


def unpack(packed):
    """n x 16 uint32 -> n x 512 bytes (0/1)."""
    packed = np.ascontiguousarray(packed, np.uint32).reshape(-1, 16)
    bits = np.unpackbits(packed.view(np.uint8).reshape(-1, 64), axis=1, bitorder="little")
    return np.ascontiguousarray(bits, np.uint8)


def heap_order(kp_xy, R):
    """Keypoints (x, y) in the order CornerDetectionParallel::non_maximum_suppression returns them: it pops a max-heap of
    (R, i, j) tuples (corner_detection_parallel.cpp:167,198-206) = descending response, then row, then column."""
    kp = np.asarray(kp_xy, np.float32).reshape(-1, 2)
    x, y = kp[:, 0].astype(np.int64), kp[:, 1].astype(np.int64)
    order = np.lexsort((-x, -y, -R[y, x].astype(np.float64)))
    return kp[order]


def dense_R(idx, val, shape):
    R = np.zeros(shape, np.float32)
    R.reshape(-1)[idx] = val
    return R


@pytest.fixture(scope="session")
def vo_small():
    """A context for the small GPU parity cases."""
    from acs_visual_odometry_b200 import VoFront
    ctx = VoFront(max_rows=480, max_cols=1241, max_keypoints=2048, n_slots=8)
    yield ctx
    ctx.close()


@pytest.fixture(scope="session")
def vo_small_unfused():
    """The same context with VO_FORCE_UNFUSED=1: extraction writes the response map and runs the stand-alone k_nms
    (the round trip the fused stencil kernel removed); both paths must give identical results."""
    import os
    from acs_visual_odometry_b200 import VoFront
    old = os.environ.get("VO_FORCE_UNFUSED")
    os.environ["VO_FORCE_UNFUSED"] = "1"          # read once, by vo_create
    try:
        ctx = VoFront(max_rows=480, max_cols=1241, max_keypoints=2048, n_slots=8)
    finally:
        if old is None:
            del os.environ["VO_FORCE_UNFUSED"]
        else:
            os.environ["VO_FORCE_UNFUSED"] = old
    yield ctx
    ctx.close()
