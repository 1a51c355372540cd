"""The C++17 adapter (include/vofront_adapter.hpp): the reference's entry points, called from C++ the way
VisualOdometry.cpp:29-36 and matching.cpp:116-117,148 call them, must give the oracle's results."""
import struct
import subprocess
from pathlib import Path

import numpy as np
import pytest

ROOT = Path(__file__).resolve().parents[1]
DRIVER = ROOT / "tests" / "cpp" / "adapter_driver"


def build_driver():
    import __graft_entry__ as g
    g.build_adapter_driver()
    return DRIVER


def test_adapter_driver_compiles():
    """CPU box: the adapter header compiles as plain C++17 against cv:: types and links to the C ABI."""
    assert build_driver().exists()


def _read_features(buf, off):
    (n,) = struct.unpack_from("<i", buf, off)
    off += 4
    kp = np.frombuffer(buf, np.float32, 2 * n, off).reshape(n, 2)
    off += 8 * n
    d = np.frombuffer(buf, np.uint8, 512 * n, off).reshape(n, 512)
    return kp, d, off + 512 * n


def _read_matches(buf, off):
    (n,) = struct.unpack_from("<i", buf, off)
    m = np.frombuffer(buf, np.int32, 2 * n, off + 4).reshape(n, 2)
    return m, off + 4 + 8 * n


@pytest.mark.gpu
def test_adapter_matches_oracle(tmp_path, oracle, factory):
    f1, f2, _ = factory
    build_driver()
    raw = tmp_path / "in.raw"
    raw.write_bytes(struct.pack("<ii", *f1.shape) + f1.tobytes() + f2.tobytes())
    out = tmp_path / "out.bin"
    r = subprocess.run([str(DRIVER), str(raw), str(out)], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stderr
    buf = out.read_bytes()
    off = 0
    # 1. GPU_settings signature: VisualOdometry preset (20000 / 3 / 2000), gray input
    for img in (f1, f2):
        kp, d, off = _read_features(buf, off)
        kp_o, d_o = oracle.extract(img, 20000.0, 3, 2000)
        assert np.array_equal(kp, kp_o)
        assert (d == d_o).mean() >= 0.9999
        if img is f1:
            d1 = d
        else:
            d2 = d
    m, off = _read_matches(buf, off)
    assert np.array_equal(m, oracle.match(d1, d2, 32, 0.75))
    # 2. thread_pool signature: CPU preset (25000 / 5 / 1500), BGR input through custom_bgr2gray
    descs = []
    for img in (f1, f2):
        kp, d, off = _read_features(buf, off)
        kp_o, d_o = oracle.extract(img, 25000.0, 5, 1500)
        assert np.array_equal(kp, kp_o)
        assert (d == d_o).mean() >= 0.9999
        descs.append(d)
    m, off = _read_matches(buf, off)
    assert np.array_equal(m, oracle.match(descs[0], descs[1], 32, 0.75))
    # 3. stage entry points chained by hand = the manager
    kp, d, off = _read_features(buf, off)
    kp_o, d_o = oracle.extract(f1, 25000.0, 5, 1500)
    assert np.array_equal(kp, kp_o) and (d == d_o).mean() >= 0.9999
    # 3b. the reference's stage classes (CornerDetectionParallel / FREAK_Parallel statics), response computed in two
    # overlapping row intervals: the same keypoints and descriptors again
    # ... and the facade's own output order is the reference's: strongest first (heap of (R, i, j),
    # corner_detection_parallel.cpp:198-206; the rule is pinned to the reference's code in tests/test_oracle.py)
    from conftest import heap_order
    (n_ranked,) = struct.unpack_from("<i", buf, off)
    ranked = np.frombuffer(buf, np.float32, 2 * n_ranked, off + 4).reshape(n_ranked, 2)
    off += 4 + 8 * n_ranked
    R_o = oracle.response(oracle.blur(f1), 25000.0)
    assert n_ranked == len(kp_o) and np.array_equal(ranked, heap_order(kp_o, R_o))
    assert not np.array_equal(ranked, kp_o)
    kp, d, off = _read_features(buf, off)
    assert np.array_equal(kp, kp_o) and (d == d_o).mean() >= 0.9999
    (npx,) = struct.unpack_from("<i", buf, off)
    off += 4
    G = np.frombuffer(buf, np.float32, 3 * npx, off).reshape(3, *f1.shape)
    off += 12 * npx
    bl = oracle.blur(f1)
    jx, jy, jxy = oracle.gradients(bl)
    want = np.zeros_like(G)
    for k, a in enumerate((jx, jy, jxy)):
        want[k, 6:39, 11:f1.shape[1] - 21] = a[6:39, 11:f1.shape[1] - 21]  # interval rows [5, 40), cols [10, cols-20): interior only
    assert np.array_equal(G, want)
    (ang,) = struct.unpack_from("<f", buf, off)
    off += 4
    assert abs(ang - oracle.orientation(bl, float(kp_o[0, 0]), float(kp_o[0, 1]))) <= 2e-6
    n_desc, same1, same2 = struct.unpack_from("<iii", buf, off)
    off += 12
    assert n_desc == len(d1) and same1 == 1 and same2 == 1
    m_edit, off = _read_matches(buf, off)
    d1e = d1.copy()
    d1e[0, 0] ^= 1
    assert np.array_equal(m_edit, oracle.match(d1e, d2, 32, 0.75))
    # 4. empty input -> empty result; oversize image -> std::runtime_error
    n_empty, threw = struct.unpack_from("<ii", buf, off)
    off += 8
    assert n_empty == 0 and threw == 1
    # 5. two threads with different image sizes through the context-less signatures (one makes the process-wide context
    # grow while the other is calling in): all 12 extractions and matches equal the single-threaded results
    n_tall, bad = struct.unpack_from("<ii", buf, off)
    assert n_tall > 0 and bad == 0
