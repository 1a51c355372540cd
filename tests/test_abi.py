"""CPU tests of the drop-in boundary: the shared library loads, exports every symbol include/vofront.h
declares, and refuses to run without a GPU (no fallback).  No compute calls."""
import ctypes
import re
from pathlib import Path

import numpy as np
import pytest

ROOT = Path(__file__).resolve().parents[1]


def declared_symbols():
    text = (ROOT / "include" / "vofront.h").read_text()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(vo_[a-z0-9_]+)\s*\(", text)))


def test_library_exports_every_declared_symbol():
    from acs_visual_odometry_b200 import lib_path, load_library
    assert lib_path().exists(), "build libvofront.so first (python -m acs_visual_odometry_b200.build)"
    lib = ctypes.CDLL(str(lib_path()))
    names = declared_symbols()
    assert len(names) >= 25
    for n in names:
        assert hasattr(lib, n), f"{n} declared in include/vofront.h but not exported"
    bound = set(load_library()._vo_symbols)
    assert set(names) == bound, f"python binding and header disagree: {set(names) ^ bound}"


def test_no_oracle_or_fallback_in_product():
    """The product must not link, import or call anything under oracle/."""
    pkg = ROOT / "acs_visual_odometry_b200"
    for f in list(pkg.rglob("*.py")) + list(pkg.rglob("*.cu")) + list(pkg.rglob("*.cuh")) + list(pkg.rglob("*.hpp")):
        src = f.read_text()
        assert "pyoracle" not in src and "vo_oracle" not in src and "liboracle" not in src and "libvoref" not in src, f


def test_presets_and_errors_without_gpu():
    from acs_visual_odometry_b200 import VoError, VoFront, device_count, load_library
    from acs_visual_odometry_b200.api import _CParams
    lib = load_library()
    p = _CParams()
    assert lib.vo_params_preset(ctypes.byref(p), 0) == 0
    assert (p.threshold, p.nms_window, p.max_keypoints, p.match_bits) == (25000.0, 5, 1500, 32) and abs(p.ratio - 0.75) < 1e-9
    assert lib.vo_params_preset(ctypes.byref(p), 1) == 0
    assert (p.threshold, p.nms_window, p.max_keypoints) == (20000.0, 3, 2000)
    assert lib.vo_params_preset(ctypes.byref(p), 7) == -1
    if device_count() == 0:
        with pytest.raises(VoError) as e:
            VoFront(376, 1241)
        assert e.value.code == -4 and "no CPU fallback" in str(e.value)
    h = ctypes.c_void_p()
    assert lib.vo_create(ctypes.byref(h), 0, 0, 10, 10, 4) == -1          # bad geometry is rejected before any CUDA call
    assert b"max_rows" in lib.vo_last_error(None)


def test_shard_ranges_cover_all_pairs():
    from acs_visual_odometry_b200.shard import shard_frames, shard_pairs
    for n in (1, 2, 3, 10, 1000, 10001):
        for world in (1, 2, 4, 8):
            got = []
            for r in range(world):
                lo, hi = shard_pairs(n, r, world)
                got += list(range(lo, hi))
                flo, fhi = shard_frames(n, r, world)
                assert (fhi - flo) == (hi - lo + 1 if hi > lo else 0)
            assert got == list(range(max(n - 1, 0)))


def test_synth_generator_is_deterministic():
    from acs_visual_odometry_b200.synth import make_sequence
    a = make_sequence(3, 120, 160, seed=5)
    b = make_sequence(3, 120, 160, seed=5)
    assert a.dtype == np.uint8 and a.shape == (3, 120, 160) and np.array_equal(a, b)
    assert not np.array_equal(a[0], a[1])


def test_documented_batch_limits_match_the_library():
    """include/vofront.h states the slot and batch limits that api.cu enforces (VO_MAX_BATCH in vo_internal.cuh)."""
    import re
    root = Path(__file__).resolve().parents[1]
    m = re.search(r"#define\s+VO_MAX_BATCH\s+(\d+)", (root / "acs_visual_odometry_b200/csrc/vo_internal.cuh").read_text())
    limit = int(m.group(1))
    header = (root / "include/vofront.h").read_text()
    assert f"2 .. {2 * limit}" in header and f"{limit} frames" in header
    assert f"2*VO_MAX_BATCH = {2 * limit}" in (root / "acs_visual_odometry_b200/csrc/api.cu").read_text()
