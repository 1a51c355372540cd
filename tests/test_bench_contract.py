"""bench.py's command line and the reference arm's JSON line (the GPU arm needs a B200 and is exercised by the driver)."""
import json
import subprocess
import sys
from pathlib import Path

import pytest

ROOT = Path(__file__).resolve().parents[1]


def test_help_lists_the_contract_flags():
    out = subprocess.run([sys.executable, str(ROOT / "bench.py"), "--help"], capture_output=True, text=True, timeout=120)
    assert out.returncode == 0, out.stderr
    for flag in ("--gpus", "--steps", "--warmup", "--impl"):
        assert flag in out.stdout


@pytest.mark.timeout(600)
def test_reference_arm_prints_one_json_line():
    from oracle import pyoracle
    if not pyoracle.reference_available():
        pyoracle.build()
    out = subprocess.run([sys.executable, str(ROOT / "bench.py"), "--impl", "reference", "--steps", "1", "--warmup", "0", "--ref-frames", "3"],
                         capture_output=True, text=True, timeout=500)
    assert out.returncode == 0, out.stderr
    lines = [l for l in out.stdout.splitlines() if l.strip()]
    assert len(lines) == 1
    d = json.loads(lines[0])
    assert d["impl"] == "reference" and d["unit"] == "frames/s" and d["value"] > 0 and d["higher_is_better"] is True
    assert d["cpu_baseline"]["kind"] in ("reference", "port") and d["cpu_baseline"]["cores"] >= 1
    assert d["e2e"]["h2d_bytes_per_step"] == 0 and d["e2e"]["d2h_bytes_per_step"] == 0 and d["e2e"]["value"] == d["value"]
    assert "workload" in d["config"]


def test_reference_arm_is_silent_on_other_ranks():
    import os
    env = dict(os.environ, RANK="1", WORLD_SIZE="2", LOCAL_RANK="1")
    out = subprocess.run([sys.executable, str(ROOT / "bench.py"), "--impl", "reference", "--gpus", "2"], capture_output=True, text=True,
                         timeout=120, env=env)
    assert out.returncode == 0 and out.stdout.strip() == ""


def test_committed_stencil_counters_are_tied_to_the_code_they_were_captured_from(tmp_path, monkeypatch):
    """roofline.traffic and the issue block come from a committed ncu capture (profiles/stencil_traffic.json).  bench.py
    quotes them only for the launch shape and the stencil code of the capture: the file records a hash of stencil.cu
    that ignores comments and white space, and other code gives `traffic: null` + `traffic_stale: true`."""
    import importlib.util
    from acs_visual_odometry_b200 import build
    spec = importlib.util.spec_from_file_location("bench_mod", ROOT / "bench.py")
    bench = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(bench)
    cap = json.loads((ROOT / "profiles" / "stencil_traffic.json").read_text())
    h = cap["stencil_code_sha1"]
    assert len(h) == 40
    shape = (cap["batch"], cap["rows"], cap["cols"])
    assert bench.capture_applies(cap, *shape, h) == (True, False)
    assert bench.capture_applies(cap, *shape, "0" * 40) == (False, True)            # same shape, other code: stale
    assert bench.capture_applies(cap, 64, cap["rows"], cap["cols"], h) == (False, False)   # other shape: simply no capture
    assert bench.capture_applies({}, *shape, h) == (False, False)
    # the hash sees code, not comments or layout
    src = (build.CSRC / "stencil.cu").read_text()
    monkeypatch.setattr(build, "CSRC", tmp_path)
    (tmp_path / "stencil.cu").write_text(src)
    base = build.source_hash("stencil.cu")
    (tmp_path / "stencil.cu").write_text("// a new comment\n" + src.replace("\n", "\n  \n") + "/* trailing\n block */\n")
    assert build.source_hash("stencil.cu") == base
    (tmp_path / "stencil.cu").write_text(src.replace("constexpr int RB = 4", "constexpr int RB = 8", 1) + "\nint extra_symbol;\n")
    assert build.source_hash("stencil.cu") != base
