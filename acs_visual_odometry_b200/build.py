"""Build libvofront.so (hand-written sm_100a kernels + the C ABI) in-tree with nvcc.

    python -m acs_visual_odometry_b200.build [--force]

nvcc cross-compiles without a GPU.  Objects are rebuilt only when a source or header is newer.
"""
import os
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor
from pathlib import Path

PKG = Path(__file__).resolve().parent
CSRC = PKG / "csrc"
OBJ = CSRC / "_obj"
LIB = PKG / "libvofront.so"
SOURCES = ["api.cu", "stencil.cu", "nms_select.cu", "freak.cu", "match.cu", "match_static.cu", "stages.cu"]
NVCC = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
         "-Xcompiler", "-fPIC", "-Xptxas", "-v", "--resource-usage"] + os.environ.get("VO_NVCC_EXTRA", "").split()


def _newest_header():
    hs = list(CSRC.glob("*.cuh")) + list(CSRC.glob("*.h")) + [PKG.parent / "include" / "vofront.h"]
    return max(h.stat().st_mtime for h in hs)


def _compile(src, force, verbose):
    obj = OBJ / (src[:-3] + ".o")
    s = CSRC / src
    if not force and obj.exists() and obj.stat().st_mtime > max(s.stat().st_mtime, _newest_header()):
        return obj, ""
    cmd = [NVCC, *FLAGS, "-c", str(s), "-o", str(obj)]
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode:
        raise RuntimeError(f"nvcc failed for {src}:\n{r.stdout}\n{r.stderr}")
    return obj, r.stderr if verbose else ""


def build(force=False, verbose=False):
    OBJ.mkdir(exist_ok=True)
    with ThreadPoolExecutor(len(SOURCES)) as ex:
        res = list(ex.map(lambda s: _compile(s, force, verbose), SOURCES))
    objs = [str(o) for o, _ in res]
    if verbose:
        for _, log in res:
            if log:
                print(log)
    if force or not LIB.exists() or any(Path(o).stat().st_mtime > LIB.stat().st_mtime for o in objs):
        cmd = [NVCC, "-shared", "-gencode", "arch=compute_100a,code=sm_100a", "-o", str(LIB), *objs]
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode:
            raise RuntimeError(f"link failed:\n{r.stdout}\n{r.stderr}")
    return LIB


def source_hash(name="stencil.cu"):
    """SHA-1 of a kernel source with comments and white space removed: changes when the code changes, not when a comment
    does.  profiles/stencil_traffic.json records it for the source its ncu counters were captured from; bench.py refuses
    to quote those counters for any other code (`traffic: null, traffic_stale: true`)."""
    import hashlib
    import re
    text = (CSRC / name).read_text()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    text = re.sub(r"//[^\n]*", "", text)
    return hashlib.sha1(re.sub(r"\s+", "", text).encode()).hexdigest()


HOST = PKG / "host"
HOST_LIB = PKG / "libvohost.so"
HOST_MAIN = PKG / "vo_main"
HOST_SOURCES = ["vo_ransac.cpp", "vo_pose.cpp", "vo_image_io.cpp", "vo_pipeline.cpp", "vo_host_abi.cpp"]
CXX = os.environ.get("CXX", "g++")
CXXFLAGS = ["-std=c++17", "-O3", "-fPIC", "-ffp-contract=off", "-Wall", "-pthread"]


def build_host(force=False):
    """libvohost.so (RANSAC, pose, formats, pipelined VisualOdometry driver; C ABI of include/vohost.h) and the
    vo_main command line, both linked against libvofront.so."""
    build(force=False)
    srcs = [HOST / s for s in HOST_SOURCES]
    deps = srcs + list(HOST.glob("*.hpp")) + [PKG.parent / "include" / "vohost.h", PKG.parent / "include" / "vofront.h", LIB]
    newest = max(d.stat().st_mtime for d in deps)
    rpath = "-Wl,-rpath,$ORIGIN"
    if force or not HOST_LIB.exists() or HOST_LIB.stat().st_mtime < newest:
        cmd = [CXX, *CXXFLAGS, "-shared", "-o", str(HOST_LIB), *map(str, srcs), f"-L{PKG}", "-lvofront", "-lz", rpath]
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode:
            raise RuntimeError(f"host library build failed:\n{r.stdout}\n{r.stderr}")
    main_src = HOST / "vo_main.cpp"
    if force or not HOST_MAIN.exists() or HOST_MAIN.stat().st_mtime < max(newest, main_src.stat().st_mtime, HOST_LIB.stat().st_mtime):
        cmd = [CXX, *CXXFLAGS, "-o", str(HOST_MAIN), str(main_src), f"-L{PKG}", "-lvohost", "-lvofront", rpath]
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode:
            raise RuntimeError(f"vo_main build failed:\n{r.stdout}\n{r.stderr}")
    return HOST_LIB


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose=True))
    print(build_host(force="--force" in sys.argv))
