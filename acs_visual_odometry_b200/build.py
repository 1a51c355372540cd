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
SOURCES = ["api.cu", "stencil.cu", "nms_select.cu", "freak.cu", "match.cu"]
NVCC = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
         "-Xcompiler", "-fPIC", "-Xptxas", "-v", "--resource-usage"]


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


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose=True))
