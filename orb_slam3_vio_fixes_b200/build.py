"""Builds liborbx.so (the C-ABI CUDA library) in-tree with nvcc for sm_100a."""
import os
import shlex
import subprocess
import sys
from pathlib import Path

PKG = Path(__file__).resolve().parent
CSRC = PKG / "csrc"
SO = PKG / "liborbx.so"
SOURCES = ["api.cu", "pyramid.cu", "fast.cu", "fast_cell.cu", "octree.cu", "describe.cu", "blur.cu", "match.cu", "stereo.cu", "search.cu", "bow.cu", "frame.cu", "voc.cu"]
NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
    # parity contract: no implicit FMA contraction anywhere (explicit __fmaf_rn where wanted)
    "--fmad=false", "-shared", "-Xcompiler", "-fPIC", "-Xcompiler", "-ffp-contract=off",
    "-cudart", "static",
]


def needs_build():
    if not SO.exists():
        return True
    t = SO.stat().st_mtime
    deps = list(CSRC.glob("*")) + [PKG.parent / "include" / "orbx.h", Path(__file__)]
    return any(d.stat().st_mtime > t for d in deps)


def build(force=False, verbose=False):
    if not force and not needs_build():
        return SO
    extra = shlex.split(os.environ.get("ORBX_NVCC_EXTRA", ""))  # experiments: -DFS_WARPS=1 ...
    cmd = ["nvcc"] + NVCC_FLAGS + extra + (["-Xptxas", "-v"] if verbose else []) + \
        ["-o", str(SO)] + [str(CSRC / s) for s in SOURCES if (CSRC / s).exists()]
    subprocess.check_call(cmd)
    return SO


if __name__ == "__main__":
    build(force=True, verbose="-v" in sys.argv)
    print(SO)
