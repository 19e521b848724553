"""In-tree build of libfvp_b200.so with nvcc for sm_100a (no torch headers, no JIT cache).

    python -m faster_voxelpose_b200.build [--force]

The library lands next to the sources (faster_voxelpose_b200/libfvp_b200.so); it is git-ignored
but travels to the GPU box with the repo snapshot.
"""
from __future__ import annotations

import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
LIB = os.path.join(HERE, "libfvp_b200.so")
SOURCES = ["fvp_abi.cu", "fvp_project.cu", "fvp_project_walk.cu", "fvp_proposal.cu", "fvp_softargmax.cu", "fvp_weightnet.cu", "fvp_convaux.cu", "fvp_splat.cu", "fvp_c2c.cu", "fvp_conv7.cu", "fvp_conv7tc.cu", "fvp_res2d.cu", "fvp_restail.cu", "fvp_peer.cu", "fvp_eval.cu"]
HEADERS = ["fvp_common.cuh", "fvp_geom.cuh", "fvp_project.cuh", os.path.join("..", "..", "include", "fvp_b200.h")]
NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-lineinfo", "-std=c++17",
    # the reference's fp32 results depend on where roundings happen: never let ptxas contract
    "-fmad=false", "-prec-div=true", "-prec-sqrt=true", "-ftz=false",
    "-Xcompiler", "-fPIC", "-Xcompiler", "-O2", "-shared", "-cudart", "static",
]


def _stale() -> bool:
    if not os.path.exists(LIB):
        return True
    t = os.path.getmtime(LIB)
    deps = [os.path.join(CSRC, s) for s in SOURCES + HEADERS] + [os.path.abspath(__file__)]
    return any(os.path.getmtime(d) > t for d in deps)


def build_paranoid() -> str:
    """libfvp_b200_paranoid.so: the same sources with -DFVP_PARANOID (device-side bounds assertions, see fvp_common.cuh).
    Load it with FVP_B200_LIB=<path> (tools/paranoid_check.sh)."""
    out = os.path.join(HERE, "libfvp_b200_paranoid.so")
    nvcc = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
    subprocess.check_call([nvcc] + NVCC_FLAGS + ["-DFVP_PARANOID", "-o", out] + [os.path.join(CSRC, s) for s in SOURCES])
    return out


def build(force: bool = False, verbose: bool = False) -> str:
    if not force and not _stale():
        return LIB
    nvcc = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
    cmd = [nvcc] + NVCC_FLAGS + (["-Xptxas", "-v"] if verbose else []) + \
          ["-o", LIB] + [os.path.join(CSRC, s) for s in SOURCES]
    subprocess.check_call(cmd)
    return LIB


if __name__ == "__main__":
    if "--paranoid" in sys.argv:
        print(build_paranoid())
    else:
        print(build(force="--force" in sys.argv, verbose="-v" in sys.argv))
