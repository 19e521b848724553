"""Recipe for ``oracle/_ref``: the reference's own modules for this path, UNMODIFIED, made available to the timed
reference arm of bench.py (and to tests) on the GPU box, where ``/root/reference`` does not exist.

    python oracle/make_ref.py            # no-op when /root/reference is absent

Copies the files the eval forward imports (lib/models/*.py, lib/core/proposal.py, lib/utils/{cameras,transforms}.py
and the package __init__ files) byte for byte from where they lie under /root/reference into oracle/_ref/lib.
oracle/_ref is git-ignored (nothing of the reference enters the history) but not gpurun-ignored, so it travels to the
GPU box like the built libfvp_b200.so does.  TEST / BENCH INFRASTRUCTURE ONLY: nothing under faster_voxelpose_b200/
imports it.  A manifest with the sha256 of every copied file is written next to them.
"""
from __future__ import annotations

import hashlib
import json
import os
import shutil

HERE = os.path.dirname(os.path.abspath(__file__))
SRC = "/root/reference/lib"
DST = os.path.join(HERE, "_ref", "lib")
FILES = [
    "models/__init__.py", "models/faster_voxelpose.py", "models/human_detection_net.py",
    "models/joint_localization_net.py", "models/project_whole.py", "models/project_individual.py",
    "models/cnns_2d.py", "models/cnns_1d.py", "models/weight_net.py", "models/resnet.py",
    "core/__init__.py", "core/proposal.py", "utils/__init__.py", "utils/cameras.py", "utils/transforms.py",
]


def make(verbose: bool = False) -> bool:
    if not os.path.isdir(SRC):
        return os.path.isdir(DST)
    manifest = {}
    for rel in FILES:
        src, dst = os.path.join(SRC, rel), os.path.join(DST, rel)
        os.makedirs(os.path.dirname(dst), exist_ok=True)
        shutil.copyfile(src, dst)
        with open(dst, "rb") as fh:
            manifest[rel] = hashlib.sha256(fh.read()).hexdigest()
    with open(os.path.join(DST, "MANIFEST.json"), "w") as fh:
        json.dump({"source": SRC, "files": manifest}, fh, indent=1)
    if verbose:
        print(f"oracle/_ref: {len(FILES)} reference files copied from {SRC}")
    return True


if __name__ == "__main__":
    make(verbose=True)
