"""Shared helpers for the parity tests: golden loading and seeded inputs."""
import functools
import os

import numpy as np

from faster_voxelpose_b200 import synthetic as syn
from faster_voxelpose_b200.config import make_cfg

GOLDEN_DIR = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
RIGS = ("panoptic", "shelf", "campus")

# gate from BASELINE.json north_star: cubes and planes within 1e-5 relative (fp32).  Values live in
# [0, 1], so the operational form is |a-b| <= ATOL + RTOL*|b| with both at 1e-5.
CUBE_ATOL = 1e-5
CUBE_RTOL = 1e-5
JOINT_MM = 0.1


@functools.lru_cache(maxsize=None)
def golden(rig):
    return dict(np.load(os.path.join(GOLDEN_DIR, f"golden_{rig}.npz")))


@functools.lru_cache(maxsize=None)
def rig_setup(rig, min_score=-1.0):
    cfg = make_cfg(rig, device="cpu", CAPTURE_SPEC__MIN_SCORE=min_score)
    cams = syn.load_rig(rig)
    g = golden(rig)
    return cfg, cams, g


@functools.lru_cache(maxsize=None)
def heatmaps(rig, fam, B=2):
    cfg, cams, _ = rig_setup(rig)
    W, H = cfg.DATASET.HEATMAP_SIZE
    if fam == "N":
        return syn.noise_heatmaps(7, B, len(cams), cfg.DATASET.NUM_JOINTS, H, W)
    return syn.gaussian_heatmaps(11, cfg, cams, B)


def close(a, b, atol=CUBE_ATOL, rtol=CUBE_RTOL):
    a = np.asarray(a, dtype=np.float64)
    b = np.asarray(b, dtype=np.float64)
    return bool(np.all(np.abs(a - b) <= atol + rtol * np.abs(b)))


def report(a, b):
    a = np.asarray(a)
    b = np.asarray(b)
    return "max|d|=%.3g, bit-identical %.4f%%" % (float(np.abs(a.astype(np.float64) - b).max()),
                                                  100.0 * float((a == b).mean()))
