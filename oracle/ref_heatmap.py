"""TEST INFRASTRUCTURE -- numpy restatement of the reference's input-heatmap generator (oracle).

Follows ``lib/dataset/JointsDataset.py:368-447`` (``generate_input_heatmap``, data augmentation
off) and ``:265-279`` (``compute_human_scale``) of the reference, written per pixel instead of with
array slices so that it doubles as the specification of the CUDA kernel
(``faster_voxelpose_b200/csrc/fvp_splat.cu``).  Only ``tests/`` and ``__graft_entry__.smoke()`` may
import it.

Parity status: PINNED by ``tests/golden/golden_splat.npz`` -- outputs of the reference's own method,
called unbound on a stub object in the authoring container (``tests/golden/make_golden_splat.py``,
numpy 2.3: the float32 ``arange`` meets ``np.float64`` scalars, so the Gaussian is evaluated in
float64 and rounded to float32 on store).
"""
from __future__ import annotations

import math

import numpy as np


def human_scale(pose_hm: np.ndarray) -> float:
    """JointsDataset.py:265-279 with joints_vis = ones: clipped squared extent of the pose."""
    ext = max(pose_hm[:, 1].max() - pose_hm[:, 1].min(), pose_hm[:, 0].max() - pose_hm[:, 0].min())
    return float(np.clip(ext ** 2, 1.0 / 4 * 96 ** 2, 4 * 96 ** 2))


def generate_input_heatmap(joints, joints_vis, sigma, image_size, heatmap_size) -> np.ndarray:
    """joints: list of [J,2] arrays (network-input pixels); joints_vis: list of [J] or None.
    Returns [J, H, W] float32.  (JointsDataset.py:368-447)"""
    W, H = int(heatmap_size[0]), int(heatmap_size[1])
    fs = np.asarray(image_size, np.float64) / np.asarray(heatmap_size, np.float64)
    J = joints[0].shape[0] if len(joints) else 0
    target = np.zeros((J, H, W), np.float32)
    for n in range(len(joints)):
        pose = np.asarray(joints[n], np.float64)[:, :2]
        hs = 2 * human_scale(pose / fs)                               # :384-385
        cur_sigma = sigma * math.sqrt(hs / (96.0 * 96.0))             # :389
        tmp = cur_sigma * 3                                           # :390
        size = 2 * tmp + 1                                            # :406
        length = int(math.ceil(size))                                 # len(np.arange(0, size, 1))
        c0 = size // 2                                                # :409
        for j in range(J):
            if joints_vis is not None and joints_vis[n][j] == 0:      # :392-393
                continue
            mu_x, mu_y = int(pose[j, 0] / fs[0]), int(pose[j, 1] / fs[1])          # :397-398
            ul = [int(mu_x - tmp), int(mu_y - tmp)]                                # :399
            br = [int(mu_x + tmp + 1), int(mu_y + tmp + 1)]                        # :400
            if ul[0] >= W or ul[1] >= H or br[0] < 0 or br[1] < 0:                 # :401-403
                continue
            for py in range(max(0, ul[1]), min(br[1], H, ul[1] + length)):         # :436-445
                for px in range(max(0, ul[0]), min(br[0], W, ul[0] + length)):
                    dx, dy = (px - ul[0]) - c0, (py - ul[1]) - c0
                    g = np.float32(math.exp(-(dx * dx + dy * dy) / (2 * cur_sigma ** 2)))   # :411
                    if g > target[j, py, px]:
                        target[j, py, px] = g
    return np.clip(target, 0, 1)                                      # :446
