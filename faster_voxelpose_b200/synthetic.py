"""Seeded synthetic inputs for tests and bench (numpy ``Generator`` only, so the same
arrays come out in the authoring container and on the GPU box).

Nothing here is on the product path: it builds camera rigs, heatmaps, proposals and
deterministic conv weights of the shapes BASELINE.json names.

Reference recipes restated here:
* resize transform: ``lib/dataset/JointsDataset.py:68-78`` (``_get_resize_transform``) via
  ``lib/utils/transforms.py:15-50,81-92`` for ``rot == 0``;
* Gaussian input heatmaps: ``lib/dataset/JointsDataset.py:368-447`` (sigma 3, max-splat,
  augmentation off);
* numpy projection: ``lib/utils/cameras.py:58-82`` (``project_point_cpu``).
"""
from __future__ import annotations

import json
import os
from typing import Dict, List, Sequence

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
RIGS_JSON = os.path.join(_HERE, "data", "rigs.json")

__all__ = [
    "load_rig", "dome_rig", "resize_transform", "noise_heatmaps", "gaussian_heatmaps",
    "synthetic_proposals", "seeded_state_dict", "distinct_peaks_map",
]


# --------------------------------------------------------------------------- rigs
def _as_cam(c: Dict) -> Dict:
    return {
        "R": np.asarray(c["R"], dtype=np.float64).reshape(3, 3),
        "T": np.asarray(c["T"], dtype=np.float64).reshape(3, 1),
        "fx": float(c["fx"]), "fy": float(c["fy"]), "cx": float(c["cx"]), "cy": float(c["cy"]),
        "k": np.asarray(c["k"], dtype=np.float64).reshape(3, 1),
        "p": np.asarray(c["p"], dtype=np.float64).reshape(2, 1),
    }


def load_rig(name: str) -> List[Dict]:
    """Camera dicts (numpy float64, the reference's convention) for 'panoptic' (the demo
    5-camera rig with real distortion), 'shelf' or 'campus'."""
    with open(RIGS_JSON, "r") as fh:
        rigs = json.load(fh)
    return [_as_cam(c) for c in rigs[name]["cameras"]]


def dome_rig(num_views: int = 31, seed: int = 0, radius: float = 5000.0,
             focal: float = 2980.0, image: Sequence[int] = (4096, 4096),
             center: Sequence[float] = (0.0, 0.0, 1000.0)) -> List[Dict]:
    """Cameras on a dome of ``radius`` mm looking at ``center`` with Panoptic-like
    distortion (BASELINE.json config 5)."""
    rng = np.random.default_rng(seed)
    cams = []
    ctr = np.asarray(center, dtype=np.float64)
    for i in range(num_views):
        az = 2.0 * np.pi * (i + 0.25 * rng.random()) / num_views
        el = np.deg2rad(15.0 + 40.0 * rng.random())
        pos = ctr + radius * np.array([np.cos(az) * np.cos(el), np.sin(az) * np.cos(el), np.sin(el)])
        fwd = ctr - pos
        fwd /= np.linalg.norm(fwd)
        right = np.cross(fwd, np.array([0.0, 0.0, 1.0]))
        right /= np.linalg.norm(right)
        down = np.cross(fwd, right)
        R = np.stack([right, down, fwd])  # world -> camera (x right, y down, z forward)
        cams.append({
            "R": R, "T": pos.reshape(3, 1),
            "fx": focal * (1.0 + 0.01 * rng.standard_normal()),
            "fy": focal * (1.0 + 0.01 * rng.standard_normal()),
            "cx": image[0] / 2.0 + 8.0 * rng.standard_normal(),
            "cy": image[1] / 2.0 + 8.0 * rng.standard_normal(),
            "k": np.array([[-0.28], [0.17], [-0.037]]) * (1.0 + 0.05 * rng.standard_normal((3, 1))),
            "p": np.array([[-2e-4], [2e-5]]) * (1.0 + 0.05 * rng.standard_normal((2, 1))),
        })
    return cams


def resize_transform(ori_image_size: Sequence[int], image_size: Sequence[int]) -> np.ndarray:
    """2x3 float64 matrix taking original-image pixels to network-input pixels (letterbox
    resize about the image centre, no rotation)."""
    w, h = float(ori_image_size[0]), float(ori_image_size[1])
    wr, hr = float(image_size[0]), float(image_size[1])
    if w / wr < h / hr:
        w_pad, h_pad = h / hr * wr, h
    else:
        w_pad, h_pad = w, w / wr * hr
    a = wr / w_pad if w_pad >= h_pad else hr / h_pad
    return np.array([[a, 0.0, wr * 0.5 - a * (w / 2.0)],
                     [0.0, a, hr * 0.5 - a * (h / 2.0)]], dtype=np.float64)


# --------------------------------------------------------------------------- heatmaps
def noise_heatmaps(seed: int, B: int, V: int, J: int, H: int, W: int) -> np.ndarray:
    """Family N: U[0,1) noise -- the worst case for fp32 sensitivity and gather locality."""
    rng = np.random.default_rng(seed)
    return rng.random((B, V, J, H, W), dtype=np.float32)


def _project_np(x: np.ndarray, cam: Dict) -> np.ndarray:
    xc = cam["R"] @ (x.T - cam["T"])
    y = xc[:2] / (xc[2] + 1e-5)
    r = np.sum(y ** 2, axis=0)
    k, p = cam["k"].reshape(3), cam["p"].reshape(2)
    d = 1 + k[0] * r + k[1] * r * r + k[2] * r * r * r
    u = y[0] * d + 2 * p[0] * y[0] * y[1] + p[1] * (r + 2 * y[0] * y[0])
    v = y[1] * d + 2 * p[1] * y[0] * y[1] + p[0] * (r + 2 * y[1] * y[1])
    return np.stack([cam["fx"] * u + cam["cx"], cam["fy"] * v + cam["cy"]], axis=1)


def gaussian_heatmaps(seed: int, cfg, cams: List[Dict], B: int, max_people: int = 6,
                      sigma: float = 3.0) -> np.ndarray:
    """Family G: 1..max_people skeletons per frame, projected into every view and splatted
    as sigma-pixel Gaussians with a max (the recipe the reference uses for 'gt' heatmaps)."""
    rng = np.random.default_rng(seed)
    V, J = len(cams), cfg.DATASET.NUM_JOINTS
    W, H = cfg.DATASET.HEATMAP_SIZE
    t = resize_transform(cfg.DATASET.ORI_IMAGE_SIZE, cfg.DATASET.IMAGE_SIZE)
    stride = np.asarray(cfg.DATASET.IMAGE_SIZE, dtype=np.float64) / np.asarray([W, H], dtype=np.float64)
    size = np.asarray(cfg.CAPTURE_SPEC.SPACE_SIZE, dtype=np.float64)
    ctr = np.asarray(cfg.CAPTURE_SPEC.SPACE_CENTER, dtype=np.float64)
    out = np.zeros((B, V, J, H, W), dtype=np.float32)
    rad = int(3 * sigma)
    ax = np.arange(-rad, rad + 1, dtype=np.float32)
    g = np.exp(-(ax[None, :] ** 2 + ax[:, None] ** 2) / (2.0 * sigma * sigma)).astype(np.float32)
    for b in range(B):
        for _ in range(int(rng.integers(1, max_people + 1))):
            root = ctr + (rng.random(3) - 0.5) * size * np.array([0.7, 0.7, 0.0])
            root[2] = 900.0 + 200.0 * rng.random()
            joints = root[None, :] + rng.normal(0.0, [220.0, 220.0, 420.0], size=(J, 3))
            joints[:, 2] = np.clip(joints[:, 2], 0.0, 1800.0)
            for v, cam in enumerate(cams):
                px = _project_np(joints, cam)
                px = px @ t[:, :2].T + t[:, 2]
                px = px / stride
                for j in range(J):
                    mx, my = int(px[j, 0]), int(px[j, 1])
                    x0, x1 = max(0, mx - rad), min(W, mx + rad + 1)
                    y0, y1 = max(0, my - rad), min(H, my + rad + 1)
                    if x0 >= x1 or y0 >= y1:
                        continue
                    tile = out[b, v, j, y0:y1, x0:x1]
                    np.maximum(tile, g[y0 - (my - rad):y1 - (my - rad), x0 - (mx - rad):x1 - (mx - rad)],
                               out=tile)
    return out


# --------------------------------------------------------------------------- proposals
def synthetic_proposals(seed: int, cfg, B: int, bbox: str = "full") -> np.ndarray:
    """Per frame MAX_PEOPLE proposal records ``[x, y, z, match, conf, bw, bh]`` on voxel
    centres of the space grid.  ``bbox``: 'full' -> 1.0 (full 64^3 cubes),
    'crop' -> U(0.4, 0.9), 'mixed' -> crops plus one empty-range and one clipped proposal."""
    rng = np.random.default_rng(seed)
    P = cfg.CAPTURE_SPEC.MAX_PEOPLE
    n = np.asarray(cfg.CAPTURE_SPEC.VOXELS_PER_AXIS, dtype=np.int64)
    size = np.asarray(cfg.CAPTURE_SPEC.SPACE_SIZE, dtype=np.float32)
    ctr = np.asarray(cfg.CAPTURE_SPEC.SPACE_CENTER, dtype=np.float32)
    scale = size / (n - 1).astype(np.float32)
    bias = ctr - size / np.float32(2.0)
    out = np.zeros((B, P, 7), dtype=np.float32)
    for b in range(B):
        flat = rng.choice(int(n[0] * n[1] * n[2]), size=P, replace=False)
        idx = np.stack(np.unravel_index(flat, tuple(n)), axis=1).astype(np.float32)
        out[b, :, 0:3] = idx * scale + bias
        out[b, :, 3] = 0.0
        out[b, :, 4] = 1.0
        if bbox == "full":
            out[b, :, 5:7] = 1.0
        else:
            out[b, :, 5:7] = rng.uniform(0.4, 0.9, size=(P, 2)).astype(np.float32)
        if bbox == "mixed" and P >= 3:
            out[b, 0, 0:3] = bias + np.array([0.0, 0.5, 0.5], np.float32) * size   # on a space face
            out[b, 1, 0:3] = bias + np.array([1.0, 1.0, 1.0], np.float32) * size   # far corner
            out[b, 2, 5:7] = 1.2                                                    # negative crop -> clamp 0
            out[b, P - 1, 0:3] = bias - np.array([3000.0, 0.0, 0.0], np.float32)    # outside -> empty range
    return out


def distinct_peaks_map(seed: int, B: int, X: int, Y: int, peaks: int = 24) -> np.ndarray:
    """A ``[B,1,X,Y]`` detection map with ``peaks`` isolated local maxima of strictly
    distinct heights over a low positive noise floor (tie-free top-k)."""
    rng = np.random.default_rng(seed)
    out = (rng.random((B, 1, X, Y), dtype=np.float32) * np.float32(0.01)).astype(np.float32)
    for b in range(B):
        heights = (0.2 + 0.75 * rng.permutation(peaks) / peaks).astype(np.float32)
        taken = np.zeros((X, Y), dtype=bool)
        placed = 0
        while placed < peaks:
            x, y = int(rng.integers(0, X)), int(rng.integers(0, Y))
            if taken[max(0, x - 2):x + 3, max(0, y - 2):y + 3].any():
                continue
            taken[x, y] = True
            out[b, 0, x, y] = heights[placed]
            placed += 1
    return out


# --------------------------------------------------------------------------- weights
def seeded_state_dict(module, seed: int = 0, conv_std: float = 0.05):
    """Deterministic weights for any torch module: every floating parameter/buffer is drawn
    from a numpy generator in sorted-key order, so this package's nets and the reference's
    (same ``state_dict`` keys) receive identical values."""
    import torch

    rng = np.random.default_rng(seed)
    sd = module.state_dict()
    new = {}
    for key in sorted(sd.keys()):
        ref = sd[key]
        if not ref.dtype.is_floating_point:
            new[key] = ref.clone()
            continue
        shape = tuple(ref.shape)
        if key.endswith("running_var"):
            arr = 0.5 + rng.random(shape)
        elif key.endswith("running_mean"):
            arr = 0.1 * rng.standard_normal(shape)
        elif key.endswith("weight") and len(shape) == 1:      # BatchNorm gamma
            arr = 0.8 + 0.4 * rng.random(shape)
        elif key.endswith("bias"):
            arr = 0.02 * rng.standard_normal(shape)
        else:
            fan_in = int(np.prod(shape[1:])) if len(shape) > 1 else shape[0]
            arr = rng.standard_normal(shape) * (conv_std if len(shape) > 2 else 1.0 / np.sqrt(fan_in))
            if len(shape) > 2:
                arr *= np.sqrt(2.0 / max(fan_in, 1)) / conv_std
        if any(tag in key for tag in ("output_layer.", "output_hm.", "output_size.")):
            arr = arr * 0.02      # heatmap-like O(0.1..1) head outputs, as trained nets produce (beta = 100 downstream)
        new[key] = torch.as_tensor(np.asarray(arr, dtype=np.float32)).reshape(shape)
    return new
