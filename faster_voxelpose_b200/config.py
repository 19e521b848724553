"""Configuration objects for the projection hot path.

The reference drives every module from one global ``edict`` (reference
``lib/core/config.py:15-144``) overlaid with a yaml (``:174-188``).  The hot path only
reads a handful of fields, so this module provides

* :func:`make_cfg` -- a nested ``types.SimpleNamespace`` with the same attribute names, so
  both this package and the unmodified reference modules accept it, and
* :data:`PRESETS` -- the five BASELINE.json workloads (values restated from the reference's
  ``configs/{panoptic,shelf,campus}/jln64.yaml`` plus the two synthetic scale-up configs).

Any object with the same attribute names (an ``easydict`` built by the reference's own
``update_config``) works as ``cfg`` everywhere in this package.
"""
from __future__ import annotations

import copy
import types
from typing import Any, Dict

__all__ = ["PRESETS", "make_cfg", "to_namespace", "from_yaml"]


def to_namespace(d: Any) -> Any:
    """Recursively turn dicts into attribute-style namespaces (lists are kept as lists)."""
    if isinstance(d, dict):
        return types.SimpleNamespace(**{k: to_namespace(v) for k, v in d.items()})
    return d


def _base(dataset: Dict[str, Any], capture: Dict[str, Any], individual=None, sigma: int = 3) -> Dict[str, Any]:
    return {
        "DEVICE": "cuda:0",
        "DATASET": dataset,
        "NETWORK": {"NUM_CHANNEL_JOINT_FEAT": 32, "NUM_CHANNEL_JOINT_HIDDEN": 64,
                    "SIGMA": sigma, "BETA": 100},
        "TRAIN": {"LAMBDA_LOSS_2D": 1.0, "LAMBDA_LOSS_1D": 1.0,
                  "LAMBDA_LOSS_BBOX": 0.1, "LAMBDA_LOSS_FUSED": 5.0},
        "CAPTURE_SPEC": capture,
        "INDIVIDUAL_SPEC": individual or {"SPACE_SIZE": [2000.0, 2000.0, 2000.0],
                                          "VOXELS_PER_AXIS": [64, 64, 64]},
    }


# (w, h) order for every *_SIZE, exactly like the reference yamls.
PRESETS: Dict[str, Dict[str, Any]] = {
    # configs/panoptic/jln64.yaml
    "panoptic": _base(
        {"CAMERA_NUM": 5, "ORI_IMAGE_SIZE": [1920, 1080], "IMAGE_SIZE": [960, 512],
         "HEATMAP_SIZE": [240, 128], "NUM_JOINTS": 15, "ROOT_JOINT_ID": 2},
        {"SPACE_SIZE": [8000.0, 8000.0, 2000.0], "SPACE_CENTER": [0.0, -500.0, 800.0],
         "VOXELS_PER_AXIS": [80, 80, 20], "MAX_PEOPLE": 10, "MIN_SCORE": 0.3}),
    # configs/shelf/jln64.yaml
    "shelf": _base(
        {"CAMERA_NUM": 5, "ORI_IMAGE_SIZE": [1032, 776], "IMAGE_SIZE": [800, 608],
         "HEATMAP_SIZE": [200, 152], "NUM_JOINTS": 17, "ROOT_JOINT_ID": [11, 12]},
        {"SPACE_SIZE": [8000.0, 8000.0, 2000.0], "SPACE_CENTER": [450.0, -320.0, 800.0],
         "VOXELS_PER_AXIS": [80, 80, 20], "MAX_PEOPLE": 10, "MIN_SCORE": 0.1}),
    # configs/campus/jln64.yaml
    "campus": _base(
        {"CAMERA_NUM": 3, "ORI_IMAGE_SIZE": [360, 288], "IMAGE_SIZE": [800, 640],
         "HEATMAP_SIZE": [200, 160], "NUM_JOINTS": 17, "ROOT_JOINT_ID": [11, 12]},
        {"SPACE_SIZE": [12000.0, 12000.0, 2000.0], "SPACE_CENTER": [3000.0, 4500.0, 1000.0],
         "VOXELS_PER_AXIS": [80, 80, 20], "MAX_PEOPLE": 5, "MIN_SCORE": 0.1},
        sigma=4),   # configs/campus/jln64.yaml:47
    # BASELINE.json config 4: dense space grid on the Panoptic rig
    "dense": _base(
        {"CAMERA_NUM": 5, "ORI_IMAGE_SIZE": [1920, 1080], "IMAGE_SIZE": [960, 512],
         "HEATMAP_SIZE": [240, 128], "NUM_JOINTS": 15, "ROOT_JOINT_ID": 2},
        {"SPACE_SIZE": [8000.0, 8000.0, 2000.0], "SPACE_CENTER": [0.0, -500.0, 800.0],
         "VOXELS_PER_AXIS": [256, 256, 64], "MAX_PEOPLE": 10, "MIN_SCORE": 0.3}),
    # BASELINE.json config 5: 31-camera stress rig (synthetic dome, see synthetic.dome_rig)
    "stress": _base(
        {"CAMERA_NUM": 31, "ORI_IMAGE_SIZE": [4096, 4096], "IMAGE_SIZE": [4096, 4096],
         "HEATMAP_SIZE": [1024, 1024], "NUM_JOINTS": 15, "ROOT_JOINT_ID": 2},
        {"SPACE_SIZE": [8000.0, 8000.0, 2000.0], "SPACE_CENTER": [0.0, 0.0, 1000.0],
         "VOXELS_PER_AXIS": [80, 80, 20], "MAX_PEOPLE": 32, "MIN_SCORE": 0.3},
        {"SPACE_SIZE": [2000.0, 2000.0, 2000.0], "VOXELS_PER_AXIS": [96, 96, 96]}),
}


def make_cfg(preset: str = "panoptic", device: str = "cuda:0", **overrides: Any):
    """Build a cfg namespace for one of :data:`PRESETS`.

    ``overrides`` uses ``SECTION__FIELD=value`` keys, e.g.
    ``make_cfg("panoptic", CAPTURE_SPEC__MIN_SCORE=-1.0)``.
    """
    d = copy.deepcopy(PRESETS[preset])
    d["DEVICE"] = device
    for key, val in overrides.items():
        sec, _, field = key.partition("__")
        if field:
            d[sec][field] = val
        else:
            d[sec] = val
    return to_namespace(d)


def from_yaml(path: str, device: str | None = None):
    """Load one of the reference's yaml files into the same namespace shape."""
    import yaml

    with open(path, "r") as fh:
        d = yaml.safe_load(fh)
    if device is not None:
        d["DEVICE"] = device
    return to_namespace(d)
