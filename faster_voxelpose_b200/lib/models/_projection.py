"""State both ProjectLayers share: packed rigs, projection constants, axis tables on device."""
from __future__ import annotations

from typing import Dict, List, Optional, Sequence

import torch
import torch.nn as nn

from ... import ops


class ProjectionBase(nn.Module):
    """Holds no parameters or buffers (like the reference layers, whose grids are plain
    attributes -- SURVEY.md section 5), so ``state_dict`` keys are unchanged by the swap."""

    def __init__(self, cfg):
        super().__init__()
        self.cfg = cfg
        self.device = torch.device(cfg.DEVICE)
        self.image_size = cfg.DATASET.IMAGE_SIZE
        self.heatmap_size = cfg.DATASET.HEATMAP_SIZE
        self.ori_image_size = cfg.DATASET.ORI_IMAGE_SIZE
        self.mode = ops.L.MODE_DIRECT          # projection kernel variant
        self.use_cached_grid = False           # A/B: read the reference-style cached grid
        # Tap tables: per rig, the sample grid with grid_sample's tap set-up already applied (20 B per view and
        # voxel), built on first use and cached like the reference's per-sequence grids
        # (project_whole.py:138-141).  Rig sets whose tables exceed the budget are projected on the fly.
        self.use_tap_table = True
        self.tap_table_budget = 16 << 30       # bytes
        self.sample_grid: Dict = {}            # per-sequence grids (cached-grid mode only)
        self._rig_key = None
        self._rigs: Optional[ops.RigTable] = None
        self._geom_key = None
        self._geom = None
        self._axes_host: List[torch.Tensor] = []
        self._axes_dev: Dict[torch.device, List[torch.Tensor]] = {}

    # -- axis tables ---------------------------------------------------------------------------
    def set_axes(self, axes: Sequence) -> None:
        """Override the voxel-centre tables (tests inject the golden tables so comparisons do not
        depend on the host's torch.linspace vector width)."""
        self._axes_host = [torch.as_tensor(a, dtype=torch.float32).clone() for a in axes]
        self._axes_dev = {}
        self.sample_grid = {}

    def axes(self, device) -> List[torch.Tensor]:
        device = torch.device(device)
        if device not in self._axes_dev:
            self._axes_dev[device] = [a.to(device) for a in self._axes_host]
        return self._axes_dev[device]

    # -- rigs / constants ----------------------------------------------------------------------
    def rigs(self, cameras: Dict, num_views: int, device) -> ops.RigTable:
        key = (id(cameras), len(cameras), num_views, str(device))
        if self._rig_key != key:
            self._rigs = ops.RigTable(cameras, num_views, torch.device(device))
            self._rig_key = key
            self.sample_grid = {}
        return self._rigs

    def geom(self, num_views: int, resize_transform):
        if isinstance(resize_transform, torch.Tensor):
            key = (id(resize_transform), resize_transform._version, num_views)
        else:
            key = (id(resize_transform), 0, num_views)
        if self._geom_key != key:
            self._geom = ops.make_geom(self.cfg, num_views, resize_transform)
            self._geom_key = key
            self.sample_grid = {}
        return self._geom

    def cached_grids(self, rigs: ops.RigTable, geom, device) -> Optional[torch.Tensor]:
        """What the projection kernels read instead of projecting, for every rig of the table:
        the reference-style grids [R][V][n][2] (``use_cached_grid``, A/B mode), else the tap tables
        [R][V][20*npad] bytes (``use_tap_table``), else None = project on the fly."""
        axes = self.axes(device)
        R, V = int(rigs.dev.shape[0]), int(rigs.dev.shape[1])
        n = int(axes[0].numel()) * int(axes[1].numel()) * int(axes[2].numel())
        if self.use_cached_grid:
            key, build = ("grid", id(rigs), str(device)), ops.sample_grid
        elif self.use_tap_table and R * V * n * ops.L.TAP_ENTRY_BYTES <= self.tap_table_budget:
            key, build = ("taps", id(rigs), str(device)), ops.tap_table
        else:
            return None
        if key not in self.sample_grid:
            self.sample_grid = {key: torch.stack([build(rigs.dev[r], geom, axes) for r in range(R)])}
        return self.sample_grid[key]
