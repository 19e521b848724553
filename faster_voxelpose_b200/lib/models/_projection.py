"""State both ProjectLayers share: packed rigs, projection constants, axis tables on device."""
from __future__ import annotations

from typing import Dict, List, Optional, Sequence

import weakref

import numpy as np
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
        self._geom_ident = None
        self._geom_ref = None
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
        """Packed camera table of ``cameras``.  The cache is keyed on CONTENT (sequence names + the packed fp32
        records, 96 bytes per view), not on ``id(cameras)``: a fresh dict at a recycled address, or a calibration
        replaced in place, rebuilds the table and drops the cached grids / tap tables that depend on it."""
        seqs = list(cameras.keys())
        packed = []
        for s in seqs:
            cams = ops._cam_list(cameras[s])
            assert len(cams) == num_views, "inconsistent number of cameras"
            packed.append(ops.pack_cameras(cams))
        blob = b"".join(p.tobytes() for p in packed)
        key = (tuple(seqs), num_views, str(device), hash(blob), len(blob))
        if self._rig_key != key or self._rigs is None or self._rigs.blob != blob:
            self._rigs = ops.RigTable(cameras, num_views, torch.device(device), packed=packed)
            self._rigs.blob = blob
            self._rig_key = key
            self.sample_grid = {}
        return self._rigs

    def geom(self, num_views: int, resize_transform):
        """Projection constants; keyed on the VALUES of ``resize_transform`` (a [2,3] matrix) -- for a device tensor
        the value check costs one 24-byte read-back, paid only when the tensor object or its version changed."""
        if isinstance(resize_transform, torch.Tensor):
            ident = (id(resize_transform), resize_transform._version, num_views)
            if self._geom_ident == ident and self._geom_ref is not None and self._geom_ref() is resize_transform:
                return self._geom
            vals = tuple(float(v) for v in resize_transform.detach().to("cpu", torch.float32).reshape(-1).tolist())
            self._geom_ident = ident
            self._geom_ref = weakref.ref(resize_transform)
        else:
            vals = tuple(float(v) for v in np.asarray(resize_transform, dtype=np.float64).astype(np.float32).reshape(-1))
            self._geom_ident, self._geom_ref = None, None
        key = (vals, num_views)
        if self._geom_key != key:
            self._geom = ops.make_geom(self.cfg, num_views, resize_transform)
            self._geom_key = key
            self.sample_grid = {}
        return self._geom

    def tap_tables_fit(self, num_rigs: int, num_views: int) -> bool:
        """Whether ``cached_grids`` will hand out tap tables for this many rigs (the row-walking planes kernel and
        the split 16 + 1 heatmap layout of 17-joint rigs need them)."""
        n = 1
        for a in self._axes_host:
            n *= int(a.numel())
        return (not self.use_cached_grid and self.use_tap_table
                and num_rigs * num_views * n * ops.L.TAP_ENTRY_BYTES <= self.tap_table_budget)

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
            out = None
            for r in range(R):   # one table at a time into a preallocated [R, ...] buffer: no second copy at the peak
                t = build(rigs.dev[r], geom, axes)
                if out is None:
                    out = torch.empty((R,) + tuple(t.shape), dtype=t.dtype, device=t.device)
                out[r].copy_(t)
                del t
            self.sample_grid = {key: out}
        return self.sample_grid[key]
