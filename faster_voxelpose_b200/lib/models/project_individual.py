"""Per-person ProjectLayer: drop-in for reference ``lib/models/project_individual.py``.

Same constructor, public attributes (``center_grid``, ``voxels_per_axis``,
``fine_voxels_per_axis``, ``scale``, ``bias``; reference :43-85,:105-110) and
``forward(heatmaps, index, meta, proposal_centers, cameras, resize_transform) -> (cubes, offset)``
(:222).  The reference caches a 164 MB fine sample grid per sequence and loops over people with a
host sync each (:272-286); here the window arithmetic runs in one small kernel and all cubes of all
frames are projected by one batched launch.
"""
from __future__ import annotations

import numpy as np
import torch

from ... import ops
from ._projection import ProjectionBase


class IndividualConsts:
    """fp32 / int32 constants of the reference constructor (:43-85), evaluated in the same order."""

    def __init__(self, cfg):
        f32 = np.float32
        self.whole_center = np.asarray(cfg.CAPTURE_SPEC.SPACE_CENTER, dtype=f32)
        self.whole_size = np.asarray(cfg.CAPTURE_SPEC.SPACE_SIZE, dtype=f32)
        self.ind_size = np.asarray(cfg.INDIVIDUAL_SPEC.SPACE_SIZE, dtype=f32)
        self.voxels = np.asarray(cfg.INDIVIDUAL_SPEC.VOXELS_PER_AXIS, dtype=np.int32)
        steps = (self.voxels - 1).astype(f32)
        self.fine = ((self.whole_size / self.ind_size) * steps).astype(np.int32) + 1
        fm1 = (self.fine - 1).astype(f32)
        self.scale = (fm1 / self.whole_size).astype(f32)
        half = (self.ind_size / f32(2.0)) / self.whole_size
        self.bias = (-half * fm1 - self.scale * (self.whole_center - self.whole_size / f32(2.0))).astype(f32)


class ProjectLayer(ProjectionBase):
    def __init__(self, cfg):
        super().__init__(cfg)
        self.consts = c = IndividualConsts(cfg)
        dev = self.device if (self.device.type != "cuda" or torch.cuda.is_available()) else torch.device("cpu")
        self.whole_space_center = torch.tensor(cfg.CAPTURE_SPEC.SPACE_CENTER, device=dev)
        self.whole_space_size = torch.tensor(cfg.CAPTURE_SPEC.SPACE_SIZE, device=dev)
        self.ind_space_size = torch.tensor(cfg.INDIVIDUAL_SPEC.SPACE_SIZE, device=dev)
        self.voxels_per_axis = torch.tensor(cfg.INDIVIDUAL_SPEC.VOXELS_PER_AXIS, device=dev, dtype=torch.int32)
        self.fine_voxels_per_axis = torch.from_numpy(c.fine.copy()).to(dev)
        self.scale = torch.from_numpy(c.scale.copy()).to(dev)
        self.bias = torch.from_numpy(c.bias.copy()).to(dev)
        # fine whole-space tables at individual-voxel pitch (reference fine_grid, :110)
        self._axes_host = ops.axis_tables(c.whole_size.tolist(), c.whole_center.tolist(), c.fine.tolist())
        self.set_center_axes(ops.axis_tables(c.ind_size.tolist(), c.whole_center.tolist(), c.voxels.tolist()), dev)

    def set_center_axes(self, axes, device=None) -> None:
        """center_grid [3, n*n, 2]: (x,y) of the xy plane, (x,z) of xz, (y,z) of yz of an
        individual-size grid centred on the SPACE centre (reference save_grid, :91-108)."""
        ax, ay, az = [torch.as_tensor(a, dtype=torch.float32) for a in axes]
        n = [int(v) for v in self.consts.voxels]
        assert n[0] == n[1] == n[2], "center_grid needs equal voxel counts per axis (as in the reference)"
        pair = lambda a, b: torch.stack(torch.meshgrid(a, b, indexing="ij"), dim=-1).reshape(-1, 2)
        grid = torch.stack([pair(ax, ay), pair(ax, az), pair(ay, az)])
        self.center_grid = grid.to(device if device is not None else self.center_grid.device)

    def project(self, hm_cl, frame_rig, rigs, geom, props, frame, valid=None, out=None):
        """Batched entry: props [N,7] of any frames (frame [N] int32) -> (cubes, offset, win)."""
        dev = hm_cl.device
        win, offset = ops.individual_windows(props, valid, frame, self.consts)
        grid = self.cached_grids(rigs, geom, dev)
        cubes = ops.project_cubes(hm_cl, rigs.dev, frame_rig, geom, self.axes(dev), self.consts.voxels.tolist(),
                                  props.shape[0], win=win, cached_grid=grid, mode=self.mode, out=out)
        return cubes, offset, win

    def project_planes(self, hm_cl, frame_rig, rigs, geom, props, frame, valid=None, out=None, nhwc=False):
        """Fused batched entry used by JointLocalizationNet: the three max-planes of every person
        cube (reference joint_localization_net.py:158-160) straight from the projection kernel;
        returns ((xy, xz, yz) each [N,J,n,n], offset [N,3])."""
        dev = hm_cl.device
        win, offset = ops.individual_windows(props, valid, frame, self.consts)
        grid = self.cached_grids(rigs, geom, dev)
        planes = ops.project_planes(hm_cl, rigs.dev, frame_rig, geom, self.axes(dev), self.consts.voxels.tolist(),
                                    props.shape[0], win=win, cached_grid=grid, mode=self.mode, out=out, nhwc=nhwc)
        return planes, offset

    def forward(self, heatmaps, index, meta, proposal_centers, cameras, resize_transform):
        n = heatmaps.shape[1]
        curr_seq = meta['seq'][index]
        assert curr_seq in cameras.keys(), "missing camera parameters for the current sequence"
        assert len(cameras[curr_seq]) == n, "inconsistent number of cameras"
        rigs = self.rigs(cameras, n, heatmaps.device)
        geom = self.geom(n, resize_transform)
        hm_cl = ops.rowpair(heatmaps)
        frame_rig = rigs.frame_rig(meta['seq'][:heatmaps.shape[0]])
        props = proposal_centers.float().contiguous()
        frame = torch.full((props.shape[0],), int(index), dtype=torch.int32, device=heatmaps.device)
        cubes, offset, _ = self.project(hm_cl, frame_rig, rigs, geom, props, frame)
        return cubes, offset
