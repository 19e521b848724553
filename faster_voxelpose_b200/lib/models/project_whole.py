"""Space-level ProjectLayer: drop-in for reference ``lib/models/project_whole.py``.

Same constructor and ``forward(heatmaps, meta, cameras, resize_transform) -> cubes [B,J,X,Y,Z]``
(reference :21,:119).  Instead of caching a sample grid per sequence and calling
``F.grid_sample`` / ``mean`` / ``clamp`` per frame (:142-167), one fused CUDA kernel projects every
voxel into every view on the fly and gathers all joints for the whole batch.
"""
from __future__ import annotations

import torch

from ... import ops
from ._projection import ProjectionBase


class ProjectLayer(ProjectionBase):
    def __init__(self, cfg):
        super().__init__(cfg)
        self.space_size = cfg.CAPTURE_SPEC.SPACE_SIZE
        self.space_center = cfg.CAPTURE_SPEC.SPACE_CENTER
        self.voxels_per_axis = cfg.CAPTURE_SPEC.VOXELS_PER_AXIS
        self._axes_host = ops.axis_tables(self.space_size, self.space_center, self.voxels_per_axis)

    @property
    def grid(self) -> torch.Tensor:
        """[X*Y*Z, 3] voxel centres (reference attribute ``self.grid``, :40) -- built on demand."""
        ax, ay, az = self._axes_host
        gx, gy, gz = torch.meshgrid(ax, ay, az, indexing="ij")
        return torch.stack([gx.reshape(-1), gy.reshape(-1), gz.reshape(-1)], dim=1).to(self.device)

    def project(self, hm_cl, frame_rig, rigs, geom, want_plane: bool = False):
        """Batched entry with the cube materialised: (cubes, xy max-plane or None)."""
        dev = hm_cl.device
        grid = self.cached_grids(rigs, geom, dev)
        cubes = ops.project_cubes(hm_cl, rigs.dev, frame_rig, geom, self.axes(dev), self.voxels_per_axis,
                                  hm_cl.shape[0], cached_grid=grid, mode=self.mode)
        plane = ops.planes3(cubes, want=(True, False, False))[0] if want_plane else None
        return cubes, plane

    def project_plane(self, hm_cl, frame_rig, rigs, geom, zslab=None, nhwc=False, out=None):
        """Fused entry used by HumanDetectionNet: the xy max-plane [B,J,X,Y] CenterNet consumes
        (reference cnns_2d.py:291), reduced inside the projection kernel -- the space cube is never
        written.  ``zslab`` restricts the launch to z in [z0,z1) (z-sharded dense grids); ``out`` is an optional
        preallocated plane (e.g. a region of a symmetric-memory buffer): [B,X,Y,Jp] with ``nhwc``, else [B,J,X,Y]."""
        dev = hm_cl.device
        grid = self.cached_grids(rigs, geom, dev)
        plane = ops.project_planes(hm_cl, rigs.dev, frame_rig, geom, self.axes(dev), self.voxels_per_axis,
                                   hm_cl.shape[0], cached_grid=grid, mode=self.mode, want=(True, False, False),
                                   zslab=zslab, nhwc=nhwc, out=None if out is None else (out, None, None))[0]
        return plane.permute(0, 3, 1, 2) if (nhwc and out is not None) else plane

    def project_columns(self, hm_cl, frame_rig, rigs, geom, flat_index):
        """feature_1d [B,P,J,Z]: the z columns at the top-k cells (reference
        human_detection_net.py:199-200), re-projected instead of gathered from a cube."""
        dev = hm_cl.device
        X, Y, Z = [int(v) for v in self.voxels_per_axis]
        B, P = flat_index.shape
        win = ops.topk_windows(flat_index, X, Y, Z)
        grid = self.cached_grids(rigs, geom, dev)
        if ops.is_split17(hm_cl):
            # split 16 + 1 layout (17-joint rigs): only the row-walking planes kernel reads it -- the xz "plane" of a
            # 1 x 1 x Z window, [C][J][1][Z], is the column itself
            cols = ops.project_planes(hm_cl, rigs.dev, frame_rig, geom, self.axes(dev), [1, 1, Z], B * P, win=win,
                                      cached_grid=grid, mode=self.mode, want=(False, True, False))[1]
        else:
            cols = ops.project_cubes(hm_cl, rigs.dev, frame_rig, geom, self.axes(dev), [1, 1, Z], B * P, win=win,
                                     cached_grid=grid, mode=self.mode)
        return cols.view(B, P, -1, Z)

    def forward(self, heatmaps, meta, cameras, resize_transform):
        batch_size, n = heatmaps.shape[0], heatmaps.shape[1]
        for i in range(batch_size):
            curr_seq = meta['seq'][i]
            assert curr_seq in cameras.keys(), "missing camera parameters for the current sequence"
            assert len(cameras[curr_seq]) == n, "inconsistent number of cameras"
        rigs = self.rigs(cameras, n, heatmaps.device)
        geom = self.geom(n, resize_transform)
        hm_cl = ops.rowpair(heatmaps)
        cubes, _ = self.project(hm_cl, rigs.frame_rig(meta['seq'][:batch_size]), rigs, geom)
        return cubes
