"""HumanDetectionNet: drop-in for reference ``lib/models/human_detection_net.py``.

Same submodule names (``project_layer, center_net, c2c_net, proposal_layer`` -- checkpoints load
unchanged) and ``forward(heatmaps, meta, cameras, resize_transform) ->
(proposal_heatmaps_2d, proposal_heatmaps_1d, proposal_centers, bbox_preds)`` (reference :157-220).
The projection, the z-max plane CenterNet consumes, NMS/top-k, both gathers and the proposal
records are CUDA kernels; CenterNet / C2CNet stay PyTorch modules.
"""
from __future__ import annotations

import torch
import torch.nn as nn

from ... import ops
from ..core.proposal import nms2D
from .cnns import C2CNet, CenterNet
from .project_whole import ProjectLayer


class ProposalLayer(nn.Module):
    """Voxel index -> mm, 7-float proposal record (reference :14-125).  Eval path only: the
    ground-truth matching used for training (``filter_proposal`` :39-80) is out of scope."""

    def __init__(self, cfg):
        super().__init__()
        self.max_people = cfg.CAPTURE_SPEC.MAX_PEOPLE
        self.min_score = cfg.CAPTURE_SPEC.MIN_SCORE
        self.device = torch.device(cfg.DEVICE)
        size = torch.tensor(cfg.CAPTURE_SPEC.SPACE_SIZE)
        self.scale = size / (torch.tensor(cfg.CAPTURE_SPEC.VOXELS_PER_AXIS) - 1)
        self.bias = torch.tensor(cfg.CAPTURE_SPEC.SPACE_CENTER) - size / 2.0

    def forward(self, topk_index2d, topk_2d_confs, proposal_heatmaps_1d, match_bbox_preds, meta=None):
        if self.training and meta is not None and 'roots_3d' in meta and 'num_person' in meta:
            raise NotImplementedError("ground-truth proposal matching (training) is outside the B200 hot path")
        return ops.proposal_records(topk_index2d, topk_2d_confs, proposal_heatmaps_1d, match_bbox_preds,
                                    self.scale.tolist(), self.bias.tolist(), self.min_score)


class HumanDetectionNet(nn.Module):
    def __init__(self, cfg):
        super().__init__()
        self.max_people = cfg.CAPTURE_SPEC.MAX_PEOPLE
        self.project_layer = ProjectLayer(cfg)
        self.center_net = CenterNet(cfg.DATASET.NUM_JOINTS, 1)
        self.c2c_net = C2CNet(cfg.DATASET.NUM_JOINTS, 1)
        self.proposal_layer = ProposalLayer(cfg)
        self.zshard = None      # see enable_zshard
        self._peer = {}

    # ------------------------------------------------------------------ z-sharded space grid
    def enable_zshard(self, group=None, combine: str = "auto") -> None:
        """Split the space volume into z slabs over the ranks of ``group`` (default: the world group).  Every rank
        holds the same frames; it projects only ``parallel.zslab(Z, rank, world)`` of the grid into a partial xy
        max-plane, and the planes are combined on every rank before CenterNet runs (replicated):

        * ``"nccl"``      ``all_reduce(MAX)`` over NVLink (the reduce-max of the north star);
        * ``"pull"``      the partial planes live in symmetric memory and ONE kernel loads every rank's plane through
                          peer pointers (``fvp_planes_combine_peers``);
        * ``"multimem"``  NVLS: the maximum is taken inside the NVSwitch (``fvp_planes_combine_multimem``);
        * ``"auto"``      multimem from 4 ranks up when the buffer has a multicast mapping, else pull.

        All three give the bits of the unsharded launch (values are >= 0 or NaN, the unsigned maximum is exact).  The
        z columns of the top-k cells are re-projected locally (``project_columns``), so nothing else is exchanged.
        ``disable_zshard()`` restores the single-GPU launch."""
        if combine not in ("auto", "nccl", "pull", "multimem"):
            raise ValueError(f"combine {combine!r}: expected 'auto', 'nccl', 'pull' or 'multimem'")
        self.zshard = {"group": group, "combine": combine}
        self._peer = {}

    def disable_zshard(self) -> None:
        self.zshard = None
        self._peer = {}

    def space_plane(self, hm_cl, frame_rig, rigs, geom):
        """The xy max-plane CenterNet consumes (reference cnns_2d.py:291), channels-last [B,Jp,X,Y]: one fused launch,
        or -- z-sharded -- this rank's slab plus the plane combine."""
        pl = self.project_layer
        if self.zshard is None:
            return pl.project_plane(hm_cl, frame_rig, rigs, geom, nhwc=True)
        from ... import parallel as par
        group, combine = self.zshard["group"], self.zshard["combine"]
        rank, world = par._world(group)
        X, Y, Z = [int(v) for v in pl.voxels_per_axis]
        z0, z1 = par.zslab(Z, rank, world)
        B, dev = int(hm_cl.shape[0]), hm_cl.device
        jp = ops.padded_joints(geom.J)
        if combine == "nccl" or (world == 1 and combine == "auto"):
            if z1 > z0:
                plane = pl.project_plane(hm_cl, frame_rig, rigs, geom, zslab=(z0, z1), nhwc=True)
            else:
                plane = torch.zeros((B, X, Y, jp), dtype=torch.float32, device=dev).permute(0, 3, 1, 2)
            if world > 1:
                import torch.distributed as dist
                dist.all_reduce(plane.permute(0, 2, 3, 1), op=dist.ReduceOp.MAX, group=group)   # the dense [B,X,Y,Jp] memory
            return plane
        key = (B, jp, X, Y, Z, str(dev), combine)
        peer = self._peer.get(key)
        if peer is None:
            peer = self._peer[key] = par.PeerPlanes(B, jp, X, Y, Z, dev, group=group, want=(True, False, False), mode=combine)
        slot = peer.step & 1
        peer.step += 1
        part = peer.views(slot)[0].view(B, X, Y, jp)          # same bytes, channels-last shape
        if z1 > z0:
            pl.project_plane(hm_cl, frame_rig, rigs, geom, zslab=(z0, z1), nhwc=True, out=part)
        else:
            part.zero_()
        full = peer.combine(slot)[0]
        return full.view(B, X, Y, jp).permute(0, 3, 1, 2)

    def forward(self, heatmaps, meta, cameras, resize_transform, hm_cl=None):
        batch_size, n = heatmaps.shape[0], heatmaps.shape[1]
        pl = self.project_layer
        for i in range(batch_size):
            curr_seq = meta['seq'][i]
            assert curr_seq in cameras.keys(), "missing camera parameters for the current sequence"
            assert len(cameras[curr_seq]) == n, "inconsistent number of cameras"
        rigs = pl.rigs(cameras, n, heatmaps.device)
        geom = pl.geom(n, resize_transform)
        if hm_cl is None:
            hm_cl = ops.rowpair_for_planes(heatmaps, layers=(pl,), num_rigs=len(cameras))
        frame_rig = rigs.frame_rig(meta['seq'][:batch_size])

        # channels-last, joints padded to a multiple of 4: CenterNet's first conv reads it as is
        plane_xy = self.space_plane(hm_cl, frame_rig, rigs, geom)
        proposal_heatmaps_2d, bbox_preds = self.center_net(plane_xy)
        topk_2d_confs, topk_2d_index, topk_2d_flatten_index = nms2D(proposal_heatmaps_2d.detach(), self.max_people)
        match_bbox_preds, _ = ops.gather_topk(bbox_preds.detach(), None, topk_2d_flatten_index)
        feature_1d = pl.project_columns(hm_cl, frame_rig, rigs, geom, topk_2d_flatten_index)
        proposal_heatmaps_1d = self.c2c_net(torch.flatten(feature_1d, 0, 1)).view(batch_size, self.max_people, -1)
        proposal_centers = self.proposal_layer(topk_2d_index, topk_2d_confs, proposal_heatmaps_1d.detach(),
                                               match_bbox_preds, meta)
        bbox_preds = torch.flatten(bbox_preds, 2, 3).permute(0, 2, 1)
        return proposal_heatmaps_2d, proposal_heatmaps_1d, proposal_centers, bbox_preds
