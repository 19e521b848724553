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
            hm_cl = ops.rowpair(heatmaps)
        frame_rig = rigs.frame_rig(meta['seq'][:batch_size])

        # channels-last, joints padded to a multiple of 4: CenterNet's first conv reads it as is
        plane_xy = pl.project_plane(hm_cl, frame_rig, rigs, geom, nhwc=True)
        proposal_heatmaps_2d, bbox_preds = self.center_net(plane_xy)
        topk_2d_confs, topk_2d_index, topk_2d_flatten_index = nms2D(proposal_heatmaps_2d.detach(), self.max_people)
        match_bbox_preds, _ = ops.gather_topk(bbox_preds.detach(), None, topk_2d_flatten_index)
        feature_1d = pl.project_columns(hm_cl, frame_rig, rigs, geom, topk_2d_flatten_index)
        proposal_heatmaps_1d = self.c2c_net(torch.flatten(feature_1d, 0, 1)).view(batch_size, self.max_people, -1)
        proposal_centers = self.proposal_layer(topk_2d_index, topk_2d_confs, proposal_heatmaps_1d.detach(),
                                               match_bbox_preds, meta)
        bbox_preds = torch.flatten(bbox_preds, 2, 3).permute(0, 2, 1)
        return proposal_heatmaps_2d, proposal_heatmaps_1d, proposal_centers, bbox_preds
