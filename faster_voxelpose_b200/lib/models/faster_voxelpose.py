"""FasterVoxelPoseNet (inference): mirrors reference ``lib/models/faster_voxelpose.py:51-162``.

``forward(backbone, views, meta, targets, input_heatmaps, cameras, resize_transform)`` returns the
reference's 5-tuple ``(fused_poses [B,P,J,5], plane_poses, proposal_centers, input_heatmaps,
loss_dict)``.  Training losses (reference :96-153) are out of scope and raise.
"""
from __future__ import annotations

import torch
import torch.nn as nn

from ... import ops
from .human_detection_net import HumanDetectionNet
from .joint_localization_net import JointLocalizationNet


class FasterVoxelPoseNet(nn.Module):
    def __init__(self, cfg):
        super().__init__()
        self.max_people = cfg.CAPTURE_SPEC.MAX_PEOPLE
        self.num_joints = cfg.DATASET.NUM_JOINTS
        self.device = torch.device(cfg.DEVICE)
        self.pose_net = HumanDetectionNet(cfg)
        self.joint_net = JointLocalizationNet(cfg)
        self.lambda_loss_2d = cfg.TRAIN.LAMBDA_LOSS_2D
        self.lambda_loss_1d = cfg.TRAIN.LAMBDA_LOSS_1D
        self.lambda_loss_bbox = cfg.TRAIN.LAMBDA_LOSS_BBOX
        self.lambda_loss_fused = cfg.TRAIN.LAMBDA_LOSS_FUSED

    def forward(self, backbone=None, views=None, meta=None, targets=None, input_heatmaps=None, cameras=None,
                resize_transform=None, proposals=None):
        """``proposals`` (extension): [B,P,7] records that replace the detected ones
        before joint localisation; detection still runs."""
        if self.training:
            raise NotImplementedError("training losses are outside the B200 inference hot path")
        if views is not None:
            num_views = views.shape[1]
            input_heatmaps = torch.stack([backbone(views[:, c]) for c in range(num_views)], dim=1)
        if not input_heatmaps.is_cuda:                       # host buffer (pinned) -> device
            input_heatmaps = input_heatmaps.to(self.device, non_blocking=True)
        input_heatmaps = input_heatmaps.contiguous()
        batch_size = input_heatmaps.shape[0]
        hm_cl = ops.rowpair_for_planes(input_heatmaps, cache=False, num_rigs=len(cameras),
                                       layers=(self.pose_net.project_layer, self.joint_net.project_layer))

        proposal_heatmaps_2d, proposal_heatmaps_1d, proposal_centers, bbox_preds = \
            self.pose_net(input_heatmaps, meta, cameras, resize_transform, hm_cl=hm_cl)
        if proposals is not None:
            proposal_centers = proposals.to(input_heatmaps.device, torch.float32).clone()
        mask = (proposal_centers[:, :, 3] >= 0)
        fused_poses, plane_poses = self.joint_net(meta, input_heatmaps, proposal_centers.detach(), mask, cameras,
                                                  resize_transform, hm_cl=hm_cl)
        fused_poses = torch.cat([fused_poses, proposal_centers[:, :, 3:5].reshape(batch_size, -1, 1, 2)
                                 .repeat(1, 1, self.num_joints, 1)], dim=3)
        return fused_poses, plane_poses, proposal_centers, input_heatmaps, None


def get(cfg):
    return FasterVoxelPoseNet(cfg)
