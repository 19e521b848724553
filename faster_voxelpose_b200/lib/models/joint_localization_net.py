"""JointLocalizationNet: drop-in for reference ``lib/models/joint_localization_net.py``.

Same submodule names (``conv_net, weight_net, project_layer, soft_argmax_layer``) and
``forward(meta, heatmaps, proposal_centers, mask, cameras, resize_transform) ->
(all_fused_pose_preds [B,P,J,3], all_pose_preds [3,B,P,J,2])`` including the in-place update of
``proposal_centers[..., 4]`` (reference :122-181).  The reference loops over frames and people with
host syncs; here all valid proposals of the batch go through one projection launch, one plane
reduction, one P2PNet / WeightNet call and two small kernels, with a single host sync (the count
of valid proposals).
"""
from __future__ import annotations

import torch
import torch.nn as nn

from ... import ops
from .cnns import P2PNet, WeightNet
from .project_individual import ProjectLayer


class SoftArgmaxLayer(nn.Module):
    """softmax(beta*x) expectation over the plane cells (reference :15-56)."""

    def __init__(self, cfg):
        super().__init__()
        self.beta = cfg.NETWORK.BETA

    def forward(self, x, grids):
        """x [3,N,J,n,n] -> (coords [3,N,J,2], confs [N])."""
        pose, maxprob = ops.softargmax(x, grids, self.beta)
        return pose, maxprob.mean(dim=(0, 2))


class JointLocalizationNet(nn.Module):
    def __init__(self, cfg):
        super().__init__()
        self.conv_net = P2PNet(cfg.DATASET.NUM_JOINTS, cfg.DATASET.NUM_JOINTS)
        self.weight_net = WeightNet(cfg)
        self.project_layer = ProjectLayer(cfg)
        self.soft_argmax_layer = SoftArgmaxLayer(cfg)
        # True: every one of the B*P proposal slots is processed and the invalid ones are masked on the device --
        # fixed shapes and no host synchronisation, which is what CUDA-graph capture needs (runtime.GraphedInference);
        # False: the valid slots are compacted first (one host sync, less conv work when few people are present)
        self.static_shapes = False

    def fuse_pose_preds(self, pose_preds, weights):
        """pose_preds [3,N,J,2] (offsets already added), weights [3N,J,1] -> [N,J,3] (reference :83-120)."""
        N, J = pose_preds.shape[1:3]
        zero = torch.zeros((N, 3), dtype=torch.float32, device=pose_preds.device)
        ones = torch.ones((3, N, J), dtype=torch.float32, device=pose_preds.device)
        fused, _ = ops.fuse(pose_preds.clone(), ones, zero, weights)
        return fused

    def forward(self, meta, heatmaps, proposal_centers, mask, cameras, resize_transform, hm_cl=None):
        device = heatmaps.device
        batch_size, max_proposals = proposal_centers.shape[0], proposal_centers.shape[1]
        n, num_joints = heatmaps.shape[1], heatmaps.shape[2]
        all_fused_pose_preds = torch.zeros((batch_size, max_proposals, num_joints, 3), device=device)
        all_pose_preds = torch.zeros((3, batch_size, max_proposals, num_joints, 2), device=device)

        valid = None
        if self.static_shapes:
            sel = None
            nv = batch_size * max_proposals
            valid = mask.reshape(-1).to(torch.uint8).contiguous()
        else:
            sel = torch.nonzero(mask.reshape(-1)).squeeze(1)   # the one host sync of the forward
            nv = int(sel.numel())
        if nv == 0:
            return all_fused_pose_preds, all_pose_preds
        pl = self.project_layer
        for i in range(batch_size):
            curr_seq = meta['seq'][i]
            assert curr_seq in cameras.keys(), "missing camera parameters for the current sequence"
            assert len(cameras[curr_seq]) == n, "inconsistent number of cameras"
        rigs = pl.rigs(cameras, n, device)
        geom = pl.geom(n, resize_transform)
        if hm_cl is None:
            hm_cl = ops.rowpair_for_planes(heatmaps, layers=(pl,), num_rigs=len(cameras))
        frame_rig = rigs.frame_rig(meta['seq'][:batch_size])

        props_all = proposal_centers.reshape(-1, 7)
        if sel is None:
            props = props_all.float().contiguous()
            frame = torch.arange(batch_size, dtype=torch.int32, device=device).repeat_interleave(max_proposals)
        else:
            props = props_all[sel].float().contiguous()
            frame = torch.div(sel, max_proposals, rounding_mode='trunc').to(torch.int32)
        nx, ny, nz = [int(v) for v in pl.consts.voxels]
        assert nx == ny == nz, "the three planes share one [3,N,J,n,n] buffer (equal voxel counts, as in the reference)"
        # the three planes of every person, channels-last with the joints padded to a multiple of 4:
        # [3,N,n,n,Jp] in memory = the [3N,Jp,n,n] channels_last batch P2PNet's first conv reads as is
        jp = ops.padded_joints(num_joints)
        planes = torch.empty((3, nv, nx, ny, jp), dtype=torch.float32, device=device)
        (_, _, _), offset = pl.project_planes(hm_cl, frame_rig, rigs, geom, props, frame.contiguous(), valid=valid,
                                              out=(planes[0], planes[1], planes[2]), nhwc=True)
        planes = planes.view(3 * nv, nx, ny, jp).permute(0, 3, 1, 2)

        joint_features = self.conv_net(planes).view(3, nv, num_joints, nx, ny)
        pose_preds, maxprob = ops.softargmax(joint_features, pl.center_grid, self.soft_argmax_layer.beta)
        weights = self.weight_net(joint_features)
        fused_pose_preds, confs = ops.fuse(pose_preds, maxprob, offset, weights)

        if sel is None:
            keep = mask.reshape(-1)
            zero = torch.zeros((), dtype=torch.float32, device=device)
            all_fused_pose_preds = torch.where(keep.view(-1, 1, 1), fused_pose_preds, zero).view(batch_size, max_proposals,
                                                                                                 num_joints, 3)
            all_pose_preds = torch.where(keep.view(1, -1, 1, 1), pose_preds, zero).view(3, batch_size, max_proposals,
                                                                                        num_joints, 2)
            pc = proposal_centers.view(-1, 7) if proposal_centers.is_contiguous() else None
            if pc is not None:
                pc[:, 4] = torch.where(keep, confs, pc[:, 4])
            else:
                proposal_centers[..., 4] = torch.where(mask, confs.view(batch_size, max_proposals), proposal_centers[..., 4])
            return all_fused_pose_preds, all_pose_preds
        all_fused_pose_preds.view(-1, num_joints, 3)[sel] = fused_pose_preds
        all_pose_preds.view(3, -1, num_joints, 2)[:, sel] = pose_preds
        if proposal_centers.is_contiguous():
            proposal_centers.view(-1, 7)[sel, 4] = confs
        else:
            b_idx = torch.div(sel, max_proposals, rounding_mode='trunc')
            proposal_centers[b_idx, sel - b_idx * max_proposals, 4] = confs
        return all_fused_pose_preds, all_pose_preds
