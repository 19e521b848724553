"""TEST INFRASTRUCTURE -- torch-CPU restatement of the reference hot path (oracle, port #2).

This file is a checker and the timed CPU baseline only.  Nothing under
``faster_voxelpose_b200/`` imports it; only ``tests/``, ``__graft_entry__.smoke()`` and
``bench.py``'s ``cpu_baseline`` / ``--impl reference`` legs do.

Why it exists next to the C oracle (``fvp_oracle.c``): the reference's arithmetic lives in
torch (``F.grid_sample``, ``torch.mm``, ``topk``, ``softmax`` ...), so the most
faithful "reference CPU torch path" is the same sequence of torch calls.  Each function cites
the reference lines it follows.  It is pinned by ``tests/golden/`` (vectors produced by the
unmodified reference modules in the authoring container, see ``tests/golden/make_golden.py``)
and, while ``/root/reference`` is present, compared directly in ``tests/test_oracle_vs_reference.py``.

Parity status: PINNED by golden vectors generated from the reference itself (the reference
ships no tests of its own).
"""
from __future__ import annotations

from typing import Dict, List, Sequence, Tuple

import torch
import torch.nn.functional as F


# ------------------------------------------------------------------ geometry (a1-a4)
def axis_tables(space_size, space_center, voxels) -> List[torch.Tensor]:
    """Per-axis voxel-centre coordinates (project_whole.py:62-70: linspace then + centre)."""
    return [torch.linspace(-space_size[a] / 2, space_size[a] / 2, int(voxels[a])) + space_center[a]
            for a in range(3)]


def world_grid(space_size, space_center, voxels) -> torch.Tensor:
    """[X*Y*Z, 3] voxel centres, x slowest / z fastest (project_whole.py:43-79)."""
    tx, ty, tz = axis_tables(space_size, space_center, voxels)
    gx, gy, gz = torch.meshgrid(tx, ty, tz, indexing="ij")
    return torch.stack([gx.reshape(-1), gy.reshape(-1), gz.reshape(-1)], dim=1)


def project_points(pts: torch.Tensor, cam: Dict) -> torch.Tensor:
    """World mm -> original-image pixels (cameras.py:11-18,30-56). No z>0 test (F5)."""
    R = torch.as_tensor(cam["R"], dtype=torch.float)
    T = torch.as_tensor(cam["T"], dtype=torch.float)
    f = torch.tensor([cam["fx"], cam["fy"]], dtype=torch.float64).to(torch.float).reshape(2, 1)
    c = torch.tensor([cam["cx"], cam["cy"]], dtype=torch.float64).to(torch.float).reshape(2, 1)
    k = torch.as_tensor(cam["k"], dtype=torch.float)
    p = torch.as_tensor(cam["p"], dtype=torch.float)
    xcam = torch.mm(R, pts.T - T)
    y = xcam[:2] / (xcam[2] + 1e-5)
    r = torch.sum(y ** 2, dim=0)
    d = 1 + k[0] * r + k[1] * r * r + k[2] * r * r * r
    u = y[0, :] * d + 2 * p[0] * y[0, :] * y[1, :] + p[1] * (r + 2 * y[0, :] * y[0, :])
    v = y[1, :] * d + 2 * p[1] * y[0, :] * y[1, :] + p[0] * (r + 2 * y[1, :] * y[1, :])
    y = torch.stack([u, v])
    return (f * y + c).T


def sample_grid(pts: torch.Tensor, cam: Dict, geom, resize_transform: torch.Tensor) -> torch.Tensor:
    """Normalised grid_sample coordinates [N, 2] for one view (project_whole.py:81-117,
    transforms.py:59-63)."""
    w, h = geom.HEATMAP_SIZE
    xy = project_points(pts, cam)
    xy = torch.clamp(xy, -1.0, max(geom.ORI_IMAGE_SIZE[0], geom.ORI_IMAGE_SIZE[1]))
    homo = torch.cat([xy, torch.ones(xy.shape[0], 1)], dim=1)
    xy = torch.t(torch.mm(resize_transform, torch.t(homo))[:2, :])
    xy = xy * torch.tensor([w, h], dtype=torch.float) / torch.tensor(geom.IMAGE_SIZE, dtype=torch.float)
    g = xy / torch.tensor([w - 1, h - 1], dtype=torch.float) * 2.0 - 1.0
    return torch.clamp(g, -1.1, 1.1)


def sample_grids(pts: torch.Tensor, cams: Sequence[Dict], geom, resize_transform) -> torch.Tensor:
    """[V, N, 2] -- what the reference caches per sequence (project_whole.py:151-156)."""
    return torch.stack([sample_grid(pts, cams[v], geom, resize_transform) for v in range(len(cams))])


# ------------------------------------------------------------------ space cube (a5, a6)
def project_whole(heatmaps: torch.Tensor, grids: torch.Tensor, voxels) -> torch.Tensor:
    """[B,V,J,H,W] + cached [V,nbins,2] -> [B,J,X,Y,Z] (project_whole.py:119-168):
    grid_sample(bilinear, zeros, align_corners=True), mean over ALL views (F4), clamp."""
    B, V, J = heatmaps.shape[:3]
    nb = grids.shape[1]
    out = torch.zeros(B, J, 1, nb, device=heatmaps.device)
    g = grids.view(V, 1, nb, 2)
    for i in range(B):
        out[i] = torch.mean(F.grid_sample(heatmaps[i], g, align_corners=True), dim=0)
    return out.clamp(0.0, 1.0).view(B, J, int(voxels[0]), int(voxels[1]), int(voxels[2]))


# ------------------------------------------------------------------ proposals (a7-a10)
def nms_topk(prob_map: torch.Tensor, max_num: int):
    """3x3 max-equality NMS + top-k (proposal.py:34-76).  flat index is divided by
    prob_map[0].shape[1] == X (F8)."""
    B = prob_map.shape[0]
    mx = F.max_pool2d(prob_map, kernel_size=3, stride=1, padding=1)
    kept = (prob_map == mx).float() * prob_map
    vals, flat = kept.reshape(B, -1).topk(max_num)
    div = prob_map[0].shape[1]
    idx = torch.stack([torch.div(flat, div, rounding_mode="trunc"), flat % div], dim=2)
    return vals, idx, flat


def gather_columns(cubes: torch.Tensor, flat: torch.Tensor) -> torch.Tensor:
    """z-columns of the space cube at the top-k cells -> [B,P,J,Z] (human_detection_net.py:199-200)."""
    B, J, X, Y, Z = cubes.shape
    cols = cubes.reshape(B, J, X * Y, Z).permute(0, 2, 1, 3)
    return torch.gather(cols, 1, flat.view(B, -1, 1, 1).expand(-1, -1, J, Z))


def proposal_records(idx3: torch.Tensor, confs: torch.Tensor, bbox: torch.Tensor, spec) -> torch.Tensor:
    """[B,P,7] = [x,y,z mm, match, conf, bw, bh] in eval mode (human_detection_net.py:23-37,82-125)."""
    scale = (torch.tensor(spec.SPACE_SIZE) / (torch.tensor(spec.VOXELS_PER_AXIS) - 1)).to(confs.device)
    bias = (torch.tensor(spec.SPACE_CENTER) - torch.tensor(spec.SPACE_SIZE) / 2.0).to(confs.device)
    B, P = confs.shape
    rec = torch.zeros(B, P, 7, device=confs.device)
    rec[:, :, 0:3] = idx3.float() * scale + bias
    rec[:, :, 4] = confs
    rec[:, :, 3] = (confs > spec.MIN_SCORE).float() - 1.0
    rec[:, :, 5:7] = bbox
    return rec


# ------------------------------------------------------------------ per-person cubes (a11-a13)
class IndividualSpec:
    """Constants of project_individual.ProjectLayer.__init__/save_grid (:29-111)."""

    def __init__(self, cfg):
        self.whole_center = torch.tensor(cfg.CAPTURE_SPEC.SPACE_CENTER)
        self.whole_size = torch.tensor(cfg.CAPTURE_SPEC.SPACE_SIZE)
        self.ind_size = torch.tensor(cfg.INDIVIDUAL_SPEC.SPACE_SIZE)
        self.voxels = torch.tensor(cfg.INDIVIDUAL_SPEC.VOXELS_PER_AXIS, dtype=torch.int32)
        self.fine = (self.whole_size / self.ind_size * (self.voxels - 1)).int() + 1
        self.scale = (self.fine.float() - 1) / self.whole_size
        self.bias = - self.ind_size / 2.0 / self.whole_size * (self.fine - 1) \
                    - self.scale * (self.whole_center - self.whole_size / 2.0)
        n = [int(v) for v in self.voxels]
        grid = world_grid(self.ind_size.tolist(), self.whole_center.tolist(), n).view(n[0], n[1], n[2], 3)
        self.center_grid = torch.stack([grid[:, :, 0, :2].reshape(-1, 2),
                                        grid[:, 0, :, ::2].reshape(-1, 2),
                                        grid[0, :, :, 1:].reshape(-1, 2)])
        self.fine_pts = world_grid(self.whole_size.tolist(), self.whole_center.tolist(),
                                   [int(v) for v in self.fine])

    def to(self, device):
        """Constants are computed on the CPU (so they match the golden run) and then moved."""
        for name in ("whole_center", "whole_size", "ind_size", "voxels", "fine", "scale", "bias", "center_grid"):
            setattr(self, name, getattr(self, name).to(device))
        return self


def individual_windows(spec: IndividualSpec, proposals: torch.Tensor):
    """centers_tl, offset, crop start/end per proposal (project_individual.py:255-269)."""
    tl = torch.round(proposals[:, :3].float() * spec.scale + spec.bias).int()
    offset = tl.float() / (spec.fine - 1) * spec.whole_size - spec.whole_size / 2.0 + spec.ind_size / 2.0
    m = ((1 - proposals[:, 5:7]) / 2 * (spec.voxels[0:2] - 1)).int()
    m[m < 0] = 0
    m = torch.cat([m, torch.zeros((proposals.shape[0], 1), dtype=torch.int32, device=proposals.device)], dim=1)
    start = torch.where(tl + m >= 0, tl + m, torch.zeros_like(tl))
    end = torch.where(tl + spec.voxels - m <= spec.fine, tl + spec.voxels - m, spec.fine)
    return tl, offset, start, end


def project_individual(heatmaps_i: torch.Tensor, fine_grids: torch.Tensor, spec: IndividualSpec,
                       proposals: torch.Tensor) -> Tuple[torch.Tensor, torch.Tensor]:
    """One frame's [V,J,H,W] heatmaps + cached fine grid [V,fx,fy,fz,2] -> cubes
    [P',J,nx,ny,nz], offset [P',3] (project_individual.py:222-293)."""
    V, J = heatmaps_i.shape[:2]
    n = [int(v) for v in spec.voxels]
    P = proposals.shape[0]
    cubes = torch.zeros(P, J, n[0], n[1], n[2], device=heatmaps_i.device)
    tl, offset, start, end = individual_windows(spec, proposals)
    for i in range(P):
        if torch.sum(start[i] >= end[i]) > 0:
            continue
        s, e, t = start[i].tolist(), end[i].tolist(), tl[i].tolist()
        g = fine_grids[:, s[0]:e[0], s[1]:e[1], s[2]:e[2]].reshape(V, 1, -1, 2)
        acc = torch.mean(F.grid_sample(heatmaps_i, g, align_corners=True), dim=0)
        cubes[i, :, s[0] - t[0]:e[0] - t[0], s[1] - t[1]:e[1] - t[1], s[2] - t[2]:e[2] - t[2]] = \
            acc.view(J, e[0] - s[0], e[1] - s[1], e[2] - s[2])
    return cubes.clamp(0.0, 1.0), offset


def planes3(cubes: torch.Tensor) -> torch.Tensor:
    """xy / xz / yz max-planes concatenated on dim 0 (joint_localization_net.py:158-160)."""
    return torch.cat([torch.max(cubes, dim=4)[0], torch.max(cubes, dim=3)[0], torch.max(cubes, dim=2)[0]])


# ------------------------------------------------------------------ soft-argmax + fusion (a14, a15)
def soft_argmax(feat: torch.Tensor, center_grid: torch.Tensor, beta: float):
    """[3,N,J,n,n] -> coords [3,N,J,2], confs [N] (joint_localization_net.py:32-56)."""
    N, J = feat.size(1), feat.size(2)
    x = F.softmax(beta * feat.reshape(3, N, J, -1, 1), dim=3)
    confs, _ = torch.max(x, dim=3)
    confs = torch.mean(confs.squeeze(3), dim=(0, 2))
    x = torch.sum(torch.mul(x, center_grid.reshape(3, 1, 1, -1, 2)), dim=3)
    return x, confs


def fuse_planes(pose: torch.Tensor, weights: torch.Tensor) -> torch.Tensor:
    """pose [3,N,J,2] (offset already added), weights [3N,J,1] -> [N,J,3]
    (joint_localization_net.py:83-120)."""
    wxy, wxz, wyz = torch.chunk(weights, 3)
    xy, xz, yz = pose[0], pose[1], pose[2]
    wx = torch.cat([wxy, wxz], dim=2)
    wy = torch.cat([wxy, wyz], dim=2)
    wz = torch.cat([wxz, wyz], dim=2)
    wx = wx / torch.sum(wx, dim=2).unsqueeze(2)
    wy = wy / torch.sum(wy, dim=2).unsqueeze(2)
    wz = wz / torch.sum(wz, dim=2).unsqueeze(2)
    px = wx[:, :, :1] * xy[:, :, :1] + wx[:, :, 1:] * xz[:, :, :1]
    py = wy[:, :, :1] * xy[:, :, 1:] + wy[:, :, 1:] * yz[:, :, :1]
    pz = wz[:, :, :1] * xz[:, :, 1:] + wz[:, :, 1:] * yz[:, :, 1:]
    return torch.cat([px, py, pz], dim=2)


def add_offsets(pose: torch.Tensor, offset: torch.Tensor) -> torch.Tensor:
    """In-place plane offsets (joint_localization_net.py:168-171)."""
    off = offset.reshape(-1, 1, 3)
    pose[0] += off[:, :, :2]
    pose[1] += off[:, :, ::2]
    pose[2] += off[:, :, 1:]
    return pose


# ------------------------------------------------------------------ whole forward (a16 + caller)
class RefTorchPipeline:
    """The reference's eval forward from heatmaps (faster_voxelpose.py:85-93,159-160,
    human_detection_net.py:157-220, joint_localization_net.py:122-181) with the conv nets
    passed in (they stay PyTorch modules on both arms).  Grids are cached per sequence the
    way the reference does (F3)."""

    def __init__(self, cfg, center_net, c2c_net, p2p_net, weight_net, device="cpu"):
        self.cfg = cfg
        self.device = torch.device(device)
        self.center_net, self.c2c_net, self.p2p_net, self.weight_net = center_net, c2c_net, p2p_net, weight_net
        self.spec = IndividualSpec(cfg).to(self.device)
        cs = cfg.CAPTURE_SPEC
        self.whole_pts = world_grid(cs.SPACE_SIZE, cs.SPACE_CENTER, cs.VOXELS_PER_AXIS)
        self._whole: Dict[str, torch.Tensor] = {}
        self._fine: Dict[str, torch.Tensor] = {}

    def grids(self, seq, cams, rt):
        if seq not in self._whole:
            rt = rt.cpu()          # grids are built on the CPU (as in the golden run), then moved
            self._whole[seq] = sample_grids(self.whole_pts, cams, self.cfg.DATASET, rt).to(self.device)
            f = [int(v) for v in self.spec.fine]
            self._fine[seq] = sample_grids(self.spec.fine_pts, cams, self.cfg.DATASET, rt) \
                .view(len(cams), f[0], f[1], f[2], 2).to(self.device)
        return self._whole[seq], self._fine[seq]

    @torch.no_grad()
    def detect(self, heatmaps, meta, cameras, rt):
        cs = self.cfg.CAPTURE_SPEC
        B, V, J = heatmaps.shape[:3]
        cubes = torch.stack([project_whole(heatmaps[i:i + 1], self.grids(meta["seq"][i], cameras[meta["seq"][i]], rt)[0],
                                           cs.VOXELS_PER_AXIS)[0] for i in range(B)])
        hm2d, bbox = self.center_net(cubes)
        vals, idx2, flat = nms_topk(hm2d.detach(), cs.MAX_PEOPLE)
        bbox_flat = torch.flatten(bbox, 2, 3).permute(0, 2, 1)
        match_bbox = torch.gather(bbox_flat, 1, flat.unsqueeze(2).repeat(1, 1, 2))
        cols = gather_columns(cubes, flat)
        hm1d = self.c2c_net(torch.flatten(cols, 0, 1)).view(B, cs.MAX_PEOPLE, -1)
        v1, i1 = hm1d.detach().topk(1)
        idx3 = torch.cat([idx2, i1], dim=2)
        rec = proposal_records(idx3, vals * v1.squeeze(2), match_bbox, cs)
        return hm2d, hm1d, rec, bbox_flat

    @torch.no_grad()
    def localize(self, meta, heatmaps, rec, mask, cameras, rt):
        B, P = rec.shape[:2]
        J = heatmaps.shape[2]
        fused = torch.zeros(B, P, J, 3, device=heatmaps.device)
        planes = torch.zeros(3, B, P, J, 2, device=heatmaps.device)
        for i in range(B):
            if torch.sum(mask[i]) == 0:
                continue
            _, fine = self.grids(meta["seq"][i], cameras[meta["seq"][i]], rt)
            cubes, offset = project_individual(heatmaps[i], fine, self.spec, rec[i, mask[i]])
            feat = torch.stack(torch.chunk(self.p2p_net(planes3(cubes)), 3), dim=0)
            pose, confs = soft_argmax(feat, self.spec.center_grid, self.cfg.NETWORK.BETA)
            pose = add_offsets(pose, offset)
            w = self.weight_net(feat)
            fused[i, mask[i]] = fuse_planes(pose, w)
            planes[:, i, mask[i]] = pose
            rec[i, mask[i], 4] = confs
        return fused, planes

    @torch.no_grad()
    def forward(self, heatmaps, meta, cameras, rt, proposals=None):
        """Returns (fused_poses [B,P,J,5], plane_poses, proposal_centers).  ``proposals``
        overrides the detected records while detection still runs."""
        hm2d, hm1d, rec, bbox = self.detect(heatmaps, meta, cameras, rt)
        if proposals is not None:
            rec = proposals.clone()
        mask = rec[:, :, 3] >= 0
        fused, planes = self.localize(meta, heatmaps, rec, mask, cameras, rt)
        B, P = rec.shape[:2]
        J = heatmaps.shape[2]
        fused = torch.cat([fused, rec[:, :, 3:5].reshape(B, -1, 1, 2).repeat(1, 1, J, 1)], dim=3)
        return fused, planes, rec
