"""CPU: the torch restatement (oracle/ref_torch.py, also the timed CPU baseline) against golden."""
import numpy as np
import pytest
import torch

from common import RIGS, golden, heatmaps, rig_setup
from faster_voxelpose_b200 import synthetic as syn
from faster_voxelpose_b200.lib.models import cnns
from oracle import ref_torch as RT


def _pipeline(cfg):
    J = cfg.DATASET.NUM_JOINTS
    from faster_voxelpose_b200.lib.models import faster_voxelpose as m
    model = m.get(cfg).eval()
    model.load_state_dict(syn.seeded_state_dict(model, seed=0))
    return RT.RefTorchPipeline(cfg, model.pose_net.center_net, model.pose_net.c2c_net, model.joint_net.conv_net,
                               model.joint_net.weight_net)


@pytest.mark.parametrize("rig", RIGS)
def test_whole_forward_matches_reference(rig):
    cfg, cams, g = rig_setup(rig)
    pipe = _pipeline(cfg)
    # inject the golden axis tables (torch.linspace low bits are ISA dependent)
    pipe.whole_pts = torch.stack([t.reshape(-1) for t in torch.meshgrid(
        torch.from_numpy(g["whole_ax"]), torch.from_numpy(g["whole_ay"]), torch.from_numpy(g["whole_az"]),
        indexing="ij")], dim=1)
    pipe.spec.fine_pts = torch.stack([t.reshape(-1) for t in torch.meshgrid(
        torch.from_numpy(g["fine_ax"]), torch.from_numpy(g["fine_ay"]), torch.from_numpy(g["fine_az"]),
        indexing="ij")], dim=1)
    pipe.spec.center_grid = torch.from_numpy(g["center_grid"])
    # the golden run had two frames; frames are independent (eval-mode BatchNorm), so panoptic checks both and the
    # other rigs the first one only -- this is the slowest CPU test of the suite
    nf = 2 if rig == "panoptic" else 1
    hm = torch.from_numpy(syn.gaussian_heatmaps(11, cfg, cams, 2))[:nf]
    rt = torch.as_tensor(g["resize_transform"], dtype=torch.float32)
    meta = {"seq": ["seq0"] * nf}
    cameras = {"seq0": cams}
    hm2d, hm1d, rec, bbox = pipe.detect(hm, meta, cameras, rt)
    assert np.allclose(hm2d.numpy(), g["model_hm2d"][:nf], rtol=1e-4, atol=1e-5)
    assert np.allclose(rec.numpy(), g["model_rec"][:nf], rtol=1e-4, atol=1e-4)
    fused, planes, rec2 = pipe.forward(hm, meta, cameras, rt)
    assert np.abs(fused.numpy()[..., :3] - g["model_fused"][:nf, ..., :3]).max() < 0.1     # mm
    assert np.allclose(fused.numpy()[..., 3:], g["model_fused"][:nf, ..., 3:], rtol=1e-4, atol=1e-5)
    assert np.abs(planes.numpy() - g["model_planes"][:, :nf]).max() < 0.1
