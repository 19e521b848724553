"""CPU, authoring container only: both oracles against the UNMODIFIED reference modules imported
from /root/reference/lib (skipped on the GPU box, where the reference tree does not exist; the
committed golden vectors carry the same comparison there)."""
import contextlib
import io
import os
import sys

import numpy as np
import pytest
import torch

from common import RIGS, rig_setup
from faster_voxelpose_b200 import synthetic as syn
from oracle import ref_torch as RT

REF = "/root/reference/lib"
pytestmark = pytest.mark.skipif(not os.path.isdir(REF), reason="reference tree not present")


@pytest.fixture(scope="module")
def ref():
    sys.path.insert(0, REF)
    import core.proposal as proposal
    import models.joint_localization_net as jln
    import models.project_individual as pind
    import models.project_whole as pwhole
    yield {"proposal": proposal, "jln": jln, "pind": pind, "pwhole": pwhole}
    sys.path.remove(REF)


def quiet(fn, *a, **k):
    with contextlib.redirect_stdout(io.StringIO()):
        return fn(*a, **k)


@pytest.mark.parametrize("rig", RIGS)
def test_torch_restatement_is_bit_identical(rig, ref):
    cfg, cams, g = rig_setup(rig)
    V, J = len(cams), cfg.DATASET.NUM_JOINTS
    W, H = cfg.DATASET.HEATMAP_SIZE
    rt = torch.as_tensor(g["resize_transform"], dtype=torch.float32)
    whole = quiet(ref["pwhole"].ProjectLayer, cfg)
    ind = quiet(ref["pind"].ProjectLayer, cfg)
    pts = RT.world_grid(cfg.CAPTURE_SPEC.SPACE_SIZE, cfg.CAPTURE_SPEC.SPACE_CENTER, cfg.CAPTURE_SPEC.VOXELS_PER_AXIS)
    assert torch.equal(pts, whole.grid)
    grids = RT.sample_grids(pts, cams, cfg.DATASET, rt)
    for v in range(V):
        assert torch.equal(grids[v], whole.project_grid(cams[v], W, H, pts.shape[0], rt, "cpu").view(-1, 2))
    spec = RT.IndividualSpec(cfg)
    assert torch.equal(spec.fine, ind.fine_voxels_per_axis) and torch.equal(spec.scale, ind.scale)
    assert torch.equal(spec.bias, ind.bias) and torch.equal(spec.center_grid, ind.center_grid)
    assert torch.equal(spec.fine_pts, ind.fine_grid)

    hm = torch.from_numpy(syn.noise_heatmaps(3, 1, V, J, H, W))
    meta, cameras = {"seq": ["q"]}, {"q": cams}
    assert torch.equal(RT.project_whole(hm, grids, cfg.CAPTURE_SPEC.VOXELS_PER_AXIS), quiet(whole.forward, hm, meta, cameras, rt))

    props = torch.from_numpy(syn.synthetic_proposals(5, cfg, 1, bbox="mixed")[0][:4])
    rc, ro = quiet(ind.forward, hm, 0, meta, props, cameras, rt)
    f = [int(x) for x in spec.fine]
    fine = RT.sample_grids(spec.fine_pts, cams, cfg.DATASET, rt).view(V, f[0], f[1], f[2], 2)
    mc, mo = RT.project_individual(hm[0], fine, spec, props)
    assert torch.equal(mc, rc) and torch.equal(mo, ro)

    X, Y = cfg.CAPTURE_SPEC.VOXELS_PER_AXIS[:2]
    pm = torch.from_numpy(syn.distinct_peaks_map(1, 2, X, Y))
    for a, b in zip(RT.nms_topk(pm, cfg.CAPTURE_SPEC.MAX_PEOPLE), ref["proposal"].nms2D(pm, cfg.CAPTURE_SPEC.MAX_PEOPLE)):
        assert torch.equal(a, b)

    n = cfg.INDIVIDUAL_SPEC.VOXELS_PER_AXIS[0]
    feat = torch.from_numpy((np.random.default_rng(2).standard_normal((3, 2, J, n, n)) * 0.03).astype(np.float32))
    rp, rcf = ref["jln"].SoftArgmaxLayer(cfg)(feat, ind.center_grid)
    mp, mcf = RT.soft_argmax(feat, spec.center_grid, cfg.NETWORK.BETA)
    assert torch.equal(mp, rp) and torch.equal(mcf, rcf)
    w = torch.rand(6, J, 1) + 0.1
    jl = quiet(ref["jln"].JointLocalizationNet, cfg)
    assert torch.equal(RT.fuse_planes(mp, w), jl.fuse_pose_preds(rp, w))


@pytest.mark.parametrize("rig", RIGS)
def test_c_oracle_cube_is_bit_identical_to_reference(rig, ref, oracle_lib):
    """Full tensors (not the golden sub-sample): C oracle == reference ProjectLayer, bit for bit."""
    O = oracle_lib
    cfg, cams, g = rig_setup(rig)
    V, J = len(cams), cfg.DATASET.NUM_JOINTS
    W, H = cfg.DATASET.HEATMAP_SIZE
    rt = torch.as_tensor(g["resize_transform"], dtype=torch.float32)
    whole = quiet(ref["pwhole"].ProjectLayer, cfg)
    hm = syn.noise_heatmaps(9, 1, V, J, H, W)
    want = quiet(whole.forward, torch.from_numpy(hm), {"seq": ["q"]}, {"q": cams}, rt).numpy()
    n = cfg.CAPTURE_SPEC.VOXELS_PER_AXIS
    grid = whole.grid.view(n[0], n[1], n[2], 3)
    axes = [grid[:, 0, 0, 0].numpy().copy(), grid[0, :, 0, 1].numpy().copy(), grid[0, 0, :, 2].numpy().copy()]
    geom = O.make_geom(cfg, V, g["resize_transform"])
    got = O.project_whole(hm, O.pack_cameras(cams)[None], np.zeros(1, np.int32), geom, axes, n)
    assert np.array_equal(got, want)


@pytest.mark.parametrize("rig", RIGS)
def test_presets_restate_the_reference_yaml(rig):
    """Every field config.PRESETS holds equals the reference's configs/<dataset>/jln64.yaml (campus
    SIGMA is 4, not 3)."""
    from faster_voxelpose_b200.config import PRESETS, from_yaml
    ref_cfg = from_yaml(f"/root/reference/configs/{rig}/jln64.yaml")
    preset = PRESETS[rig]
    checked = 0
    for section, fields in preset.items():
        if not isinstance(fields, dict):
            continue
        ref_section = getattr(ref_cfg, section)
        for key, val in fields.items():
            assert getattr(ref_section, key) == val, (rig, section, key, getattr(ref_section, key), val)
            checked += 1
    assert checked >= 20
