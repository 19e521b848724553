"""CPU: host-side logic of the package -- conv nets against golden, constants, C-ABI exports."""
import ctypes
import os
import re

import numpy as np
import pytest
import torch

from common import RIGS, golden, rig_setup
from faster_voxelpose_b200 import _lib as L
from faster_voxelpose_b200 import ops, synthetic as syn
from faster_voxelpose_b200.config import make_cfg
from faster_voxelpose_b200.lib.models import faster_voxelpose as fvp_model
from faster_voxelpose_b200.lib.models.project_individual import IndividualConsts

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_library_exports_every_declared_symbol():
    """The C-ABI library loads and exports exactly what include/fvp_b200.h declares."""
    from faster_voxelpose_b200 import build
    build.build()
    header = open(os.path.join(ROOT, "include", "fvp_b200.h")).read()
    declared = set(re.findall(r"\b(fvp_[a-z0-9_]+)\s*\(", header))
    assert declared == set(L.EXPORTS), declared ^ set(L.EXPORTS)
    lib = ctypes.CDLL(L.LIB_PATH)
    for name in declared:
        assert hasattr(lib, name), name
    assert L.load().fvp_abi_version() == L.ABI_VERSION


def test_bad_arguments_are_reported_not_crashed():
    lib = L.load()
    rc = lib.fvp_planes3(None, 1, 4, 4, 4, None, None, None, None)
    assert rc == -1 and b"null" in lib.fvp_last_error()
    with pytest.raises(L.FvpError):
        L.check(rc, "fvp_planes3")


def test_no_cpu_fallback():
    with pytest.raises(L.FvpError):
        ops.channels_last(torch.zeros(1, 2, 3, 4, 5))


@pytest.mark.parametrize("rig", RIGS)
def test_individual_constants_match_reference(rig):
    cfg, cams, g = rig_setup(rig)
    c = IndividualConsts(cfg)
    assert np.array_equal(c.fine, g["fine_voxels"])
    assert np.array_equal(c.scale, g["ind_scale"])
    assert np.array_equal(c.bias, g["ind_bias"])


@pytest.mark.parametrize("rig", RIGS)
def test_conv_nets_match_reference(rig):
    """Same state_dict keys as the reference nets and the same outputs on seeded inputs."""
    cfg, cams, g = rig_setup(rig)
    J = cfg.DATASET.NUM_JOINTS
    X, Y = cfg.CAPTURE_SPEC.VOXELS_PER_AXIS[:2]
    n = cfg.INDIVIDUAL_SPEC.VOXELS_PER_AXIS[0]
    model = fvp_model.get(cfg).eval()
    model.load_state_dict(syn.seeded_state_dict(model, seed=0))
    with torch.no_grad():
        xin = torch.from_numpy(np.random.default_rng(31).random((1, J, X, Y), dtype=np.float32))
        chm, csz = model.pose_net.center_net(xin)
        c1 = model.pose_net.c2c_net(torch.from_numpy(np.random.default_rng(32).random((2, J, 20), dtype=np.float32)))
        pin = torch.from_numpy(np.random.default_rng(33).random((3, J, n, n), dtype=np.float32))
        pout = model.joint_net.conv_net(pin)
        wout = model.joint_net.weight_net(torch.stack(torch.chunk(pout, 3), dim=0))
    tol = dict(rtol=1e-4, atol=1e-5)
    assert np.allclose(chm.numpy(), g["cnn_center_hm"], **tol)
    assert np.allclose(csz.numpy()[:, :, ::4, ::4], g["cnn_center_size"], **tol)
    assert np.allclose(c1.numpy(), g["cnn_c2c"], **tol)
    assert np.allclose(pout.numpy()[:, :, ::4, ::4], g["cnn_p2p"], **tol)
    assert np.allclose(wout.numpy(), g["cnn_weight"], **tol)


def test_rig_table_and_geom_host_logic():
    cfg = make_cfg("shelf", device="cpu")
    cams = syn.load_rig("shelf")
    cameras = {"a": cams, "b": {i: c for i, c in enumerate(cams)}}     # list and int-keyed dict forms
    table = ops.RigTable(cameras, 5, torch.device("cpu"))
    assert table.dev.shape == (2, 5, L.CAM_STRIDE)
    assert torch.equal(table.dev[0], table.dev[1])
    assert table.frame_rig(["b", "a", "b"]).tolist() == [1, 0, 1]
    with pytest.raises(AssertionError):
        table.frame_rig(["missing"])
    with pytest.raises(AssertionError):
        ops.RigTable(cameras, 4, torch.device("cpu"))
    geom = ops.make_geom(cfg, 5, syn.resize_transform(cfg.DATASET.ORI_IMAGE_SIZE, cfg.DATASET.IMAGE_SIZE))
    assert (geom.W, geom.H, geom.J, geom.V) == (200, 152, 17, 5)
    assert geom.clamp_max == 1032.0 and ops.padded_joints(17) == 20


def test_rig_and_geometry_caches_are_keyed_on_content():
    """a fresh ``cameras`` dict at a recycled address, or a calibration replaced in place, must rebuild the
    packed camera table (and drop the grids / tap tables cached from it); the same holds for ``resize_transform``."""
    import copy
    from faster_voxelpose_b200.lib.models.project_whole import ProjectLayer
    cfg = make_cfg("panoptic", device="cpu")
    cams_a = syn.load_rig("panoptic")
    cams_b = copy.deepcopy(cams_a)
    cams_b[2]["T"] = cams_b[2]["T"] + 10.0
    pl = ProjectLayer(cfg)
    t1 = pl.rigs({"s": cams_a}, 5, "cpu")
    pl.sample_grid = {"sentinel": 1}
    assert pl.rigs({"s": cams_a}, 5, "cpu") is t1 and pl.sample_grid            # same content, new dict: cache hit
    t2 = pl.rigs({"s": cams_b}, 5, "cpu")                                        # new content (possibly same id): rebuilt
    assert t2 is not t1 and not np.array_equal(t2.host, t1.host) and pl.sample_grid == {}
    holder = {"s": cams_b}
    t3 = pl.rigs(holder, 5, "cpu")
    holder["s"] = cams_a                                                         # replaced in place
    t4 = pl.rigs(holder, 5, "cpu")
    assert t3 is t2 and t4 is not t3 and np.array_equal(t4.host, t1.host)
    rt = syn.resize_transform(cfg.DATASET.ORI_IMAGE_SIZE, cfg.DATASET.IMAGE_SIZE)
    g1 = pl.geom(5, torch.as_tensor(rt, dtype=torch.float32))
    assert pl.geom(5, torch.as_tensor(rt, dtype=torch.float32)) is g1            # equal values, new tensor: cache hit
    rt2 = torch.as_tensor(rt, dtype=torch.float32).clone()
    g2 = pl.geom(5, rt2)
    assert g2 is g1
    rt2[0, 2] += 1.0                                                             # in-place edit bumps the version
    assert pl.geom(5, rt2) is not g1
