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


def test_filter_packing_layouts_of_the_tensor_core_kernels():
    """Host-side packing (no GPU): res2d_pack lays filters out as [tap][cin/4][cout][4] (the K-major operand planes of
    fvp_res2d_block_tf32), conv7_fragments as mma B fragments with the k4 tail for 20-channel inputs; both round to TF32."""
    gen = torch.Generator().manual_seed(3)

    def rna(t):
        return ((t.contiguous().view(torch.int32) + 0x1000) & ~0x1FFF).view(torch.float32)
    for cin in (16, 32):
        w1, w2 = torch.randn((32, cin, 3, 3), generator=gen), torch.randn((32, 32, 3, 3), generator=gen)
        ws = torch.randn((32, cin, 1, 1), generator=gen)
        b1, b2 = torch.randn(32, generator=gen), torch.randn(32, generator=gen)
        p = ops.res2d_pack((w1, b1), (w2, b2), ws)
        assert p["cin"] == cin and p["w1"].shape == (9, cin // 4, 32, 4) and p["w2"].shape == (9, 8, 32, 4) and p["ws"].shape == (cin // 4, 32, 4)
        for t, pl, o, i in ((0, 0, 0, 0), (4, cin // 4 - 1, 31, 3), (8, 1, 17, 2), (5, 2, 3, 1)):
            assert p["w1"][t, pl, o, i] == rna(w1)[o, 4 * pl + i, t // 3, t % 3]
            assert p["ws"][pl, o, i] == rna(ws)[o, 4 * pl + i, 0, 0]
        assert p["w2"][7, 6, 30, 1] == rna(w2)[30, 25, 2, 1]
        assert torch.equal(p["b1"], b1) and torch.equal(p["b2"], b2)
        assert ops.res2d_pack((w1, b1), (w2, b2), None)["ws"] is None
    for cin in (16, 20):
        w = torch.randn((16, cin, 7, 7), generator=gen)
        f = ops.conv7_fragments(w)
        assert f.numel() == 49 * 2 * 32 * 4 + (49 * 32 * 2 if cin == 20 else 0)
        main = f[:49 * 2 * 32 * 4].view(49, 2, 8, 4, 2, 2)            # [tap][ks][g][t][nt][h]
        for tap, ks, g, t, nt, h in ((0, 0, 0, 0, 0, 0), (48, 1, 7, 3, 1, 1), (20, 1, 2, 1, 0, 1)):
            assert main[tap, ks, g, t, nt, h] == rna(w)[nt * 8 + g, ks * 8 + h * 4 + t, tap // 7, tap % 7]
        if cin == 20:
            tail = f[49 * 2 * 32 * 4:].view(49, 8, 4, 2)               # [tap][g][t][nt]
            for tap, g, t, nt in ((0, 0, 0, 0), (48, 7, 3, 1), (11, 4, 2, 0)):
                assert tail[tap, g, t, nt] == rna(w)[nt * 8 + g, 16 + t, tap // 7, tap % 7]
    with pytest.raises(L.FvpError):
        ops.conv7_fragments(torch.zeros(16, 24, 7, 7))
