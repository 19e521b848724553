"""CPU: the C oracle (oracle/fvp_oracle.c) against the golden vectors produced by the unmodified
reference (tests/golden/make_golden.py).  This is what pins the oracle."""
import numpy as np
import pytest

from common import RIGS, close, golden, heatmaps, report, rig_setup
from faster_voxelpose_b200 import synthetic as syn


def _whole_axes(g):
    return [g["whole_ax"], g["whole_ay"], g["whole_az"]]


def _fine_axes(g):
    return [g["fine_ax"], g["fine_ay"], g["fine_az"]]


@pytest.mark.parametrize("rig", RIGS)
def test_resize_transform_matches_reference(rig):
    cfg, cams, g = rig_setup(rig)
    mine = syn.resize_transform(cfg.DATASET.ORI_IMAGE_SIZE, cfg.DATASET.IMAGE_SIZE)
    assert np.array_equal(mine.astype(np.float32), g["resize_transform"].astype(np.float32))


@pytest.mark.parametrize("rig", RIGS)
def test_sample_grid_bit_exact(rig, oracle_lib):
    O = oracle_lib
    cfg, cams, g = rig_setup(rig)
    geom = O.make_geom(cfg, len(cams), g["resize_transform"])
    packed = O.pack_cameras(cams)
    wg = O.sample_grid(packed, geom, _whole_axes(g), cfg.CAPTURE_SPEC.VOXELS_PER_AXIS)
    assert np.array_equal(wg[:, g["whole_grid_idx"]], g["whole_grid_val"]), report(wg[:, g["whole_grid_idx"]], g["whole_grid_val"])
    assert np.allclose(wg.astype(np.float64).sum(1), g["whole_grid_sum"], rtol=0, atol=1e-6)
    fg = O.sample_grid(packed, geom, _fine_axes(g), g["fine_voxels"])
    assert np.array_equal(fg[:, g["fine_grid_idx"]], g["fine_grid_val"])
    assert np.allclose(fg.astype(np.float64).sum(1), g["fine_grid_sum"], rtol=0, atol=1e-5)


@pytest.mark.parametrize("rig", RIGS)
def test_individual_constants(rig, oracle_lib):
    O = oracle_lib
    cfg, cams, g = rig_setup(rig)
    spec = O.IndividualSpec(cfg)
    assert np.array_equal(spec.fine, g["fine_voxels"])
    assert np.array_equal(spec.scale, g["ind_scale"])
    assert np.array_equal(spec.bias, g["ind_bias"])
    # center_grid depends on torch.linspace's ISA-dependent low bits: ulp-level tolerance
    assert np.allclose(spec.center_grid, g["center_grid"], rtol=0, atol=1e-3)


@pytest.mark.parametrize("rig", RIGS)
@pytest.mark.parametrize("fam", ["N", "G"])
def test_space_cube(rig, fam, oracle_lib):
    O = oracle_lib
    cfg, cams, g = rig_setup(rig)
    hm = heatmaps(rig, fam)
    geom = O.make_geom(cfg, len(cams), g["resize_transform"])
    packed = O.pack_cameras(cams)[None]
    dims = cfg.CAPTURE_SPEC.VOXELS_PER_AXIS
    cube = O.project_whole(hm, packed, np.zeros(hm.shape[0], np.int32), geom, _whole_axes(g), dims)
    got = cube.reshape(-1)[g[f"{fam}_cube_idx"]]
    assert close(got, g[f"{fam}_cube_val"]), report(got, g[f"{fam}_cube_val"])
    assert np.allclose(cube.astype(np.float64).sum(axis=(2, 3, 4)), g[f"{fam}_cube_sum"], rtol=1e-7, atol=1e-4)
    xy, _, _ = O.planes3(cube)
    assert close(xy.reshape(-1)[g[f"{fam}_xy_idx"]], g[f"{fam}_xy_val"])
    # the mean's reduction order in torch is shape dependent; everything else is bit-exact
    assert (got == g[f"{fam}_cube_val"]).mean() > 0.99


@pytest.mark.parametrize("rig", RIGS)
@pytest.mark.parametrize("fam", ["N", "G"])
def test_individual_cubes(rig, fam, oracle_lib):
    O = oracle_lib
    cfg, cams, g = rig_setup(rig)
    hm = heatmaps(rig, fam)
    geom = O.make_geom(cfg, len(cams), g["resize_transform"])
    packed = O.pack_cameras(cams)[None]
    spec = O.IndividualSpec(cfg)
    spec.fine_axes = _fine_axes(g)
    props = syn.synthetic_proposals(21, cfg, hm.shape[0], bbox="mixed")
    for b in range(hm.shape[0]):
        P = props.shape[1]
        cubes, offset = O.project_individual(hm, packed, np.zeros(hm.shape[0], np.int32), geom, spec, props[b],
                                             np.full(P, b, np.int32))
        assert np.array_equal(offset, g[f"{fam}_ind{b}_offset"])
        assert np.array_equal((cubes != 0).sum(axis=(1, 2, 3, 4)), g[f"{fam}_ind{b}_nnz"])
        got = cubes.reshape(-1)[g[f"{fam}_ind{b}_idx"]]
        assert close(got, g[f"{fam}_ind{b}_val"]), report(got, g[f"{fam}_ind{b}_val"])
        assert np.allclose(cubes.astype(np.float64).sum(axis=(2, 3, 4)), g[f"{fam}_ind{b}_sum"], rtol=1e-7, atol=1e-4)
        xy, xz, yz = O.planes3(cubes)
        pl3 = np.concatenate([xy, xz, yz]).reshape(-1)
        assert close(pl3[g[f"{fam}_pl3_{b}_idx"]], g[f"{fam}_pl3_{b}_val"])


@pytest.mark.parametrize("rig", RIGS)
def test_nms_topk_bit_exact(rig, oracle_lib):
    O = oracle_lib
    cfg, cams, g = rig_setup(rig)
    X, Y = cfg.CAPTURE_SPEC.VOXELS_PER_AXIS[:2]
    pm = syn.distinct_peaks_map(3, 3, X, Y)
    vals, idx2, flat = O.nms_topk(pm, cfg.CAPTURE_SPEC.MAX_PEOPLE)
    assert np.array_equal(flat, g["nms_flat"])
    assert np.array_equal(idx2, g["nms_idx2"])
    assert np.array_equal(vals, g["nms_vals"])


@pytest.mark.parametrize("rig", RIGS)
def test_proposal_records(rig, oracle_lib):
    O = oracle_lib
    cfg, cams, g = rig_setup(rig, min_score=None)
    from faster_voxelpose_b200.config import make_cfg
    cfg = make_cfg(rig, device="cpu")
    idx3, confs, bbox = g["rec_idx3"], g["rec_confs"], g["rec_bbox"]
    Z = cfg.CAPTURE_SPEC.VOXELS_PER_AXIS[2]
    # build 1-D maps whose argmax is idx3[...,2] and whose max is 1 so conf2d*conf1d == confs
    hm1d = np.zeros(confs.shape + (Z,), np.float32)
    np.put_along_axis(hm1d, idx3[..., 2:3], 1.0, axis=2)
    rec = O.proposal_records(idx3[..., :2], confs, hm1d, bbox, cfg)
    assert np.array_equal(rec, g["rec_out"])


@pytest.mark.parametrize("rig", RIGS)
def test_softargmax_and_fuse(rig, oracle_lib):
    O = oracle_lib
    cfg, cams, g = rig_setup(rig)
    J = cfg.DATASET.NUM_JOINTS
    n = cfg.INDIVIDUAL_SPEC.VOXELS_PER_AXIS[0]
    N = 3
    feat = (np.random.default_rng(5).standard_normal((3, N, J, n, n)) * 0.02).astype(np.float32)
    peaks = np.random.default_rng(6).integers(0, n, size=(3, N, J, 2))
    for a in range(3):
        for i in range(N):
            for j in range(J):
                feat[a, i, j, peaks[a, i, j, 0], peaks[a, i, j, 1]] += 0.08
    pose, conf = O.softargmax(feat, g["center_grid"], cfg.NETWORK.BETA)
    assert np.abs(pose - g["sa_pose"]).max() < 0.02      # mm; gate is 0.1 mm
    assert np.allclose(conf, g["sa_conf"], rtol=1e-5, atol=1e-7)
    wts = np.random.default_rng(8).random((3 * N, J, 1), dtype=np.float32) + np.float32(0.05)
    offs = (np.random.default_rng(9).standard_normal((N, 3)) * 900).astype(np.float32)
    pose_off, fused = O.fuse(g["sa_pose"], offs, wts)
    assert np.array_equal(pose_off, g["sa_pose_off"])
    assert np.abs(fused - g["sa_fused"]).max() < 1e-3
