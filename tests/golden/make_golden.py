#!/usr/bin/env python
"""Generate the golden vectors under tests/golden/ by running the UNMODIFIED reference
modules (imported from /root/reference/lib) on CPU torch.

Run in the authoring container only (the GPU box has no /root/reference):

    python tests/golden/make_golden.py

Outputs
* faster_voxelpose_b200/data/rigs.json   -- the three calibration rigs + reference resize transforms
* tests/golden/golden_<rig>.npz          -- sub-sampled outputs of every hot-path function
Inputs are NOT stored: they are regenerated from seeds by faster_voxelpose_b200.synthetic.
"""
import contextlib
import io
import json
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
sys.path.insert(0, "/root/reference/lib")

from faster_voxelpose_b200 import synthetic as syn          # noqa: E402
from faster_voxelpose_b200.config import make_cfg            # noqa: E402

from core.proposal import nms2D                              # noqa: E402  (reference)
from models.human_detection_net import HumanDetectionNet, ProposalLayer   # noqa: E402
from models.joint_localization_net import JointLocalizationNet, SoftArgmaxLayer  # noqa: E402
from models.faster_voxelpose import FasterVoxelPoseNet       # noqa: E402
from models.project_individual import ProjectLayer as RefIndividual   # noqa: E402
from models.project_whole import ProjectLayer as RefWhole    # noqa: E402
from utils.transforms import get_affine_transform, get_scale  # noqa: E402

CAL = {
    "panoptic": ("/root/reference/demo/calibration.json", "customized_sequence"),
    "shelf": ("/root/reference/data/Shelf/calibration_shelf.json", None),
    "campus": ("/root/reference/data/Campus/calibration_campus.json", None),
}
NSUB = 4096


def quiet(fn, *a, **k):
    with contextlib.redirect_stdout(io.StringIO()):
        return fn(*a, **k)


def read_cams(name):
    path, key = CAL[name]
    d = json.load(open(path))
    if key is not None:
        return list(d[key])
    return [d[str(i)] for i in range(len(d))]


def ref_resize(cfg):
    ori, img = cfg.DATASET.ORI_IMAGE_SIZE, cfg.DATASET.IMAGE_SIZE
    c = np.array([ori[0] / 2.0, ori[1] / 2.0])
    s = get_scale((ori[0], ori[1]), img)
    return get_affine_transform(c, s, 0, img)


def write_rigs():
    rigs = {}
    for name in CAL:
        cfg = make_cfg(name, device="cpu")
        cams = read_cams(name)
        rigs[name] = {
            "cameras": [{k: (np.asarray(c[k], dtype=np.float64).tolist() if k in "RTkp" else float(c[k]))
                         for k in ("R", "T", "fx", "fy", "cx", "cy", "k", "p")} for c in cams],
            "resize_transform": ref_resize(cfg).tolist(),
        }
    os.makedirs(os.path.dirname(syn.RIGS_JSON), exist_ok=True)
    with open(syn.RIGS_JSON, "w") as fh:
        json.dump(rigs, fh, indent=1)


def sub(rng, n, k=NSUB):
    return np.sort(rng.choice(n, size=min(k, n), replace=False)).astype(np.int64)


def golden_for(name):
    cfg = make_cfg(name, device="cpu", CAPTURE_SPEC__MIN_SCORE=-1.0)
    cams = syn.load_rig(name)
    V, J = len(cams), cfg.DATASET.NUM_JOINTS
    W, H = cfg.DATASET.HEATMAP_SIZE
    rt64 = ref_resize(cfg)
    rt = torch.as_tensor(rt64, dtype=torch.float32)
    rng = np.random.default_rng(1234)
    out = {"resize_transform": rt64}
    cameras = {"seq0": cams}
    B = 2
    meta = {"seq": ["seq0"] * B}

    whole = quiet(RefWhole, cfg)
    ind = quiet(RefIndividual, cfg)
    nb = whole.grid.shape[0]
    nf = ind.fine_grid.shape[0]

    # ---- a1-a4: sample grids (sub-sampled) + checksums of the full grids
    gi, fi = sub(rng, nb), sub(rng, nf)
    wg = torch.stack([whole.project_grid(cams[v], W, H, nb, rt, "cpu").view(nb, 2) for v in range(V)])
    fg = torch.stack([ind.project_grid(cams[v], W, H, nf, rt, "cpu").view(nf, 2) for v in range(V)])
    out.update(whole_grid_idx=gi, whole_grid_val=wg[:, gi].numpy(), whole_grid_sum=wg.double().sum(1).numpy(),
               fine_grid_idx=fi, fine_grid_val=fg[:, fi].numpy(), fine_grid_sum=fg.double().sum(1).numpy(),
               fine_voxels=ind.fine_voxels_per_axis.numpy(), ind_scale=ind.scale.numpy(), ind_bias=ind.bias.numpy(),
               center_grid=ind.center_grid.numpy())

    # axis tables exactly as torch.linspace produced them here (ISA-dependent low bits; tests
    # inject these so comparisons are bit-stable on any host)
    vx = cfg.CAPTURE_SPEC.VOXELS_PER_AXIS
    wgrid = whole.grid.view(vx[0], vx[1], vx[2], 3)
    fv = [int(v) for v in ind.fine_voxels_per_axis]
    fgrid = ind.fine_grid.view(fv[0], fv[1], fv[2], 3)
    out.update(whole_ax=wgrid[:, 0, 0, 0].numpy().copy(), whole_ay=wgrid[0, :, 0, 1].numpy().copy(),
               whole_az=wgrid[0, 0, :, 2].numpy().copy(), fine_ax=fgrid[:, 0, 0, 0].numpy().copy(),
               fine_ay=fgrid[0, :, 0, 1].numpy().copy(), fine_az=fgrid[0, 0, :, 2].numpy().copy())

    for fam in ("N", "G"):
        if fam == "N":
            hm = syn.noise_heatmaps(7, B, V, J, H, W)
        else:
            hm = syn.gaussian_heatmaps(11, cfg, cams, B)
        hm_t = torch.from_numpy(hm)
        # ---- a5/a6: space cube and its xy max-plane
        cube = quiet(whole.forward, hm_t, meta, cameras, rt)
        flat = cube.reshape(-1)
        ci = sub(rng, flat.numel(), 4 * NSUB)
        plane = torch.max(cube, dim=4)[0].reshape(-1)
        pi = sub(rng, plane.numel())
        out.update({f"{fam}_cube_idx": ci, f"{fam}_cube_val": flat[ci].numpy(),
                    f"{fam}_cube_sum": cube.double().sum(dim=(2, 3, 4)).numpy(),
                    f"{fam}_xy_idx": pi, f"{fam}_xy_val": plane[pi].numpy()})
        # ---- a12/a13: per-person cubes, offsets, three max-planes
        props = syn.synthetic_proposals(21, cfg, B, bbox="mixed")
        for b in range(B):
            p_t = torch.from_numpy(props[b])
            cubes, offset = quiet(ind.forward, hm_t, b, meta, p_t, cameras, rt)
            pl3 = torch.cat([torch.max(cubes, dim=4)[0], torch.max(cubes, dim=3)[0], torch.max(cubes, dim=2)[0]])
            cf = cubes.reshape(-1)
            cj = sub(rng, cf.numel(), 4 * NSUB)
            pf = pl3.reshape(-1)
            pj = sub(rng, pf.numel(), 2 * NSUB)
            out.update({f"{fam}_ind{b}_idx": cj, f"{fam}_ind{b}_val": cf[cj].numpy(),
                        f"{fam}_ind{b}_sum": cubes.double().sum(dim=(2, 3, 4)).numpy(),
                        f"{fam}_ind{b}_nnz": (cubes != 0).sum(dim=(1, 2, 3, 4)).numpy(),
                        f"{fam}_ind{b}_offset": offset.numpy(),
                        f"{fam}_pl3_{b}_idx": pj, f"{fam}_pl3_{b}_val": pf[pj].numpy()})

    # ---- a7: NMS + top-k on a tie-free map
    X, Y = cfg.CAPTURE_SPEC.VOXELS_PER_AXIS[:2]
    P = cfg.CAPTURE_SPEC.MAX_PEOPLE
    pm = syn.distinct_peaks_map(3, 3, X, Y)
    vals, idx2, flat = nms2D(torch.from_numpy(pm), P)
    out.update(nms_vals=vals.numpy(), nms_idx2=idx2.numpy(), nms_flat=flat.numpy())

    # ---- a9/a10: proposal records
    pl = ProposalLayer(cfg).eval()
    idx3 = torch.from_numpy(rng.integers(0, 20, size=(3, P, 3)).astype(np.int64))
    confs = torch.from_numpy(rng.random((3, P), dtype=np.float32) - np.float32(0.3))
    bbox = torch.from_numpy(rng.random((3, P, 2), dtype=np.float32))
    cfg2 = make_cfg(name, device="cpu")
    out.update(rec_idx3=idx3.numpy(), rec_confs=confs.numpy(), rec_bbox=bbox.numpy(),
               rec_out=ProposalLayer(cfg2).eval()(idx3, confs, bbox, {}).numpy())

    # ---- a14/a15: soft-argmax and plane fusion
    N = 3
    n = cfg.INDIVIDUAL_SPEC.VOXELS_PER_AXIS[0]
    feat = (np.random.default_rng(5).standard_normal((3, N, J, n, n)) * 0.02).astype(np.float32)
    peaks = np.random.default_rng(6).integers(0, n, size=(3, N, J, 2))
    for a in range(3):
        for i in range(N):
            for j in range(J):
                feat[a, i, j, peaks[a, i, j, 0], peaks[a, i, j, 1]] += 0.08
    sal = SoftArgmaxLayer(cfg)
    pose, conf = sal(torch.from_numpy(feat), ind.center_grid)
    wts = torch.from_numpy(np.random.default_rng(8).random((3 * N, J, 1), dtype=np.float32) + np.float32(0.05))
    offs = torch.from_numpy((np.random.default_rng(9).standard_normal((N, 3)) * 900).astype(np.float32))
    pose2 = pose.clone()
    o3 = offs.reshape(-1, 1, 3)
    pose2[0] += o3[:, :, :2]
    pose2[1] += o3[:, :, ::2]
    pose2[2] += o3[:, :, 1:]
    jln = quiet(JointLocalizationNet, cfg)
    fused = jln.fuse_pose_preds(pose2, wts)
    out.update(sa_pose=pose.numpy(), sa_conf=conf.numpy(), sa_pose_off=pose2.numpy(), sa_fused=fused.numpy())

    # ---- whole model (a16 + caller): seeded weights, Gaussian heatmaps, MIN_SCORE=-1
    model = quiet(FasterVoxelPoseNet, cfg).eval()
    model.load_state_dict(syn.seeded_state_dict(model, seed=0))
    hm_t = torch.from_numpy(syn.gaussian_heatmaps(11, cfg, cams, B))
    with torch.no_grad():
        hm2d, hm1d, rec, bbox_flat = quiet(model.pose_net, hm_t, meta, cameras, rt)
        fused_p, plane_p, rec2, _, _ = quiet(model, None, None, meta, None, hm_t, cameras, rt)
    out.update(model_hm2d=hm2d.numpy(), model_hm1d=hm1d.numpy(), model_rec=rec.numpy(),
               model_bbox=bbox_flat.numpy()[:, ::37].copy(), model_fused=fused_p.numpy(),
               model_planes=plane_p.numpy(), model_rec_after=rec2.numpy())
    # conv nets alone on seeded inputs (pins faster_voxelpose_b200.lib.models.cnns)
    with torch.no_grad():
        xin = torch.from_numpy(np.random.default_rng(31).random((1, J, X, Y), dtype=np.float32))
        chm, csz = model.pose_net.center_net(xin.unsqueeze(-1))
        c1 = model.pose_net.c2c_net(torch.from_numpy(np.random.default_rng(32).random((2, J, 20), dtype=np.float32)))
        pin = torch.from_numpy(np.random.default_rng(33).random((3, J, n, n), dtype=np.float32))
        pout = model.joint_net.conv_net(pin)
        wout = model.joint_net.weight_net(torch.stack(torch.chunk(pout, 3), dim=0))
    out.update(cnn_center_hm=chm.numpy(), cnn_center_size=csz.numpy()[:, :, ::4, ::4].copy(), cnn_c2c=c1.numpy(),
               cnn_p2p=pout.numpy()[:, :, ::4, ::4].copy(), cnn_weight=wout.numpy())
    np.savez_compressed(os.path.join(ROOT, "tests", "golden", f"golden_{name}.npz"), **out)
    print(name, "golden written:", len(out), "arrays")


if __name__ == "__main__":
    torch.set_num_threads(os.cpu_count() or 1)
    write_rigs()
    for rig in CAL:
        golden_for(rig)
