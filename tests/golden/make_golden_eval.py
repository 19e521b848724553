"""Golden vectors for faster_voxelpose_b200/evaluation.py: the reference's own evaluate methods
(lib/dataset/panoptic.py:214-312, shelf.py:162-256, campus.py:138-230) run on seeded synthetic poses.
Run in the authoring container only (needs /root/reference):  python tests/golden/make_golden_eval.py

The dataset classes are instantiated without their __init__ (no data on disk): Panoptic gets a synthetic
``db``; Shelf/Campus read a synthetic ``actorsGT.mat`` written to a temporary directory."""
import importlib.util
import os
import sys
import tempfile
import types

import numpy as np
import scipy.io as scio
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, "/root/reference/lib")
sys.modules.setdefault("json_tricks", types.ModuleType("json_tricks"))   # imported by the dataset modules, unused here


def load(name):
    spec = importlib.util.spec_from_file_location("ref_" + name, f"/root/reference/lib/dataset/{name}.py")
    mod = importlib.util.module_from_spec(spec)
    sys.modules["dataset"] = types.ModuleType("dataset")
    jd = importlib.util.spec_from_file_location("dataset.JointsDataset", "/root/reference/lib/dataset/JointsDataset.py")
    jm = importlib.util.module_from_spec(jd)
    jd.loader.exec_module(jm)
    sys.modules["dataset.JointsDataset"] = jm
    spec.loader.exec_module(mod)
    return mod


def people(rng, n, J):
    root = rng.uniform([-3000, -3000, 800], [3000, 3000, 1200], size=(n, 1, 3))
    return root + rng.normal(0, 250, size=(n, J, 3))


def panoptic_case(rng, frames=40, J=15, P=10):
    preds, gts, viss = [], [], []
    for _ in range(frames):
        n = int(rng.integers(0, 5))
        gt = people(rng, n, J)
        vis = (rng.random((n, J)) > 0.15).astype(np.float64)
        pred = np.zeros((P, J, 5), np.float32)
        pred[:, :, 3] = -1
        k = 0
        for g in gt:                      # detections of varying quality, some missed, some duplicated
            for _ in range(int(rng.choice([0, 1, 1, 1, 2]))):
                if k < P:
                    pred[k, :, :3] = g + rng.normal(0, rng.choice([10, 40, 120, 400]), size=(J, 3))
                    pred[k, :, 3] = 0
                    pred[k, :, 4] = rng.random()
                    k += 1
        if k < P and rng.random() < 0.3:  # a false positive
            pred[k, :, :3] = people(rng, 1, J)[0]
            pred[k, :, 3] = 0
            pred[k, :, 4] = rng.random() * 0.5
        preds.append(pred)
        gts.append(gt)
        viss.append(vis)
    return preds, gts, viss


def pcp_case(rng, frames=30, actors=4, P=10, empty_frames=()):
    J = 17
    preds = []
    gt = [[None] * frames for _ in range(actors)]
    for f in range(frames):
        pred = np.zeros((P, J, 5), np.float32)
        pred[:, :, 3] = -1
        n = 0 if f in empty_frames else int(rng.integers(1, 4))
        coco = people(rng, max(n, 1), J)
        for k in range(n):
            pred[k, :, :3] = coco[k]
            pred[k, :, 3] = 0
            pred[k, :, 4] = rng.random()
        preds.append(pred)
        for a in range(actors):
            if rng.random() < 0.7:
                base = coco[int(rng.integers(0, max(n, 1)))]
                gt[a][f] = base, float(rng.choice([15, 60, 200]))
    return preds, gt


def main():
    rng = np.random.default_rng(77)
    out = {}
    pan = load("panoptic")
    preds, gts, viss = panoptic_case(rng)
    obj = pan.Panoptic.__new__(pan.Panoptic)
    obj.db_size = len(preds)
    obj.db = [{"meta": {"num_person": len(g), "joints_3d": g, "joints_3d_vis": v}} for g, v in zip(gts, viss)]
    metric, msg = pan.Panoptic.evaluate(obj, [torch.from_numpy(p) for p in preds])
    out["panoptic_preds"] = np.stack(preds)
    out["panoptic_gt_count"] = np.array([len(g) for g in gts])
    out["panoptic_gt"] = np.concatenate(gts) if sum(len(g) for g in gts) else np.zeros((0, 15, 3))
    out["panoptic_vis"] = np.concatenate(viss)
    out["panoptic_metric"] = np.array(metric)
    out["panoptic_msg"] = np.array(msg)

    for name, cls_name, skip in (("shelf", "Shelf", False), ("campus", "Campus", True)):
        mod = load(name)
        cls = getattr(mod, cls_name)
        conv = getattr(cls, "coco2shelf3D" if name == "shelf" else "coco2campus3D")
        preds, gt = pcp_case(rng, empty_frames=(3, 11) if skip else ())
        frames, actors = len(preds), len(gt)
        gt_mm = [[None] * frames for _ in range(actors)]
        cell = np.empty((1, actors), dtype=object)
        for a in range(actors):
            col = np.empty((frames, 1), dtype=object)
            for f in range(frames):
                if gt[a][f] is None:
                    col[f, 0] = np.zeros((1, 0))
                else:
                    base, noise = gt[a][f]
                    g = conv(base) + rng.normal(0, noise, size=(14, 3))
                    gt_mm[a][f] = g
                    col[f, 0] = g / 1000.0
            cell[0, a] = col
        with tempfile.TemporaryDirectory() as d:
            scio.savemat(os.path.join(d, "actorsGT.mat"), {"actor3D": cell})
            obj = cls.__new__(cls)
            obj.dataset_dir = d
            obj.frame_range = list(range(frames))
            metric, msg = cls.evaluate(obj, [torch.from_numpy(p) for p in preds])
        out[f"{name}_preds"] = np.stack(preds)
        gt_arr = np.full((actors, frames, 14, 3), np.nan)
        for a in range(actors):
            for f in range(frames):
                if gt_mm[a][f] is not None:
                    gt_arr[a, f] = gt_mm[a][f]
        out[f"{name}_gt"] = gt_arr
        out[f"{name}_metric"] = np.array(metric)
        out[f"{name}_msg"] = np.array(msg)
        sample = people(rng, 5, 17)
        out[f"{name}_convert_in"] = sample
        out[f"{name}_convert_out"] = np.stack([conv(p) for p in sample])
    np.savez_compressed(os.path.join(HERE, "golden_eval.npz"), **out)
    for k in ("panoptic_msg", "shelf_msg", "campus_msg"):
        print(str(out[k]))


if __name__ == "__main__":
    main()
