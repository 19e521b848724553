"""Accuracy metrics of the reference's dataset adapters, as functions of arrays.

The reference computes them inside its dataset classes, tied to the files on disk:
``lib/dataset/panoptic.py:214-312`` (AP@25..150 mm, recall@500 mm, MPJPE@500 mm) and
``lib/dataset/shelf.py:162-256`` / ``lib/dataset/campus.py:138-230`` (PCP per actor and per bone group,
recall@500 mm, after mapping the 17 COCO joints to the 14 joints of those datasets).  Here the same
numbers come from the fused poses this package produces (``[P,J,5]`` per frame: x, y, z in mm, a match
flag that is >= 0 for real detections, the detection score) and the ground truth passed as arrays, so
the README tables of the reference can be reproduced wherever the data sets are available.

Host-side numpy: a few thousand poses per data set, not a hot path -- except that the poses live on the device: when
``evaluate_panoptic`` is handed one CUDA tensor [F,P,J,5] the pose-to-person matching runs there (``ops.pose_match`` /
``fvp_pose_match``, double precision) and only errors and indices come back; the AP bookkeeping is sequential and stays here.
"""
from __future__ import annotations

from collections import OrderedDict
from typing import Dict, List, Optional, Sequence, Tuple

import numpy as np

__all__ = ["evaluate_panoptic", "evaluate_pcp", "coco_to_shelf", "coco_to_campus", "ap_and_recall"]

PANOPTIC_THRESHOLDS_MM = (25, 50, 75, 100, 125, 150)
# limbs of the 14-joint Shelf/Campus skeleton scored by PCP; the tenth "bone" is hip centre -> head bottom
PCP_LIMBS = ((0, 1), (1, 2), (3, 4), (4, 5), (6, 7), (7, 8), (9, 10), (10, 11), (12, 13))
PCP_BONE_GROUPS = OrderedDict([("Head", [8]), ("Torso", [9]), ("Upper arms", [5, 6]), ("Lower arms", [4, 7]),
                               ("Upper legs", [1, 2]), ("Lower legs", [0, 3])])


def _as_numpy(x) -> np.ndarray:
    if hasattr(x, "detach"):
        x = x.detach().cpu().numpy()
    return np.asarray(x)


# ------------------------------------------------------------------------------------------ Panoptic
def _match_records(preds, gt_joints, gt_vis) -> Tuple[np.ndarray, np.ndarray, np.ndarray, int]:
    """Every valid predicted pose -> (error to its closest ground-truth person, score, global id of
    that person); frames without ground truth are skipped as in panoptic.py:224-229."""
    err, score, gid = [], [], []
    total_gt = 0
    for pred, gt, vis in zip(preds, gt_joints, gt_vis):
        gt = np.asarray(gt, dtype=np.float64)
        if gt.shape[0] == 0:
            continue
        pred = _as_numpy(pred)
        pred = pred[pred[:, 0, 3] >= 0]
        seen = np.asarray(vis) > 0.1                                   # [n_gt, J]
        if pred.shape[0]:
            d = np.sqrt(((pred[:, None, :, 0:3] - gt[None, :, :, 0:3]) ** 2).sum(-1))   # [n_pred, n_gt, J]
            with np.errstate(invalid="ignore", divide="ignore"):
                mean = (d * seen[None]).sum(-1) / seen.sum(-1)[None]   # a person without visible joints -> nan
            mean = np.where(np.isnan(mean), np.inf, mean) if np.isnan(mean).any() else mean
            best = mean.argmin(1)
            err.append(mean[np.arange(len(best)), best])
            score.append(pred[:, 0, 4])
            gid.append(total_gt + best)
        total_gt += gt.shape[0]
    if not err:
        return np.zeros(0), np.zeros(0), np.zeros(0, np.int64), total_gt
    return np.concatenate(err), np.concatenate(score), np.concatenate(gid).astype(np.int64), total_gt


def _match_records_device(preds, gt_joints, gt_vis) -> Tuple[np.ndarray, np.ndarray, np.ndarray, int]:
    """``_match_records`` for poses that are still on the device (``preds``: one CUDA tensor [F,P,J,5]): the
    pose-to-person distances are computed by ``ops.pose_match`` (one launch for the whole set, double precision) and only
    the [F,P] errors / indices come back."""
    import torch

    from . import ops
    F, P, J = int(preds.shape[0]), int(preds.shape[1]), int(preds.shape[2])
    counts = np.array([len(g) for g in gt_joints], dtype=np.int32)
    G = max(1, int(counts.max()) if F else 1)
    gt = np.zeros((F, G, J, 3), dtype=np.float64)
    vis = np.zeros((F, G, J), dtype=np.float32)
    for i, (g, v) in enumerate(zip(gt_joints, gt_vis)):
        if counts[i]:
            gt[i, :counts[i]] = np.asarray(g, dtype=np.float64)[:, :, :3]
            vis[i, :counts[i]] = np.asarray(v, dtype=np.float32)
    dev = preds.device
    err, best = ops.pose_match(preds.float().contiguous(), torch.from_numpy(gt).to(dev), torch.from_numpy(vis).to(dev),
                               torch.from_numpy(counts).to(dev))
    err, best, score = err.cpu().numpy(), best.cpu().numpy(), preds[:, :, 0, 4].cpu().numpy()
    offset = np.concatenate(([0], np.cumsum(counts)[:-1])) if F else np.zeros(0, np.int64)
    keep = best >= 0                                              # valid pose in a frame that has ground truth
    gid = (offset[:, None] + best)[keep].astype(np.int64)
    return err[keep], score[keep].astype(np.float64), gid, int(counts.sum())


def _greedy_hits(err, gid, threshold) -> np.ndarray:
    """In score order: a pose is a hit if it is closer than ``threshold`` to a person nobody claimed yet."""
    hit = np.zeros(len(err), dtype=bool)
    claimed = set()
    for i in range(len(err)):
        if err[i] < threshold and int(gid[i]) not in claimed:
            hit[i] = True
            claimed.add(int(gid[i]))
    return hit


def ap_and_recall(err, score, gid, total_gt: int, threshold: float) -> Tuple[float, float]:
    """Average precision (area under the monotone precision envelope) and final recall at one threshold
    (panoptic.py:266-293)."""
    order = np.argsort(-np.asarray(score), kind="stable")
    hit = _greedy_hits(np.asarray(err)[order], np.asarray(gid)[order], threshold)
    tp = np.cumsum(hit)
    fp = np.cumsum(~hit)
    recall = tp / (total_gt + 1e-5)
    precision = tp / (tp + fp + 1e-5)
    precision = np.maximum.accumulate(precision[::-1])[::-1] if len(precision) else precision
    precision = np.concatenate(([0.0], precision, [0.0]))
    recall = np.concatenate(([0.0], recall, [1.0]))
    step = np.where(recall[1:] != recall[:-1])[0]
    return float(np.sum((recall[step + 1] - recall[step]) * precision[step + 1])), float(recall[-2])


def evaluate_panoptic(preds: Sequence, gt_joints: Sequence, gt_vis: Sequence) -> Tuple[float, str, Dict]:
    """preds[i]: [P,J,5]; gt_joints[i]: [n_i,J,3] mm; gt_vis[i]: [n_i,J].  Returns (mean AP, the reference's
    message line, details) -- ``Panoptic.evaluate`` (panoptic.py:214-264)."""
    assert len(preds) == len(gt_joints) == len(gt_vis), "number mismatch"
    on_device = hasattr(preds, "is_cuda") and preds.is_cuda and preds.dim() == 4
    err, score, gid, total_gt = (_match_records_device if on_device else _match_records)(preds, gt_joints, gt_vis)
    aps, recs = [], []
    for t in PANOPTIC_THRESHOLDS_MM:
        ap, rec = ap_and_recall(err, score, gid, total_gt, t)
        aps.append(ap)
        recs.append(rec)
    order = np.argsort(-score, kind="stable")
    hit500 = _greedy_hits(err[order], gid[order], 500)
    mpjpe = float(err[order][hit500].mean()) if hit500.any() else float("inf")
    recall = len(np.unique(gid[err < 500])) / total_gt if total_gt else 0.0
    msg = "Evaluation results on Panoptic dataset:\n" + "\t".join(
        f"ap@{t}: {a:.4f}" for t, a in zip(PANOPTIC_THRESHOLDS_MM, aps)) + \
        f"\trecall@500mm: {recall:.4f}\tmpjpe@500mm: {mpjpe:.3f}"
    details = {"ap": dict(zip(PANOPTIC_THRESHOLDS_MM, aps)), "recall_at_ap": dict(zip(PANOPTIC_THRESHOLDS_MM, recs)),
               "recall@500mm": recall, "mpjpe@500mm": mpjpe, "total_gt": total_gt, "num_pred": int(len(err))}
    return float(np.mean(aps)), msg, details


# ------------------------------------------------------------------------------------- Shelf / Campus
_COCO_LIMB_ORDER = np.array([16, 14, 12, 11, 13, 15, 10, 8, 6, 5, 7, 9])


def _head_points(coco: np.ndarray) -> Tuple[np.ndarray, np.ndarray]:
    mid_shoulder = (coco[5] + coco[6]) / 2
    head_center = (coco[3] + coco[4]) / 2                      # between the ears
    bottom = (mid_shoulder + head_center) / 2
    return bottom, bottom + (head_center - bottom) * 2


def coco_to_campus(coco_pose) -> np.ndarray:
    """17 COCO joints -> the 14 Campus joints: limbs re-ordered, head bottom/top interpolated from
    shoulders and ears (campus.py:214-230)."""
    coco = np.asarray(coco_pose, dtype=np.float64)[:, :3]
    out = np.zeros((14, 3))
    out[:12] = coco[_COCO_LIMB_ORDER]
    out[12], out[13] = _head_points(coco)
    return out


def coco_to_shelf(coco_pose) -> np.ndarray:
    """17 COCO joints -> the 14 Shelf joints; the two head joints blend a nose-based estimate (weight 0.75)
    with the ear-based one (shelf.py:229-256)."""
    coco = np.asarray(coco_pose, dtype=np.float64)[:, :3]
    out = np.zeros((14, 3))
    out[:12] = coco[_COCO_LIMB_ORDER]
    bottom, top = _head_points(coco)
    neck = (out[8] + out[9]) / 2                              # middle of the shoulders
    nose = coco[0]
    top_nose = neck + (nose - neck) * np.array([0.75, 0.75, 1.5])
    bottom_nose = neck + (nose - neck) * np.array([0.5, 0.5, 0.5])
    out[13] = top_nose * 0.75 + top * 0.25
    out[12] = bottom_nose * 0.75 + bottom * 0.25
    return out


def evaluate_pcp(preds: Sequence, actors_gt: Sequence[Sequence[Optional[np.ndarray]]], convert=coco_to_shelf,
                 recall_threshold: float = 500.0, alpha: float = 0.5, skip_empty_frames: bool = False
                 ) -> Tuple[float, str, Dict]:
    """Percentage of correct parts (shelf.py:162-227, campus.py:138-212).

    preds[i]: [P,17,5] fused poses of evaluated frame i; actors_gt[a][i]: [14,3] mm of actor a in that frame
    or None when the actor is absent.  ``convert`` maps a COCO pose to the data set's 14 joints.
    ``skip_empty_frames``: Campus skips frames without any detection (campus.py:156-157); Shelf does not.
    Returns (mean PCP of the first three actors, the reference's message line, details)."""
    num_actor = len(actors_gt)
    correct = np.zeros(num_actor)
    total = np.zeros(num_actor)
    bone_correct = np.zeros((num_actor, 10))
    total_gt = match_gt = 0
    for i, pred in enumerate(preds):
        pred = _as_numpy(pred)
        pred = pred[pred[:, 0, 3] >= 0, :, :3]
        if len(pred) == 0:
            if skip_empty_frames:
                continue
            raise ValueError(f"frame {i}: no detection to evaluate (the reference's Shelf.evaluate fails here too)")
        poses = np.stack([convert(p) for p in pred])                      # [n, 14, 3]
        for a in range(num_actor):
            gt = actors_gt[a][i]
            if gt is None or len(gt) == 0:
                continue
            gt = np.asarray(gt, dtype=np.float64)
            mpjpe = np.sqrt(((gt[None] - poses) ** 2).sum(-1)).mean(-1)
            best = int(mpjpe.argmin())
            match_gt += int(mpjpe[best] < recall_threshold)
            total_gt += 1
            p = poses[best]
            starts = [p[s] for s, _ in PCP_LIMBS] + [(p[2] + p[3]) / 2]
            ends = [p[e] for _, e in PCP_LIMBS] + [p[12]]
            gstarts = [gt[s] for s, _ in PCP_LIMBS] + [(gt[2] + gt[3]) / 2]
            gends = [gt[e] for _, e in PCP_LIMBS] + [gt[12]]
            for b in range(10):
                total[a] += 1
                err = (np.linalg.norm(starts[b] - gstarts[b]) + np.linalg.norm(ends[b] - gends[b])) / 2.0
                if err <= alpha * np.linalg.norm(gstarts[b] - gends[b]):
                    correct[a] += 1
                    bone_correct[a, b] += 1
    actor_pcp = correct / (total + 1e-8)
    avg = float(np.mean(actor_pcp[:3]))
    bone_pcp = OrderedDict((k, bone_correct[:, v].sum(-1) / (total / 10 * len(v) + 1e-8)) for k, v in PCP_BONE_GROUPS.items())
    recall = match_gt / (total_gt + 1e-8)
    cells = "".join(f"  {v * 100:.2f}  |" for v in list(actor_pcp[:3]) + [avg])
    msg = "     | Actor 1 | Actor 2 | Actor 3 | Average | \n PCP |" + cells + f"\t Recall@500mm: {recall:.4f}"
    return avg, msg, {"actor_pcp": actor_pcp, "bone_pcp": bone_pcp, "recall@500mm": recall, "total_gt": total_gt}
