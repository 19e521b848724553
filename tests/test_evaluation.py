"""CPU: faster_voxelpose_b200/evaluation.py against the outputs of the reference's own evaluate methods
(tests/golden/golden_eval.npz, made by tests/golden/make_golden_eval.py)."""
import os

import numpy as np
import pytest

from faster_voxelpose_b200 import evaluation as ev

G = np.load(os.path.join(os.path.dirname(__file__), "golden", "golden_eval.npz"))


def test_panoptic_ap_recall_mpjpe():
    counts = G["panoptic_gt_count"]
    off = np.concatenate(([0], np.cumsum(counts)))
    gts = [G["panoptic_gt"][off[i]:off[i + 1]] for i in range(len(counts))]
    vis = [G["panoptic_vis"][off[i]:off[i + 1]] for i in range(len(counts))]
    metric, msg, details = ev.evaluate_panoptic(list(G["panoptic_preds"]), gts, vis)
    assert msg == str(G["panoptic_msg"])
    assert abs(metric - float(G["panoptic_metric"])) < 1e-12
    assert details["total_gt"] == int(counts.sum()) and 0 < details["recall@500mm"] < 1


def test_panoptic_edge_cases():
    J = 15
    none = np.zeros((10, J, 5), np.float32)
    none[:, :, 3] = -1
    gt = [np.zeros((2, J, 3))]
    metric, msg, d = ev.evaluate_panoptic([none], gt, [np.ones((2, J))])       # nothing detected
    assert metric == 0.0 and d["mpjpe@500mm"] == float("inf") and d["recall@500mm"] == 0.0
    metric, _, d = ev.evaluate_panoptic([none], [np.zeros((0, J, 3))], [np.zeros((0, J))])   # nothing to find
    assert metric == 0.0 and d["total_gt"] == 0
    perfect = none.copy()
    perfect[:2, :, :3] = 7.0
    perfect[:2, :, 3] = 0
    perfect[:2, :, 4] = np.array([0.9, 0.8])[:, None]
    gt = [np.stack([np.full((J, 3), 7.0), np.full((J, 3), 5000.0)])]
    metric, _, d = ev.evaluate_panoptic([perfect], gt, [np.ones((2, J))])      # duplicate detection of one person
    assert d["recall@500mm"] == 0.5 and d["mpjpe@500mm"] == 0.0 and 0.4 < metric < 0.6


@pytest.mark.parametrize("name", ["shelf", "campus"])
def test_pcp_and_joint_conversion(name):
    convert = ev.coco_to_shelf if name == "shelf" else ev.coco_to_campus
    got = np.stack([convert(p) for p in G[f"{name}_convert_in"]])
    assert np.allclose(got, G[f"{name}_convert_out"], rtol=0, atol=1e-9)
    gt = G[f"{name}_gt"]
    actors = [[None if np.isnan(gt[a, f]).any() else gt[a, f] for f in range(gt.shape[1])] for a in range(gt.shape[0])]
    metric, msg, details = ev.evaluate_pcp(list(G[f"{name}_preds"]), actors, convert=convert,
                                           skip_empty_frames=(name == "campus"))
    assert msg == str(G[f"{name}_msg"])
    assert abs(metric - float(G[f"{name}_metric"])) < 1e-9
    assert set(details["bone_pcp"]) == {"Head", "Torso", "Upper arms", "Lower arms", "Upper legs", "Lower legs"}


def test_shelf_rejects_frames_without_detection():
    preds = np.zeros((1, 10, 17, 5), np.float32)
    preds[:, :, :, 3] = -1
    with pytest.raises(ValueError):
        ev.evaluate_pcp(list(preds), [[np.zeros((14, 3))]], convert=ev.coco_to_shelf)
    metric, _, d = ev.evaluate_pcp(list(preds), [[np.zeros((14, 3))]] * 3, convert=ev.coco_to_campus, skip_empty_frames=True)
    assert metric == 0.0 and d["total_gt"] == 0
