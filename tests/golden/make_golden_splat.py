"""Golden vectors for the input-heatmap producer: the reference's own
``JointsDataset.generate_input_heatmap`` (lib/dataset/JointsDataset.py:368-447), called unbound on
a stub object, on seeded 2-D poses.  Run in the authoring container only (needs /root/reference):

    python tests/golden/make_golden_splat.py
"""
import os
import sys
import types

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, "/root/reference/lib")
# dataset/__init__.py pulls in json_tricks (absent here); load the one module file directly
import importlib.util  # noqa: E402

_spec = importlib.util.spec_from_file_location("ref_joints_dataset", "/root/reference/lib/dataset/JointsDataset.py")
_mod = importlib.util.module_from_spec(_spec)
_spec.loader.exec_module(_mod)
JointsDataset = _mod.JointsDataset


def stub(image_size, heatmap_size, sigma=3):
    s = types.SimpleNamespace(image_size=np.array(image_size), heatmap_size=np.array(heatmap_size), sigma=sigma,
                              data_augmentation=False)
    s.compute_human_scale = lambda pose, vis: JointsDataset.compute_human_scale(s, pose, vis)
    return s


def seeded_poses(rng, n_people, J, image_size):
    """People of assorted sizes, some partly or wholly outside the image (clipping / skip paths)."""
    joints, vis = [], []
    w, h = image_size
    for _ in range(n_people):
        extent = rng.choice([20.0, 90.0, 300.0, 900.0])
        root = rng.uniform([-0.1 * w, -0.1 * h], [1.1 * w, 1.1 * h])
        pose = (root + rng.normal(0, extent / 3, size=(J, 2))).astype(np.float32)
        joints.append(pose)
        vis.append((rng.random(J) > 0.2).astype(np.uint8))
    return joints, vis


def main():
    rng = np.random.default_rng(20240607)
    out = {}
    cases = [("panoptic", (960, 512), (240, 128), 15), ("shelf", (800, 608), (200, 152), 17),
             ("campus", (800, 640), (200, 160), 17)]
    for name, image_size, heatmap_size, J in cases:
        s = stub(image_size, heatmap_size)
        for k, n_people in enumerate((0, 1, 4, 9)):
            joints, vis = seeded_poses(rng, n_people, J, image_size)
            use_vis = k % 2 == 1
            if n_people == 0:
                hm = np.zeros((J, heatmap_size[1], heatmap_size[0]), np.float32)
            else:
                # the reference takes float64 lists; float32-representable values keep both sides on the same input
                hm = JointsDataset.generate_input_heatmap(s, [j.astype(np.float64) for j in joints],
                                                          joints_vis=vis if use_vis else None)
            out[f"{name}_{k}_joints"] = np.stack(joints) if n_people else np.zeros((0, J, 2), np.float32)
            out[f"{name}_{k}_vis"] = np.stack(vis) if n_people else np.zeros((0, J), np.uint8)
            out[f"{name}_{k}_use_vis"] = np.array(use_vis)
            out[f"{name}_{k}_heatmap"] = hm.astype(np.float32)
        out[f"{name}_sizes"] = np.array([*image_size, *heatmap_size, J])
    # Campus with the sigma its yaml really uses (configs/campus/jln64.yaml:47 SIGMA: 4); own generator, so the
    # fixtures above stay byte-identical
    rng4 = np.random.default_rng(20241105)
    name, image_size, heatmap_size, J = "campus4", (800, 640), (200, 160), 17
    s = stub(image_size, heatmap_size, sigma=4)
    for k, n_people in enumerate((0, 1, 4, 9)):
        joints, vis = seeded_poses(rng4, n_people, J, image_size)
        use_vis = k % 2 == 1
        hm = np.zeros((J, heatmap_size[1], heatmap_size[0]), np.float32) if n_people == 0 else \
            JointsDataset.generate_input_heatmap(s, [j.astype(np.float64) for j in joints], joints_vis=vis if use_vis else None)
        out[f"{name}_{k}_joints"] = np.stack(joints) if n_people else np.zeros((0, J, 2), np.float32)
        out[f"{name}_{k}_vis"] = np.stack(vis) if n_people else np.zeros((0, J), np.uint8)
        out[f"{name}_{k}_use_vis"] = np.array(use_vis)
        out[f"{name}_{k}_heatmap"] = hm.astype(np.float32)
    out[f"{name}_sizes"] = np.array([*image_size, *heatmap_size, J])
    out[f"{name}_sigma"] = np.array(4)
    np.savez_compressed(os.path.join(HERE, "golden_splat.npz"), **out)
    print("wrote golden_splat.npz", {k: v.shape for k, v in out.items() if k.endswith("heatmap")})


if __name__ == "__main__":
    main()
