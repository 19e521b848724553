"""The reference's own eval forward, from its own UNMODIFIED modules (``oracle/_ref/lib``, see oracle/make_ref.py;
in the authoring container also straight from ``/root/reference/lib``).

TEST / BENCH INFRASTRUCTURE ONLY (the timed reference arm of bench.py, cross-checks in tests/): nothing under
faster_voxelpose_b200/ imports this module.

``RefModulePipeline`` composes ``HumanDetectionNet`` and ``JointLocalizationNet`` exactly as
``FasterVoxelPoseNet.forward`` does (lib/models/faster_voxelpose.py:85-93) -- the only addition is that proposal
records can be injected between the two nets (random-init weights give degenerate proposals), which is
also what this package's own bench arm does.
"""
from __future__ import annotations

import contextlib
import io
import os
import sys

import torch

HERE = os.path.dirname(os.path.abspath(__file__))
CANDIDATES = [os.path.join(HERE, "_ref", "lib"), "/root/reference/lib"]


def ref_lib_path():
    for p in CANDIDATES:
        if os.path.isfile(os.path.join(p, "models", "human_detection_net.py")):
            return p
    return None


def available() -> bool:
    return ref_lib_path() is not None


@contextlib.contextmanager
def _ref_on_path():
    """The reference's packages are called ``models``, ``core``, ``utils`` -- generic names.  They are imported with
    the reference lib directory FIRST on sys.path and are left in sys.modules afterwards (the reference's modules
    import each other lazily by those names)."""
    p = ref_lib_path()
    if p is None:
        raise RuntimeError("reference modules not available: run oracle/make_ref.py where /root/reference exists")
    sys.path.insert(0, p)
    try:
        yield p
    finally:
        sys.path.remove(p)


class RefModulePipeline:
    def __init__(self, cfg, device="cpu"):
        with _ref_on_path(), contextlib.redirect_stdout(io.StringIO()):
            from models.human_detection_net import HumanDetectionNet
            from models.joint_localization_net import JointLocalizationNet
            self.device = torch.device(device)
            self.cfg = cfg
            self.pose_net = HumanDetectionNet(cfg).to(self.device).eval()
            self.joint_net = JointLocalizationNet(cfg).to(self.device).eval()

    def load_weights_from(self, model) -> None:
        """Copy the parameters of this package's FasterVoxelPoseNet (same state_dict keys by construction)."""
        self.pose_net.load_state_dict(model.pose_net.state_dict(), strict=True)
        self.joint_net.load_state_dict(model.joint_net.state_dict(), strict=True)

    @torch.no_grad()
    def forward(self, heatmaps, meta, cameras, rt, proposals=None):
        """-> (fused_poses [B,P,J,5], plane_poses [3,B,P,J,2], proposal_centers [B,P,7]), faster_voxelpose.py:85-93,159-160."""
        with contextlib.redirect_stdout(io.StringIO()):     # the reference prints when it caches a grid
            _, _, centers, _ = self.pose_net(heatmaps, meta, cameras, rt)
            if proposals is not None:
                centers = proposals.clone()
            mask = centers[:, :, 3] >= 0
            fused, planes = self.joint_net(meta, heatmaps, centers.detach(), mask, cameras, rt)
        B, P = centers.shape[:2]
        J = heatmaps.shape[2]
        fused = torch.cat([fused, centers[:, :, 3:5].reshape(B, -1, 1, 2).repeat(1, 1, J, 1)], dim=3)
        return fused, planes, centers
