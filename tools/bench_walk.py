"""A/B of the per-person projection kernels on the bench workload (Panoptic, B frames, 10 x 64^3 cubes, bbox full).

    python tools/bench_walk.py [--variants -1,0,1,2,...] [--families G,N] [--batch 32] [--reps 20]

Every variant's planes are compared bit for bit with the generic brick kernel (variant -1, itself pinned to
the oracle by tests/test_gpu_parity.py); prints one JSON line per (family, variant) with the CUDA-event time,
the live voxel count (sum of prod(end-start) over non-skipped windows) and the on-chip data-path fraction.
Not part of the product; the numbers quoted in DESIGN.md come from here.
"""
from __future__ import annotations

import argparse
import json
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from faster_voxelpose_b200 import _lib as L, ops, synthetic as syn   # noqa: E402
from faster_voxelpose_b200.config import make_cfg                    # noqa: E402
from faster_voxelpose_b200.lib.models.project_individual import ProjectLayer   # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--variants", default="-1,1,2,3,4,5,6,7")
    ap.add_argument("--families", default="G,N")
    ap.add_argument("--workload", default="panoptic")
    ap.add_argument("--batch", type=int, default=32)
    ap.add_argument("--reps", type=int, default=20)
    ap.add_argument("--bbox", default="full")
    ap.add_argument("--no-flush", action="store_true", help="do not flush L2 between repetitions")
    ap.add_argument("--occ", default="1", help="comma list of 0/1: hand the occupancy words to the kernel or not")
    args = ap.parse_args()
    dev = torch.device("cuda", 0)
    cfg = make_cfg(args.workload, device=str(dev))
    cams = syn.dome_rig(31, seed=0) if args.workload == "stress" else syn.load_rig(args.workload)
    V, J, B = len(cams), cfg.DATASET.NUM_JOINTS, args.batch
    W, H = cfg.DATASET.HEATMAP_SIZE
    pl = ProjectLayer(cfg)
    rt = torch.as_tensor(syn.resize_transform(cfg.DATASET.ORI_IMAGE_SIZE, cfg.DATASET.IMAGE_SIZE), dtype=torch.float32, device=dev)
    cameras = {"s": cams}
    rigs, geom = pl.rigs(cameras, V, dev), pl.geom(V, rt)
    frame_rig = rigs.frame_rig(["s"] * B)
    props = torch.from_numpy(syn.synthetic_proposals(101, cfg, B, bbox=args.bbox)).to(dev).reshape(-1, 7).contiguous()
    N = props.shape[0]
    frame = torch.arange(B, dtype=torch.int32, device=dev).repeat_interleave(N // B)
    win, _ = ops.individual_windows(props, None, frame, pl.consts)
    w = win.cpu().numpy()
    live = int(sum(int(np.prod(r[6:9] - r[3:6])) for r in w if r[9] == 0))
    n = int(pl.consts.voxels[0])
    jp = ops.padded_joints(J)
    sms = L.sm_count()
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)

    for fam in args.families.split(","):
        hm_np = syn.gaussian_heatmaps(100, cfg, cams, B) if fam == "G" else syn.noise_heatmaps(100, B, V, J, H, W)
        hm_rp = ops.rowpair(torch.from_numpy(hm_np).to(dev), cache=False)
        occ_words = hm_rp.occ
        nz_frac = float(occ_words[:, 0].sum().item()) / (B * V * 2 * (H // 2 + 1) * W)
        ref = None
        for var, use_occ in [(int(v), int(o)) for v in args.variants.split(",") for o in args.occ.split(",")]:
            L.set_option(L.OPT_WALK_VARIANT, var)
            hm_rp.occ = occ_words if use_occ else None
            planes = torch.empty((3, N, n, n, jp), dtype=torch.float32, device=dev)
            out = (planes[0], planes[1], planes[2])

            def run():
                pl.project_planes(hm_rp, frame_rig, rigs, geom, props, frame, out=out, nhwc=True)
            try:
                for _ in range(3):
                    run()
                torch.cuda.synchronize()
                ms = []
                for _ in range(args.reps):
                    if not args.no_flush:
                        flush.zero_()                   # 256 MB > 126 MB L2
                    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                    s.record(); run(); e.record()
                    torch.cuda.synchronize()
                    ms.append(s.elapsed_time(e))
            except Exception as ex:  # a variant that cannot launch (shared memory, registers) is reported, not fatal
                print(json.dumps({"family": fam, "variant": var, "error": str(ex)[:200]}), flush=True)
                continue
            if ref is None:
                ref = planes.clone()
                same = True
            else:
                same = bool(torch.equal(planes.view(torch.int32), ref.view(torch.int32)))
            med = float(np.median(ms))
            tap_bytes = 4.0 * 4 * jp * V * live
            l1_peak = sms * 128 * 1.965e9
            print(json.dumps({"family": fam, "variant": var, "occ": use_occ, "nonzero_records": round(nz_frac, 4), "ms_median": round(med, 4), "ms_min": round(min(ms), 4),
                              "bit_identical_to_first": same, "cubes": N, "live_voxels": live,
                              "voxels_in_window_frac": round(live / (N * n ** 3), 4),
                              "onchip_frac_upper": round(tap_bytes / (med * 1e-3) / l1_peak, 4),
                              "gvoxel_views_per_s": round(live * V / (med * 1e-3) / 1e9, 2)}), flush=True)
    L.set_option(L.OPT_WALK_VARIANT, 0)


if __name__ == "__main__":
    main()
