"""Where the fused residual-block kernel (csrc/fvp_res2d.cu) spends its time, on the shapes of the bench:
960 planes of 64 x 64 (P2PNet: front block 16 -> 32 with a 1x1 skip, skip_res1 32 -> 32) against the cuDNN launches it
replaces.  Prints CUDA-event times and the per-phase cycle counters of the first CTA (fvp_res2d_profile).

    python tools/res2d_phases.py [--planes 960] [--size 64]
"""
from __future__ import annotations

import argparse
import json
import os
import sys

import torch
import torch.nn.functional as F

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from faster_voxelpose_b200 import _lib as L, ops   # noqa: E402

PHASES = ["fill", "issue1", "wait1", "epilogue1", "issue2", "wait2", "epilogue2"]


def timed(fn, reps=10):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    s.record()
    for _ in range(reps):
        fn()
    e.record()
    torch.cuda.synchronize()
    return s.elapsed_time(e) / reps


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--planes", type=int, default=960)
    ap.add_argument("--size", type=int, default=64)
    args = ap.parse_args()
    dev = torch.device("cuda", 0)
    torch.backends.cudnn.benchmark = True
    gen = torch.Generator().manual_seed(1)
    N, S = args.planes, args.size
    for cin, with_skip in ((16, True), (32, False)):
        w1 = (torch.randn((32, cin, 3, 3), generator=gen) * 0.08).to(dev).contiguous(memory_format=torch.channels_last)
        w2 = (torch.randn((32, 32, 3, 3), generator=gen) * 0.08).to(dev).contiguous(memory_format=torch.channels_last)
        ws = (torch.randn((32, cin, 1, 1), generator=gen) * 0.2).to(dev).contiguous(memory_format=torch.channels_last) if with_skip else None
        b1, b2 = torch.randn((32,), generator=gen).to(dev), torch.randn((32,), generator=gen).to(dev)
        x = torch.randn((N, cin, S, S), generator=gen).to(dev).contiguous(memory_format=torch.channels_last)
        packed = ops.res2d_pack((w1, b1), (w2, b2), ws)

        def library():
            h = torch.cudnn_convolution_relu(x, w1, b1, (1, 1), (1, 1), (1, 1), 1)
            z = x if ws is None else F.conv2d(x, ws)
            return torch.cudnn_convolution_add_relu(h, w2, z, 1.0, b2, (1, 1), (1, 1), (1, 1), 1)

        ms_lib = timed(library)
        ms_ours = timed(lambda: ops.res2d_block_tf32(x, packed))
        counters = torch.zeros(8, dtype=torch.int64, device=dev)
        L.check(L.load().fvp_res2d_profile(L.ptr(counters)), "fvp_res2d_profile")
        ops.res2d_block_tf32(x, packed)
        torch.cuda.synchronize()
        L.check(L.load().fvp_res2d_profile(None), "fvp_res2d_profile")
        c = counters.cpu().tolist()
        tiles = max(c[7], 1)
        diff = float((ops.res2d_block_tf32(x, packed) - library()).abs().max())
        print(json.dumps({"block": f"{cin}->32->32 {'1x1 skip' if with_skip else 'identity skip'}", "planes": N, "size": S,
                          "ms_cudnn_3_launches" if with_skip else "ms_cudnn_2_launches": round(ms_lib, 4),
                          "ms_tcgen05_fused": round(ms_ours, 4), "max_abs_diff_vs_cudnn_tf32": diff,
                          "cta0_tiles": c[7], "cycles_per_tile": {k: round(v / tiles) for k, v in zip(PHASES, c[:7])},
                          "cycles_per_tile_total": round(sum(c[:7]) / tiles)}), flush=True)


if __name__ == "__main__":
    main()
