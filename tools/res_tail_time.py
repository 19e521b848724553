"""The tcgen05 tail kernel of P2PNet (csrc/fvp_restail.cu) against what it replaces -- cuDNN's bare 3x3 convolution +
fvp_pointwise_planar_res -- on the shapes of the bench: 960 planes of 64 x 64 x 32, 15 output maps.

    python tools/res_tail_time.py [--planes 960] [--size 64] [--cout 15]
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

PHASES = ["drain:wait_mma", "drain:tmem+stage", "drain:fetch_next", "drain:1x1+store", "issue:wait_raster", "issue:wait_acc",
          "issue:mma", "load:wait_slot", "load:issue", "load:wait_data"]


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
    ap.add_argument("--cout", type=int, default=15)
    args = ap.parse_args()
    dev = torch.device("cuda", 0)
    torch.backends.cudnn.benchmark = True
    gen = torch.Generator().manual_seed(1)
    N, S, C = args.planes, args.size, args.cout
    cl = lambda t: t.to(dev).contiguous(memory_format=torch.channels_last)   # noqa: E731
    w2 = cl(torch.randn((32, 32, 3, 3), generator=gen) * 0.08)
    b2 = (torch.randn((32,), generator=gen) * 0.3).numpy().copy()
    wo = (torch.randn((C, 32), generator=gen) * 0.2).numpy().copy()
    bo = (torch.randn((C,), generator=gen) * 0.3).numpy().copy()
    h = cl(torch.randn((N, 32, S, S), generator=gen).relu())
    x = cl(torch.randn((N, 32, S, S), generator=gen))
    up = cl(torch.randn((N, 128, S // 2, S // 2), generator=gen))
    packed = ops.res_tail_pack(w2)

    def old():
        return ops.pointwise_planar(F.conv2d(h, w2, None, padding=1), up, wo, bo, resid=x, resid_bias=b2)

    ms_old = timed(old)
    ms_conv = timed(lambda: F.conv2d(h, w2, None, padding=1))
    ms_new = timed(lambda: ops.res_tail_tf32(h, x, up, packed, b2, wo, bo))
    diff = float((ops.res_tail_tf32(h, x, up, packed, b2, wo, bo) - old()).abs().max())
    counters = torch.zeros(12, dtype=torch.int64, device=dev)
    L.check(L.load().fvp_res_tail_profile(L.ptr(counters)), "fvp_res_tail_profile")
    ops.res_tail_tf32(h, x, up, packed, b2, wo, bo)
    torch.cuda.synchronize()
    L.check(L.load().fvp_res_tail_profile(None), "fvp_res_tail_profile")
    c = counters.cpu().tolist()
    tiles = max(c[11], 1)
    gb = (3 * N * S * S * 32 + N * S * S * C) * 4 / 1e9
    print(json.dumps({"planes": N, "size": S, "cout": C, "ms_cudnn_conv_plus_glue_kernel": round(ms_old, 4),
                      "ms_cudnn_conv_alone": round(ms_conv, 4), "ms_tcgen05_tail": round(ms_new, 4),
                      "algorithmic_gb": round(gb, 3), "tail_gbs": round(gb / ms_new * 1e3, 1), "max_abs_diff": diff,
                      "cta0_tiles": c[11], "cycles_per_tile": {k: round(v / tiles) for k, v in zip(PHASES, c[:10])}}), flush=True)


if __name__ == "__main__":
    main()
