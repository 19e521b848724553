"""Probe: does the per-person projection (L1/LSU-bound) overlap with P2PNet's convolutions (tensor/HBM-bound)
when the people of a batch are split into chunks that alternate between two streams?
    python tools/overlap_probe.py [chunks ...]"""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from faster_voxelpose_b200 import ops, synthetic as syn          # noqa: E402
from faster_voxelpose_b200.config import make_cfg                # noqa: E402
from faster_voxelpose_b200.lib.models import faster_voxelpose as m  # noqa: E402

dev = torch.device("cuda", 0)
torch.backends.cudnn.benchmark = True
cfg = make_cfg("panoptic", device=str(dev), CAPTURE_SPEC__MIN_SCORE=-1.0)
cams = syn.load_rig("panoptic")
B, V, J = 32, len(cams), cfg.DATASET.NUM_JOINTS
torch.manual_seed(0)
model = m.get(cfg).eval().to(dev)
jl = model.joint_net
pl = jl.project_layer
hm = torch.from_numpy(syn.gaussian_heatmaps(100, cfg, cams, B)).to(dev)
props = torch.from_numpy(syn.synthetic_proposals(101, cfg, B, bbox="full")).to(dev).view(-1, 7).float().contiguous()
rt = torch.as_tensor(syn.resize_transform(cfg.DATASET.ORI_IMAGE_SIZE, cfg.DATASET.IMAGE_SIZE), dtype=torch.float32, device=dev)
cameras = {"s": cams}
rigs, geom = pl.rigs(cameras, V, dev), pl.geom(V, rt)
frame_rig = rigs.frame_rig(["s"] * B)
hm_cl = ops.rowpair(hm)
frame = torch.arange(B, dtype=torch.int32, device=dev).repeat_interleave(10).contiguous()
N, n, jp = props.shape[0], 64, ops.padded_joints(J)
streams = [torch.cuda.Stream(dev), torch.cuda.Stream(dev)]


def chunked(k):
    outs = []
    step = N // k
    cur = torch.cuda.current_stream(dev)
    for s in streams:
        s.wait_stream(cur)
    for c in range(k):
        s = streams[c % 2]
        with torch.cuda.stream(s), torch.no_grad():
            sl = slice(c * step, (c + 1) * step)
            planes = torch.empty((3, step, n, n, jp), dtype=torch.float32, device=dev)
            pl.project_planes(hm_cl, frame_rig, rigs, geom, props[sl].contiguous(), frame[sl].contiguous(),
                              out=(planes[0], planes[1], planes[2]), nhwc=True)
            outs.append(jl.conv_net(planes.view(3 * step, n, n, jp).permute(0, 3, 1, 2)))
    for s in streams:
        cur.wait_stream(s)
    return outs


def serial(k):
    outs = []
    step = N // k
    with torch.no_grad():
        for c in range(k):
            sl = slice(c * step, (c + 1) * step)
            planes = torch.empty((3, step, n, n, jp), dtype=torch.float32, device=dev)
            pl.project_planes(hm_cl, frame_rig, rigs, geom, props[sl].contiguous(), frame[sl].contiguous(),
                              out=(planes[0], planes[1], planes[2]), nhwc=True)
            outs.append(jl.conv_net(planes.view(3 * step, n, n, jp).permute(0, 3, 1, 2)))
    return outs


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


for k in [int(a) for a in sys.argv[1:]] or [1, 2, 4, 8]:
    print(f"chunks {k}: serial {timed(lambda: serial(k)):.3f} ms, two streams {timed(lambda: chunked(k)):.3f} ms", flush=True)
