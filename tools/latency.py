"""Small-batch latency: eager forward vs CUDA-graph replay (runtime.GraphedInference), host heatmaps in, poses out.
    python tools/latency.py [workload] [batch]"""
import json
import sys
import time

import torch

sys.path.insert(0, "/root/repo")
from faster_voxelpose_b200 import synthetic as syn
from faster_voxelpose_b200.config import make_cfg
from faster_voxelpose_b200.lib.models import faster_voxelpose as m
from faster_voxelpose_b200.runtime import GraphedInference

workload = sys.argv[1] if len(sys.argv) > 1 else "panoptic"
B = int(sys.argv[2]) if len(sys.argv) > 2 else 1
dev = "cuda:0"
torch.backends.cudnn.benchmark = True
cfg = make_cfg(workload, device=dev, CAPTURE_SPEC__MIN_SCORE=-1.0)
cams = syn.load_rig(workload)
model = m.get(cfg).eval()
model.load_state_dict(syn.seeded_state_dict(model, seed=0))
model.to(dev)
rt = torch.as_tensor(syn.resize_transform(cfg.DATASET.ORI_IMAGE_SIZE, cfg.DATASET.IMAGE_SIZE), dtype=torch.float32, device=dev)
cameras = {"s": cams}
hm = torch.from_numpy(syn.gaussian_heatmaps(3, cfg, cams, B)).pin_memory()
props = torch.from_numpy(syn.synthetic_proposals(4, cfg, B, bbox="full")).to(dev)
meta = {"seq": ["s"] * B}
host_out = None


def eager():
    global host_out
    with torch.no_grad():
        out = model(meta=meta, input_heatmaps=hm.to(dev, non_blocking=True), cameras=cameras, resize_transform=rt, proposals=props)[0]
    if host_out is None:
        host_out = torch.empty(out.shape, dtype=out.dtype).pin_memory()
    host_out.copy_(out, non_blocking=True)
    torch.cuda.synchronize()


def wall(fn, n=50):
    for _ in range(5):
        fn()
    t0 = time.perf_counter()
    for _ in range(n):
        fn()
    return (time.perf_counter() - t0) / n * 1e3


e = wall(eager)
drv = GraphedInference(model, cameras, rt, dev, hm, proposals=props)
g = wall(lambda: drv(hm))
print(json.dumps({"workload": workload, "batch": B, "eager_ms_per_call": e, "graph_ms_per_call": g,
                  "eager_frames_per_s": B / e * 1e3, "graph_frames_per_s": B / g * 1e3,
                  "what": "wall clock per call, pinned host heatmaps in, pinned host poses out, synchronised"}))
