"""torchrun --nproc-per-node N tools/zslab_breakdown.py : where the time of the z-sharded space plane goes
(BASELINE config 4, 256x256x64, 4 frames replicated on every rank).  Per combine mode, CUDA events, max over ranks:
the slab launch alone (memset + row-walking kernel on this rank's z slab), the combine alone, and both.
Evidence for DESIGN.md section 6; not part of the product."""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist

from faster_voxelpose_b200 import ops, parallel as par, synthetic as syn
from faster_voxelpose_b200.config import make_cfg
from faster_voxelpose_b200.lib.models.human_detection_net import HumanDetectionNet

rank, world, local = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
dist.init_process_group("nccl", device_id=dev)
B = int(os.environ.get("ZB", 4))
cfg = make_cfg("dense", device=str(dev), CAPTURE_SPEC__MIN_SCORE=-1.0)
cams = syn.load_rig("panoptic")
torch.manual_seed(0)
net = HumanDetectionNet(cfg).eval().to(dev)
hm = torch.from_numpy(syn.gaussian_heatmaps(5, cfg, cams, B)).to(dev)
rt = torch.as_tensor(syn.resize_transform(cfg.DATASET.ORI_IMAGE_SIZE, cfg.DATASET.IMAGE_SIZE), dtype=torch.float32, device=dev)
pl = net.project_layer
rigs, geom = pl.rigs({"s": cams}, len(cams), dev), pl.geom(len(cams), rt)
fr = rigs.frame_rig(["s"] * B)
hm_rp = ops.rowpair(hm, cache=False)
X, Y, Z = [int(v) for v in pl.voxels_per_axis]
z0, z1 = par.zslab(Z, rank, world)
jp = ops.padded_joints(geom.J)


def timed_max(fn, reps=30):
    for _ in range(5):
        fn()
    dist.barrier()
    torch.cuda.synchronize(dev)
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    s.record()
    for _ in range(reps):
        fn()
    e.record()
    torch.cuda.synchronize(dev)
    t = torch.tensor([s.elapsed_time(e) / reps], device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return round(float(t.item()), 4)


out = {"n_gpus": world, "frames": B, "slab": [z0, z1]}
with torch.no_grad():
    ref = net.space_plane(hm_rp, fr, rigs, geom).clone()
    out["ms_unsharded"] = timed_max(lambda: net.space_plane(hm_rp, fr, rigs, geom))
    part = torch.empty((B, X, Y, jp), dtype=torch.float32, device=dev)
    out["ms_slab_launch_alone"] = timed_max(lambda: pl.project_plane(hm_rp, fr, rigs, geom, zslab=(z0, z1), nhwc=True, out=part))
    out["ms_memset_plane_alone"] = timed_max(lambda: part.zero_())
    for mode in ("nccl", "pull", "multimem"):
        try:
            net.enable_zshard(combine=mode)
            same = bool(torch.equal(net.space_plane(hm_rp, fr, rigs, geom), ref))
            rec = {"ms_total": timed_max(lambda: net.space_plane(hm_rp, fr, rigs, geom)), "bit_identical": same}
            if mode == "nccl":
                buf = part.clone()
                rec["ms_combine_alone"] = timed_max(lambda: dist.all_reduce(buf, op=dist.ReduceOp.MAX))
            else:
                peer = next(p for k, p in net._peer.items() if k[-1] == mode)
                rec["ms_combine_alone"] = timed_max(lambda: peer.combine(0))
                rec["ms_barrier_alone"] = timed_max(lambda: peer.handle.barrier(channel=0), 50)
            out[mode] = rec
        except Exception as exc:  # a mode the box cannot run is reported, not fatal
            out[mode] = {"unavailable": str(exc)[:200]}
    net.disable_zshard()
if rank == 0:
    print(json.dumps(out), flush=True)
dist.destroy_process_group()
