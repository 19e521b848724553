"""torchrun --nproc-per-node N tools/zslab_check.py : z-sharded fused projection of a dense space grid
(BASELINE config 4: 256x256x64 over 8000x8000x2000 mm, Panoptic rig) with the NCCL plane combine,
checked on every rank against the unsharded planes, and timed."""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch, torch.distributed as dist
from faster_voxelpose_b200 import ops, parallel as par, synthetic as syn
from faster_voxelpose_b200.config import make_cfg
from faster_voxelpose_b200.lib.models.project_whole import ProjectLayer

rank, world, local = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
torch.cuda.set_device(local); dev = torch.device("cuda", local)
if world > 1:
    dist.init_process_group("nccl", device_id=dev)
B = int(os.environ.get("ZB", 4))
cfg = make_cfg("panoptic", device=str(dev), CAPTURE_SPEC__VOXELS_PER_AXIS=[256, 256, 64])
cams = syn.load_rig("panoptic")
layer = ProjectLayer(cfg)
hm = torch.from_numpy(syn.gaussian_heatmaps(5, cfg, cams, B)).to(dev)        # same frames on every rank
rt = torch.as_tensor(syn.resize_transform(cfg.DATASET.ORI_IMAGE_SIZE, cfg.DATASET.IMAGE_SIZE), dtype=torch.float32, device=dev)
cameras = {"s": cams}
rigs = layer.rigs(cameras, len(cams), dev); geom = layer.geom(len(cams), rt)
hm_cl = ops.rowpair(hm, cache=False); fr = rigs.frame_rig(["s"] * B)
full = ops.project_planes(hm_cl, rigs.dev, fr, geom, layer.axes(dev), [256, 256, 64], B)
# PEER=pull|multimem|auto (1 = auto): planes staged in symmetric memory and combined by one peer-memory kernel
# instead of NCCL
pm = os.environ.get("PEER", "0")
peer = par.PeerPlanes(B, 15, 256, 256, 64, dev, mode="auto" if pm == "1" else pm) if (pm != "0" and world > 1) else None
def step():
    return par.project_space_planes_zsharded(layer, hm_cl, fr, rigs, geom, peer=peer)
got = step()
ok = all(torch.equal(a, b) for a, b in zip(got, full))
for _ in range(3): ok = ok and all(torch.equal(a, b) for a, b in zip(step(), full))   # both staging slots
if world > 1: dist.barrier()
torch.cuda.synchronize(); s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
s.record(); n = 10
for _ in range(n): step()
e.record(); torch.cuda.synchronize()
ms = torch.tensor([s.elapsed_time(e) / n], device=dev); okt = torch.tensor([int(ok)], device=dev)
if world > 1:
    dist.all_reduce(ms, op=dist.ReduceOp.MAX); dist.all_reduce(okt, op=dist.ReduceOp.MIN)
if rank == 0:
    print(json.dumps({"check": "zslab dense 256x256x64", "combine": f"peer-memory kernel ({peer.mode}, symmetric memory)" if peer else "NCCL all_reduce(MAX) + all_gather", "n_gpus": world, "frames": B, "bit_identical_to_unsharded": bool(okt.item()),
                      "ms_per_step_max_over_ranks": float(ms.item()), "frames_per_s": B / (float(ms.item()) * 1e-3)}))
if world > 1: dist.destroy_process_group()
