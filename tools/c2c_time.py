import torch, time, os, sys
sys.path.insert(0, '/root/repo')
from faster_voxelpose_b200.config import make_cfg
from faster_voxelpose_b200.lib.models import cnns
cfg = make_cfg("panoptic", device="cuda:0")
net = cnns.C2CNet(15, 1).cuda().eval()
x = torch.rand(320, 15, 20, device="cuda")
with torch.no_grad():
    for S in (1, 2, 3, 4):
        os.environ["FVP_C2C_S"] = str(S)
        for _ in range(3): net(x)
        torch.cuda.synchronize()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        for _ in range(20): net(x)
        b.record(); torch.cuda.synchronize()
        print("S", S, a.elapsed_time(b) / 20, "ms")
    cnns.FUSED_INFERENCE = False
    for _ in range(3): net(x)
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(20): net(x)
    b.record(); torch.cuda.synchronize()
    print("torch", a.elapsed_time(b) / 20, "ms")
