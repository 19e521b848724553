"""Time the front 7x7 conv: hand-written TF32 kernel vs cuDNN (TF32, benchmark mode) on the Panoptic P2PNet shape."""
import sys
import torch
sys.path.insert(0, "/root/repo")
from faster_voxelpose_b200 import ops

torch.backends.cudnn.benchmark = True
x = torch.randn(960, 16, 64, 64, device="cuda").contiguous(memory_format=torch.channels_last)
w = (torch.randn(16, 16, 7, 7, device="cuda") * 0.05).contiguous(memory_format=torch.channels_last)
b = torch.randn(16, device="cuda")
frag = ops.conv7_fragments(w)


def timeit(fn, n=20):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    a, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(n):
        fn()
    e.record()
    torch.cuda.synchronize()
    return a.elapsed_time(e) / n


print("ours  ", timeit(lambda: ops.conv7x7_c16_tf32(x, frag, b, relu=True)), "ms")
print("cudnn ", timeit(lambda: torch.cudnn_convolution_relu(x, w, b, (1, 1), (3, 3), (1, 1), 1)), "ms")
