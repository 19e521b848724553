"""Print the memory format of every intermediate of the fused P2PNet trunk (GPU only)."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch, torch.nn.functional as F
from faster_voxelpose_b200.lib.models import cnns

def fmt(name, t):
    cl = t.is_contiguous(memory_format=torch.channels_last); nc = t.is_contiguous()
    print(f"{name:28s} {tuple(t.shape)} channels_last={cl} nchw={nc} strides={t.stride()}")

net = cnns.P2PNet(15, 15).eval().cuda()
x = torch.rand(64, 15, 64, 64, device="cuda")
with torch.no_grad():
    st = cnns._FusedTrunk(net)
    x = x.contiguous(memory_format=torch.channels_last); fmt("input", x)
    y = cnns._conv_relu(x, st.front, st.front_pad); fmt("front conv_relu", y)
    c1, c2, skip = st.front_res
    h = cnns._conv_relu(y, c1, 1); fmt("res conv_relu", h)
    z = cnns._conv(y, skip, 0); fmt("skip conv2d", z)
    o = torch.cudnn_convolution_add_relu(h, c2[0], z, 1.0, c2[1], (1, 1), (1, 1), (1, 1), 1); fmt("conv_add_relu", o)
    p = F.max_pool2d(o, 2, 2); fmt("max_pool", p)
    e1 = st._run_residual(p, st.res["encoder_res1"]); fmt("encoder_res1", e1)
    e2 = st._run_residual(F.max_pool2d(e1, 2, 2), st.res["encoder_res2"]); fmt("encoder_res2", e2)
    u = F.conv_transpose2d(e2, st.up2[0], st.up2[1], stride=2); fmt("conv_transpose2d", u)
    fmt("weight front", st.front[0]); fmt("weight up2", st.up2[0])
    import time
    for name, fn in (("conv_relu 32->32 3x3 @64", lambda: cnns._conv_relu(o, st.res["skip_res1"][0], 1)),
                     ("front 7x7 15->16", lambda: cnns._conv_relu(x, st.front, st.front_pad)),
                     ("F.conv2d+relu 32->32", lambda: F.relu_(F.conv2d(o, *st.res["skip_res1"][0], padding=1))),
                     ("deconv 128->64", lambda: F.conv_transpose2d(e2, st.up2[0], st.up2[1], stride=2))):
        for _ in range(3): fn()
        torch.cuda.synchronize(); t0 = time.perf_counter()
        for _ in range(10): fn()
        torch.cuda.synchronize(); print(f"{name:28s} {(time.perf_counter()-t0)*100:.3f} ms (B=64)")
