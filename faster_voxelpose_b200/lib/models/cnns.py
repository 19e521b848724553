"""The small conv nets that sit between the projection kernels.

BASELINE.json's north_star keeps these as plain PyTorch modules ("The HDN/JLN 2D/1D
convolutions stay as the reference's PyTorch modules"); they are re-stated here only so the
package runs where the reference tree is absent (the GPU box).  One dimension-generic
implementation replaces the reference's duplicated 1-D / 2-D files:

* ``CenterNet`` / ``P2PNet``  -- reference ``lib/models/cnns_2d.py:187-304``
* ``C2CNet``                  -- reference ``lib/models/cnns_1d.py:165-241``
* ``WeightNet``               -- reference ``lib/models/weight_net.py:36-88``

Attribute names (and therefore ``state_dict`` keys) equal the reference's, so published
checkpoints load unchanged (``tests/test_host_cpu.py::test_conv_nets_match_reference`` checks key-for-key
equality and numerics against the golden vectors).
"""
from __future__ import annotations

import torch
import torch.nn as nn
import torch.nn.functional as F

_CONV = {1: nn.Conv1d, 2: nn.Conv2d}
_DECONV = {1: nn.ConvTranspose1d, 2: nn.ConvTranspose2d}
_BN = {1: nn.BatchNorm1d, 2: nn.BatchNorm2d}
_POOL = {1: F.max_pool1d, 2: F.max_pool2d}


class _ConvBnRelu(nn.Module):
    def __init__(self, dim, cin, cout, k):
        super().__init__()
        self.block = nn.Sequential(_CONV[dim](cin, cout, kernel_size=k, stride=1, padding=(k - 1) // 2),
                                   _BN[dim](cout), nn.ReLU(True))

    def forward(self, x):
        return self.block(x)


class _Residual(nn.Module):
    def __init__(self, dim, cin, cout):
        super().__init__()
        conv, bn = _CONV[dim], _BN[dim]
        self.res_branch = nn.Sequential(conv(cin, cout, kernel_size=3, stride=1, padding=1), bn(cout), nn.ReLU(True),
                                        conv(cout, cout, kernel_size=3, stride=1, padding=1), bn(cout))
        self.skip_con = nn.Sequential() if cin == cout else \
            nn.Sequential(conv(cin, cout, kernel_size=1, stride=1, padding=0), bn(cout))

    def forward(self, x):
        return F.relu(self.res_branch(x) + self.skip_con(x), True)


class _Up2(nn.Module):
    def __init__(self, dim, cin, cout):
        super().__init__()
        self.block = nn.Sequential(_DECONV[dim](cin, cout, kernel_size=2, stride=2, padding=0, output_padding=0),
                                   _BN[dim](cout), nn.ReLU(True))

    def forward(self, x):
        return self.block(x)


class _Hourglass(nn.Module):
    """Two-level encoder/decoder with residual skips, 32 -> 64 -> 128 channels."""

    def __init__(self, dim):
        super().__init__()
        self._pool = _POOL[dim]
        self.encoder_res1 = _Residual(dim, 32, 64)
        self.encoder_res2 = _Residual(dim, 64, 128)
        self.mid_res = _Residual(dim, 128, 128)
        self.decoder_res2 = _Residual(dim, 128, 128)
        self.decoder_upsample2 = _Up2(dim, 128, 64)
        self.decoder_res1 = _Residual(dim, 64, 64)
        self.decoder_upsample1 = _Up2(dim, 64, 32)
        self.skip_res1 = _Residual(dim, 32, 32)
        self.skip_res2 = _Residual(dim, 64, 64)

    def forward(self, x):
        s1 = self.skip_res1(x)
        x = self.encoder_res1(self._pool(x, 2, 2))
        s2 = self.skip_res2(x)
        x = self.encoder_res2(self._pool(x, 2, 2))
        x = self.decoder_res2(self.mid_res(x))
        x = self.decoder_res1(self.decoder_upsample2(x) + s2)
        return self.decoder_upsample1(x) + s1


def _init_small(module, kinds):
    for m in module.modules():
        if isinstance(m, kinds):
            nn.init.normal_(m.weight, 0, 0.001)
            nn.init.constant_(m.bias, 0)


class _Trunk(nn.Module):
    def __init__(self, dim, cin):
        super().__init__()
        self.front_layers = nn.Sequential(_ConvBnRelu(dim, cin, 16, 7), _Residual(dim, 16, 32))
        self.encoder_decoder = _Hourglass(dim)

    def features(self, x):
        cin = self.front_layers[0].block[0].in_channels
        if x.shape[1] > cin:          # channel-padded input from the projection kernel (pad = 0)
            x = x[:, :cin]
        return self.encoder_decoder(self.front_layers(x))


class P2PNet(_Trunk):
    def __init__(self, input_channels, output_channels):
        super().__init__(2, input_channels)
        self.output_channels = output_channels
        self.output_layer = nn.Conv2d(32, output_channels, kernel_size=1, stride=1, padding=0)
        _init_small(self, (nn.Conv2d, nn.ConvTranspose2d))

    def forward(self, x):
        if _use_fused(self, x):
            def build(m):
                w, b = _fold(m.output_layer)
                return (_FusedTrunk(m), (w, b), w.reshape(w.shape[0], -1).cpu().numpy().copy(), b.cpu().numpy().copy())
            trunk, out_wb, w_host, b_host = _fused_state(self, build)
            if w_host.shape == (w_host.shape[0], 32) and w_host.shape[0] <= 20 and x.shape[-1] % 4 == 0 \
                    and x.shape[-2] % 4 == 0:
                # last up-sampling + skip add + 1x1 output layer in one kernel, planar maps out
                from ... import ops
                up, skip = trunk(x, split_last=True, out_channels=int(w_host.shape[0]))
                if isinstance(skip, dict):      # the block's second convolution is still to do: one tcgen05 kernel for the tail
                    return ops.res_tail_tf32(skip["h"], skip["x"], up, skip["w2"], skip["b2"], w_host, b_host)
                if isinstance(skip, tuple):     # (conv without epilogue, block input, bias): the block is finished in the kernel
                    return ops.pointwise_planar(skip[0], up, w_host, b_host, resid=skip[1], resid_bias=skip[2])
                return ops.pointwise_planar(skip, up, w_host, b_host)
            return _conv(trunk(x), out_wb, 0).contiguous()
        return self.output_layer(self.features(x))


class CenterNet(_Trunk):
    """``forward`` accepts either the space cube ``[B,J,X,Y,Z]`` (the reference's contract,
    z-max taken here) or the already reduced xy max-plane ``[B,J,X,Y]`` that the fused
    projection kernel emits."""

    def __init__(self, input_channels, output_channels, head_conv=32):
        super().__init__(2, input_channels)
        self.output_channels = output_channels
        self.output_hm = nn.Sequential(nn.Conv2d(32, head_conv, kernel_size=3, padding=1), nn.ReLU(inplace=True),
                                       nn.Conv2d(head_conv, output_channels, kernel_size=1, padding=0))
        self.output_size = nn.Sequential(nn.Conv2d(32, head_conv, kernel_size=3, padding=1, bias=True),
                                         nn.ReLU(inplace=True),
                                         nn.Conv2d(head_conv, 2, kernel_size=1, padding=0, bias=True))
        _init_small(self, (nn.Conv2d, nn.ConvTranspose2d))

    def forward(self, x):
        if x.dim() == 5:
            x, _ = torch.max(x, dim=4)
        if _use_fused(self, x):
            # the two 3x3 heads share their input: one conv+bias+ReLU with concatenated filters
            def build(m):
                (w1, b1), (w2, b2) = _fold(m.output_hm[0]), _fold(m.output_size[0])
                heads = (torch.cat([w1, w2]).contiguous(memory_format=torch.channels_last), torch.cat([b1, b2]))
                # ... and the two 1x1 output layers are one 1x1 conv with block-diagonal filters over the concatenated
                # activation (slicing the channels-last tensor per head cost two layout copies and a second launch)
                (wh, bh), (ws, bs) = _fold(m.output_hm[2]), _fold(m.output_size[2])
                c1, c2 = w1.shape[0], w2.shape[0]
                wo = torch.zeros((wh.shape[0] + ws.shape[0], c1 + c2, 1, 1), dtype=wh.dtype, device=wh.device)
                wo[:wh.shape[0], :c1] = wh
                wo[wh.shape[0]:, c1:] = ws
                outs = (wo.contiguous(memory_format=torch.channels_last), torch.cat([bh, bs]))
                return _FusedTrunk(m), heads, wh.shape[0], outs
            trunk, heads, n_hm, outs = _fused_state(self, build)
            o = _conv(_conv_relu(trunk(x), heads, 1), outs, 0)
            return o[:, :n_hm].contiguous(), o[:, n_hm:].contiguous()
        x = self.features(x)
        return self.output_hm(x), self.output_size(x)


class C2CNet(_Trunk):
    def __init__(self, input_channels, output_channels, head_conv=32):
        super().__init__(1, input_channels)
        self.output_channels = output_channels
        self.output_hm = nn.Conv1d(32, output_channels, kernel_size=1, stride=1, padding=0)
        _init_small(self, (nn.Conv1d, nn.ConvTranspose1d))

    def forward(self, x):
        if _use_fused(self, x) and x.dim() == 3 and x.shape[-1] == 20 and x.shape[1] <= 32 \
                and self.output_hm.out_channels == 1:
            from ... import ops
            weights, table = _fused_state(self, _c2c_program)
            return ops.c2c_forward(x, weights, table)
        return self.output_hm(self.features(x))


class WeightNet(nn.Module):
    """One confidence weight per (plane, person, joint) from the joint's 2-D feature map."""

    def __init__(self, cfg):
        super().__init__()
        self.voxels_per_axis = cfg.INDIVIDUAL_SPEC.VOXELS_PER_AXIS
        self.num_joints = cfg.DATASET.NUM_JOINTS
        feat, hidden = cfg.NETWORK.NUM_CHANNEL_JOINT_FEAT, cfg.NETWORK.NUM_CHANNEL_JOINT_HIDDEN
        self.heatmap_feature_net = nn.Sequential(nn.Conv2d(1, feat, 3, stride=1, padding=1), nn.BatchNorm2d(feat),
                                                 nn.MaxPool2d(2), nn.ReLU(inplace=True))
        self.output = nn.Sequential(nn.Linear(feat, hidden), nn.ReLU(inplace=True), nn.Linear(hidden, 1),
                                    nn.Sigmoid())

    def forward(self, x):
        x = torch.flatten(x, 0, 1)
        n = x.shape[0]
        x = x.reshape(n * self.num_joints, 1, self.voxels_per_axis[0], self.voxels_per_axis[1])
        if _use_fused(self, x):
            # conv + folded BN + max-pool + ReLU + global mean in one CUDA kernel (the [M,F,n,n]
            # activation never exists); the MLP head stays torch
            def build(m):
                w, b = _fold(m.heatmap_feature_net[0], m.heatmap_feature_net[1])
                return (w.reshape(w.shape[0], 9).cpu().numpy().copy(), b.cpu().numpy().copy())
            w_host, b_host = _fused_state(self, build)
            if w_host.shape[0] in (16, 32, 64) and x.shape[-1] == x.shape[-2]:
                from ... import ops
                tf32 = TF32_WEIGHTNET and torch.backends.cudnn.allow_tf32 and x.shape[-1] % 8 == 0 and x.shape[-1] <= 160
                feats = ops.weightnet_features(x.reshape(-1, x.shape[-2], x.shape[-1]), w_host, b_host, tf32=tf32)
                return self.output(feats).reshape(n, self.num_joints, 1)
        x = F.adaptive_avg_pool2d(self.heatmap_feature_net(x), 1).reshape(n * self.num_joints, -1)
        return self.output(x).reshape(n, self.num_joints, 1)


# =================================================================================================
# Inference path of the same modules: BatchNorm folded into the preceding
# conv, channels-last activations, cuDNN's fused conv+bias+ReLU / conv+add+bias+ReLU kernels.  The
# modules, their parameters and state_dict keys are untouched; eval-mode CUDA forwards of the 2-D
# nets route through here when ``FUSED_INFERENCE`` is on (training and CPU use the plain Sequential
# path above).  In eval mode BN is the affine map y = (x - mean) * gamma / sqrt(var + eps) + beta, so
# folding changes results only by fp32 rounding of the weights.
# =================================================================================================
FUSED_INFERENCE = True
TF32_FRONT_CONV = True   # front 7x7 conv on the hand-written TF32 kernel when torch allows TF32 convolutions
TF32_WEIGHTNET = True    # WeightNet's 3x3 conv on TF32 tensor cores under the same condition
TCGEN05_RES2D = True     # residual blocks with 32 output channels as one tcgen05 kernel (both convs, skip, ReLUs), same condition
TCGEN05_CONV7 = True     # front 7x7 conv on the tcgen05 kernel where the image fits its tiles (else the mma.sync one)
TCGEN05_RES_TAIL = True  # P2PNet: skip_res1's second convolution, add + ReLU, last up-sampling add and output layer in one tcgen05 kernel
DEFER_SKIP_RES1 = True   # P2PNet: skip_res1's add + ReLU are finished by the output kernel (ops.pointwise_planar)
TCGEN05_RES2D_CIN = (16,)   # input channel counts routed to it (32 is supported by the kernel, measured slower than cuDNN)


def _fold(conv, bn=None, transposed=False):
    """(weight, bias) of conv followed by eval-mode bn, weight in channels-last memory format."""
    w = conv.weight.detach().float()
    b = conv.bias.detach().float() if conv.bias is not None else torch.zeros(
        w.shape[1] if transposed else w.shape[0], device=w.device)
    if bn is not None:
        scale = bn.weight.detach().float() / torch.sqrt(bn.running_var.detach().float() + bn.eps)
        shape = (1, -1, 1, 1) if transposed else (-1, 1, 1, 1)
        w = w * scale.view(shape)
        b = (b - bn.running_mean.detach().float()) * scale + bn.bias.detach().float()
    return w.contiguous(memory_format=torch.channels_last), b.contiguous()


def _conv_relu(x, wb, pad):
    return torch.cudnn_convolution_relu(x, wb[0], wb[1], (1, 1), (pad, pad), (1, 1), 1)


def _conv(x, wb, pad):
    return F.conv2d(x, wb[0], wb[1], stride=1, padding=pad)


class _FusedTrunk:
    """Folded weights of one 2-D trunk (front layers + hourglass) and its functional forward."""

    def __init__(self, trunk: "_Trunk"):
        f0, f1 = trunk.front_layers[0], trunk.front_layers[1]
        self.front = _fold(f0.block[0], f0.block[1])
        self.front_pad = f0.block[0].padding[0]
        self._padded = {}
        self._front_frag = None
        self._front_pairs = None
        self.front_res = self._residual(f1)
        h = trunk.encoder_decoder
        self.res = {n: self._residual(getattr(h, n)) for n in
                    ("encoder_res1", "encoder_res2", "mid_res", "decoder_res2", "decoder_res1", "skip_res1", "skip_res2")}
        self.up2 = self._upsample(h.decoder_upsample2)
        self.up1 = self._upsample(h.decoder_upsample1)

    @staticmethod
    def _residual(r: _Residual):
        br = r.res_branch
        c1, c2 = _fold(br[0], br[1]), _fold(br[3], br[4])
        skip = None
        if len(r.skip_con):
            # the 1x1 skip conv runs without bias; its bias joins the second conv's (one add kernel less)
            skip, skip_bias = _fold(r.skip_con[0], r.skip_con[1])
            c2 = (c2[0], (c2[1] + skip_bias).contiguous())
        return [c1, c2, skip]      # + the packed filters of the tcgen05 kernel, added on first use

    @staticmethod
    def _run_residual(x, w, defer=False, out_channels=None):
        """``defer`` (identity-skip blocks whose only consumer can finish them): return (second convolution without bias,
        add or ReLU; the block's input; the bias as a host array) instead of the block's output -- or, when the consumer is
        the tcgen05 tail kernel (``out_channels`` = the channels of its 1x1 layer), the operands of that kernel."""
        c1, c2, skip = w[:3]
        if TCGEN05_RES2D and torch.backends.cudnn.allow_tf32 and c2[0].shape[0] == 32 and c1[0].shape[1] in TCGEN05_RES2D_CIN:
            # 16 -> 32 channels (the front block): cuDNN needs 3 launches that are HBM-bound on the block's intermediate
            # tensors (0.92 ms on 960 planes of 64 x 64); one tcgen05 kernel keeps them on the SM (csrc/fvp_res2d.cu,
            # 0.54 ms).  The 32 -> 32 block (skip_res1) has no room to double-buffer its tile and stays on cuDNN.
            from ... import ops
            if ops.res2d_supported(int(c1[0].shape[1]), 32, int(x.shape[-2]), int(x.shape[-1]), skip is not None):
                if len(w) < 4:
                    w.append(None)
                if w[3] is None:
                    w[3] = ops.res2d_pack(c1, c2, skip)
                return ops.res2d_block_tf32(x, w[3])
        h = _conv_relu(x, c1, 1)
        if defer and skip is None and DEFER_SKIP_RES1:
            # cuDNN's fused conv + add + ReLU costs 0.44 ms on 960 planes of 64 x 64 x 32, the bare convolution 0.20 ms
            if len(w) < 5:
                w.extend([None] * (5 - len(w)))
                w[4] = c2[1].detach().cpu().numpy().copy()
            from ... import ops
            if TCGEN05_RES_TAIL and torch.backends.cudnn.allow_tf32 and tuple(c2[0].shape) == (32, 32, 3, 3) \
                    and out_channels is not None and ops.res_tail_supported(int(x.shape[-2]), int(x.shape[-1]), out_channels):
                # second convolution, skip add, ReLU, last up-sampling add and the 1x1 output layer in one tcgen05 kernel
                # (csrc/fvp_restail.cu): hand over what it needs
                if len(w) < 6:
                    w.append(ops.res_tail_pack(c2[0]))
                return {"h": h, "x": x, "w2": w[5], "b2": w[4]}
            return (F.conv2d(h, c2[0], None, stride=1, padding=1), x, w[4])
        z = x if skip is None else F.conv2d(x, skip)
        # relu(conv(h) + 1.0 * z + bias): the block's add + ReLU ride in the conv epilogue
        return torch.cudnn_convolution_add_relu(h, c2[0], z, 1.0, c2[1], (1, 1), (1, 1), (1, 1), 1)

    @staticmethod
    def _upsample(up: _Up2):
        """ConvTranspose(k=2, s=2) + BN as a 1x1 conv to 4*Cout channels ordered (a, b, cout): non-overlapping
        taps, so out[2h+a, 2w+b] depends on in[h, w] alone; ``ops.upshuffle_add`` un-shuffles it."""
        w, b = _fold(up.block[0], up.block[1], transposed=True)          # [Cin, Cout, 2, 2]
        assert tuple(up.block[0].kernel_size) == (2, 2) and tuple(up.block[0].stride) == (2, 2)
        w1 = w.permute(2, 3, 1, 0).reshape(4 * w.shape[1], w.shape[0], 1, 1)
        return w1.contiguous(memory_format=torch.channels_last), b.repeat(4).contiguous()

    def _front_weights(self, cin):
        """Front conv filters zero-padded to ``cin`` input channels: the projection kernel hands over
        planes whose joints are padded to a multiple of 4 (tensor-core friendly, pad = 0)."""
        w, b = self.front
        if cin == w.shape[1]:
            return self.front
        if cin not in self._padded:
            wp = torch.zeros((w.shape[0], cin, w.shape[2], w.shape[3]), dtype=w.dtype, device=w.device)
            wp[:, :w.shape[1]] = w
            self._padded[cin] = (wp.contiguous(memory_format=torch.channels_last), b)
        return self._padded[cin]

    def __call__(self, x, split_last=False, out_channels=None):
        """``split_last``: return (un-shuffled last up-sampling [N,4*32,H/2,W/2], skip [N,32,H,W]) instead of
        their sum, for a consumer that fuses the shuffle + add (``ops.pointwise_planar``); with ``out_channels`` (of the
        consumer's 1x1 layer) the skip may come back as the operands of ``ops.res_tail_tf32`` instead."""
        from ... import ops
        x = x.contiguous(memory_format=torch.channels_last)
        fw, fb = self._front_weights(x.shape[1])
        if TF32_FRONT_CONV and torch.backends.cudnn.allow_tf32 and tuple(fw.shape) in ((16, 16, 7, 7), (16, 20, 7, 7)) and self.front_pad == 3:
            # 7x7, 16 (or 20: 17-joint rigs) -> 16 channels: hand-written TF32 tensor-core kernel (cuDNN's best engine is 3x slower);
            # only where torch itself would use TF32 convolutions
            if TCGEN05_CONV7 and ops.conv7x7_tc_supported(int(fw.shape[1]), int(x.shape[-2]), int(x.shape[-1])):
                # tcgen05 form (tap pairs in the two halves of an M = 32 accumulator): csrc/fvp_conv7tc.cu
                if self._front_pairs is None or self._front_pairs[0] != fw.shape[1]:
                    self._front_pairs = (fw.shape[1], ops.conv7_pairs(fw))
                x = ops.conv7x7_tc_tf32(x, self._front_pairs[1], fb, relu=True)
            else:
                if self._front_frag is None or self._front_frag[0] != fw.shape[1]:
                    self._front_frag = (fw.shape[1], ops.conv7_fragments(fw))
                x = ops.conv7x7_c16_tf32(x, self._front_frag[1], fb, relu=True)
        else:
            x = _conv_relu(x, (fw, fb), self.front_pad)
        x = self._run_residual(x, self.front_res)
        r = self.res
        s1 = self._run_residual(x, r["skip_res1"], defer=split_last, out_channels=out_channels)
        x = self._run_residual(ops.maxpool2_nhwc(x), r["encoder_res1"])
        s2 = self._run_residual(x, r["skip_res2"])
        x = self._run_residual(ops.maxpool2_nhwc(x), r["encoder_res2"])
        x = self._run_residual(self._run_residual(x, r["mid_res"]), r["decoder_res2"])
        x = self._run_residual(ops.upshuffle_add(_conv_relu(x, self.up2, 0), s2), r["decoder_res1"])
        up = _conv_relu(x, self.up1, 0)
        return (up, s1) if split_last else ops.upshuffle_add(up, s1)


def _fold1d(conv, bn=None, transposed=False):
    """1-D ``_fold``: (weight, bias) with eval-mode BatchNorm folded in; weight keeps its torch layout."""
    w = conv.weight.detach().float()
    nout = w.shape[1] if transposed else w.shape[0]
    b = conv.bias.detach().float() if conv.bias is not None else torch.zeros(nout, device=w.device)
    if bn is not None:
        scale = bn.weight.detach().float() / torch.sqrt(bn.running_var.detach().float() + bn.eps)
        w = w * scale.view((1, -1, 1) if transposed else (-1, 1, 1))
        b = (b - bn.running_mean.detach().float()) * scale + bn.bias.detach().float()
    return w, b


def _c2c_program(net: "C2CNet"):
    """Flatten C2CNet into the op list + flat weight buffer of ``fvp_c2c_forward`` (one kernel for the net).

    Activation buffers: 0-3 hold 32 rows x 20, 4-7 hold 64 x 10, 8-11 hold 128 x 5; the input columns arrive
    in buffer 0.  Conv filters are stored [cin][k][cout], transposed-conv filters [cin][2][cout]."""
    import numpy as np
    CONV, POOL, UP, OUT = 0, 1, 2, 3
    chunks, ops, size = [], [], [0]

    def put(t):
        off = size[0]
        chunks.append(t.reshape(-1))
        size[0] += t.numel()
        return off

    def conv(m, bn, src, dst, add=-1, relu=True):
        w, b = _fold1d(m, bn)                                   # [cout, cin, k]
        cout, cin, k = w.shape
        assert cout & (cout - 1) == 0, "channel counts must be powers of two"
        ops.append([CONV, src, dst, add, cin, cout.bit_length() - 1, k, int(relu), put(w.permute(1, 2, 0).contiguous()), put(b)])

    def residual(r, src, tmp_h, tmp_z, dst):
        br = r.res_branch
        conv(br[0], br[1], src, tmp_h)
        skip = src
        if len(r.skip_con):
            conv(r.skip_con[0], r.skip_con[1], src, tmp_z, relu=False)
            skip = tmp_z
        conv(br[3], br[4], tmp_h, dst, add=skip)

    def up(u, src, dst, add):
        w, b = _fold1d(u.block[0], u.block[1], transposed=True)     # [cin, cout, 2]
        cin, cout, k = w.shape
        assert k == 2 and tuple(u.block[0].stride) == (2,)
        ops.append([UP, src, dst, add, cin, cout.bit_length() - 1, 2, 1, put(w.permute(0, 2, 1).contiguous()), put(b)])

    A, B, C = (0, 1, 2, 3), (4, 5, 6, 7), (8, 9, 10, 11)
    f0, f1 = net.front_layers[0], net.front_layers[1]
    h = net.encoder_decoder
    conv(f0.block[0], f0.block[1], A[0], A[1])
    residual(f1, A[1], A[2], A[0], A[3])                    # x  -> A3
    residual(h.skip_res1, A[3], A[2], A[0], A[1])           # s1 -> A1
    ops.append([POOL, A[3], B[0], -1, 32, 0, 2, 0, 0, 0])
    residual(h.encoder_res1, B[0], B[1], B[2], B[3])        # x  -> B3
    residual(h.skip_res2, B[3], B[1], B[0], B[2])           # s2 -> B2
    ops.append([POOL, B[3], C[0], -1, 64, 0, 2, 0, 0, 0])
    residual(h.encoder_res2, C[0], C[1], C[2], C[3])        # x  -> C3
    residual(h.mid_res, C[3], C[1], C[0], C[2])             # x  -> C2
    residual(h.decoder_res2, C[2], C[1], C[0], C[3])        # x  -> C3
    up(h.decoder_upsample2, C[3], B[0], B[2])               # relu(up) + s2 -> B0
    residual(h.decoder_res1, B[0], B[1], B[2], B[3])        # x  -> B3
    up(h.decoder_upsample1, B[3], A[0], A[1])               # relu(up) + s1 -> A0
    w, b = _fold1d(net.output_hm)
    ops.append([OUT, A[0], 0, -1, w.shape[1], 0, 1, 0, put(w.reshape(-1)), put(b)])
    return torch.cat(chunks).contiguous(), np.ascontiguousarray(np.asarray(ops, dtype=np.int32))


def _param_key(module):
    return tuple((t.data_ptr(), t._version) for t in list(module.parameters()) + list(module.buffers()))


def _fused_state(module, build):
    """Folded weights cached on the module, rebuilt when any parameter/buffer changed or moved."""
    key = _param_key(module)
    st = module.__dict__.get("_fused_state")
    if st is None or st[0] != key:
        with torch.no_grad():
            st = (key, build(module))
        module.__dict__["_fused_state"] = st
    return st[1]


def _use_fused(module, x):
    return FUSED_INFERENCE and not module.training and x.is_cuda and x.dtype == torch.float32 and \
        not torch.is_grad_enabled()
