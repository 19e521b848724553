"""The small conv nets that sit between the projection kernels.

BASELINE.json's north_star keeps these as plain PyTorch modules ("The HDN/JLN 2D/1D
convolutions stay as the reference's PyTorch modules"); they are re-stated here only so the
package runs where the reference tree is absent (the GPU box).  One dimension-generic
implementation replaces the reference's duplicated 1-D / 2-D files:

* ``CenterNet`` / ``P2PNet``  -- reference ``lib/models/cnns_2d.py:187-304``
* ``C2CNet``                  -- reference ``lib/models/cnns_1d.py:165-241``
* ``WeightNet``               -- reference ``lib/models/weight_net.py:36-88``

Attribute names (and therefore ``state_dict`` keys) equal the reference's, so published
checkpoints load unchanged (``tests/test_cnns.py`` checks key-for-key equality and numerics
against the golden vectors).
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
        return self.encoder_decoder(self.front_layers(x))


class P2PNet(_Trunk):
    def __init__(self, input_channels, output_channels):
        super().__init__(2, input_channels)
        self.output_channels = output_channels
        self.output_layer = nn.Conv2d(32, output_channels, kernel_size=1, stride=1, padding=0)
        _init_small(self, (nn.Conv2d, nn.ConvTranspose2d))

    def forward(self, x):
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
        x = self.features(x)
        return self.output_hm(x), self.output_size(x)


class C2CNet(_Trunk):
    def __init__(self, input_channels, output_channels, head_conv=32):
        super().__init__(1, input_channels)
        self.output_channels = output_channels
        self.output_hm = nn.Conv1d(32, output_channels, kernel_size=1, stride=1, padding=0)
        _init_small(self, (nn.Conv1d, nn.ConvTranspose1d))

    def forward(self, x):
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
        x = F.adaptive_avg_pool2d(self.heatmap_feature_net(x), 1).reshape(n * self.num_joints, -1)
        return self.output(x).reshape(n, self.num_joints, 1)
