"""SimpleUNet for ERA5-like weather grids; same module/parameter names as the reference
(/root/reference/scripts/01_data_parallel_ddp/multinode_ddp_unet.py:171-214; duplicate in
scripts/02_fully_sharded_fsdp/multinode_fsdp_unet.py:69-112): enc1-3, bottleneck, up1-3, dec1-3, out, each block
``Conv3x3 -> BN -> ReLU -> Conv3x3 -> BN -> ReLU`` (Sequential indices 0,1,3,4 carry parameters), bilinear resize to
the skip size so odd grids (lat = 181) survive the 2x pool / 2x transposed-conv round trip."""
from __future__ import annotations

import torch
from torch import nn

from ..ops import conv as cops
from .layers import BatchNorm2d, Conv2d, ConvTranspose2x2, MaxPool2d


class _Block(nn.Sequential):
    """Indices match the reference's nn.Sequential(Conv, BN, ReLU, Conv, BN, ReLU); ReLU is fused into BN."""

    def __init__(self, in_ch, out_ch):
        super().__init__(Conv2d(in_ch, out_ch, 3, padding=1), BatchNorm2d(out_ch), nn.Identity(),
                         Conv2d(out_ch, out_ch, 3, padding=1), BatchNorm2d(out_ch), nn.Identity())

    def forward(self, x):
        x = self[1](self[0](x), relu=True)
        return self[4](self[3](x), relu=True)


class SimpleUNet(nn.Module):
    def __init__(self, in_channels=65, out_channels=65, base_dim=64):
        super().__init__()
        b = base_dim
        self.enc1, self.enc2, self.enc3 = _Block(in_channels, b), _Block(b, b * 2), _Block(b * 2, b * 4)
        self.bottleneck = _Block(b * 4, b * 8)
        self.up3, self.dec3 = ConvTranspose2x2(b * 8, b * 4), _Block(b * 8, b * 4)
        self.up2, self.dec2 = ConvTranspose2x2(b * 4, b * 2), _Block(b * 4, b * 2)
        self.up1, self.dec1 = ConvTranspose2x2(b * 2, b), _Block(b * 2, b)
        self.out = Conv2d(b, out_channels, kernel_size=1)
        self.pool = MaxPool2d(2)

    def forward(self, x):
        if x.is_cuda:
            x = x.contiguous(memory_format=torch.channels_last)
        e1 = self.enc1(x)
        e2 = self.enc2(self.pool(e1))
        e3 = self.enc3(self.pool(e2))
        bt = self.bottleneck(self.pool(e3))
        up3 = cops.bilinear_resize(self.up3(bt), e3.shape[2:])
        d3 = self.dec3(torch.cat([up3, e3], dim=1))
        up2 = cops.bilinear_resize(self.up2(d3), e2.shape[2:])
        d2 = self.dec2(torch.cat([up2, e2], dim=1))
        up1 = cops.bilinear_resize(self.up1(d2), e1.shape[2:])
        d1 = self.dec1(torch.cat([up1, e1], dim=1))
        return self.out(d1)


latitude_weighted_mse = cops.latitude_weighted_mse
