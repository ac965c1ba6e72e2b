"""nn.Module wrappers over ``b200dist.ops`` with torch / torchvision-compatible parameter names."""
from __future__ import annotations

import math

import torch
from torch import nn

from .. import ops
from ..ops import conv as cops


class Linear(nn.Linear):
    """nn.Linear parameters (weight, bias) on the tcgen05 GEMM."""

    def forward(self, x, residual=None):
        return ops.linear(x, self.weight, self.bias, residual)


class LayerNorm(nn.LayerNorm):
    def forward(self, x):
        return ops.layer_norm(x, self.weight, self.bias, self.eps)


class GELU(nn.Module):
    def forward(self, x):
        return ops.gelu(x)


class ReLU(nn.Module):
    def __init__(self, inplace: bool = False):
        super().__init__()

    def forward(self, x):
        return ops.relu(x)


class Conv2d(nn.Conv2d):
    """nn.Conv2d parameters; square kernels, symmetric padding, dilation 1, groups 1."""

    def forward(self, x):
        assert self.groups == 1 and self.dilation == (1, 1) and self.stride[0] == self.stride[1] \
            and self.padding[0] == self.padding[1], "b200dist Conv2d supports square stride/padding, groups=1"
        return cops.conv2d(x, self.weight, self.bias, self.stride[0], self.padding[0])


class ConvTranspose2x2(nn.ConvTranspose2d):
    def __init__(self, in_ch, out_ch, kernel_size=2, stride=2):
        assert kernel_size == 2 and stride == 2
        super().__init__(in_ch, out_ch, 2, 2)

    def forward(self, x):
        return cops.conv_transpose2x2(x, self.weight, self.bias)


class BatchNorm2d(nn.BatchNorm2d):
    """nn.BatchNorm2d state (weight, bias, running_mean, running_var, num_batches_tracked); statistics and
    affine parameters stay fp32 even when the model is cast to bf16; optional fused residual + ReLU."""

    def _apply(self, fn, recurse=True):
        super()._apply(fn, recurse)
        for store in (self._parameters, self._buffers):
            for k, t in store.items():
                if t is not None and t.is_floating_point() and t.dtype != torch.float32:
                    t.data = t.data.float()
        return self

    def forward(self, x, residual=None, relu: bool = False):
        if self.training and self.track_running_stats and self.num_batches_tracked is not None:
            self.num_batches_tracked.add_(1)
        return cops.batch_norm(x, self.weight, self.bias, self.running_mean, self.running_var,
                               self.training or not self.track_running_stats, self.momentum or 0.1, self.eps,
                               residual=residual, relu=relu)


class MaxPool2d(nn.Module):
    def __init__(self, kernel_size, stride=None, padding=0):
        super().__init__()
        self.kernel_size, self.stride, self.padding = kernel_size, stride or kernel_size, padding

    def forward(self, x):
        return cops.max_pool2d(x, self.kernel_size, self.stride, self.padding)


class GlobalAvgPool(nn.Module):
    def forward(self, x):
        return cops.global_avg_pool(x)
