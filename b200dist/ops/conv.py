"""Convolution-network ops (conv2d, BatchNorm, pooling, transposed conv, bilinear resize, lat-MSE).

Tensors keep the user-facing NCHW *shape* of the reference models but are stored channels-last
(NHWC in memory), so a 1x1 convolution is a plain tcgen05 GEMM over ``[N*H*W, C]`` and a kxk
convolution is im2col + GEMM, with dgrad = GEMM + col2im and wgrad = GEMM (MN-major operands).
CUDA tensors run the kernels of ``csrc/conv_ops.cu`` + ``csrc/gemm_tcgen05.cu``; CPU tensors use
the PyTorch expressions (unit tests / gloo config 1).

Reference: torchvision ResNets in /root/reference/scripts/main.py:249, UNet blocks in
scripts/01_data_parallel_ddp/multinode_ddp_unet.py:171-229, patch embedding in tensor_parallel_vit.py:82-90.
"""
from __future__ import annotations

import os

import ctypes as C
from typing import Optional

import torch
import torch.nn.functional as F
from torch.autograd import Function

from . import _lib as L
from . import native as N

I, LL, Fl = C.c_int, C.c_longlong, C.c_float
_BF16 = torch.bfloat16


def _nhwc(x: torch.Tensor) -> torch.Tensor:
    """[N,C,H,W] (any memory format) -> contiguous NHWC view [N,H,W,C] of a channels-last tensor."""
    return x.contiguous(memory_format=torch.channels_last).permute(0, 2, 3, 1)


def _nchw_view(y_nhwc: torch.Tensor) -> torch.Tensor:
    return y_nhwc.permute(0, 3, 1, 2)  # logical NCHW, channels-last strides


def _s():
    return L.stream_ptr()


def _conv_geom(H, W, KH, KW, stride, pad):
    return (H + 2 * pad - KH) // stride + 1, (W + 2 * pad - KW) // stride + 1


def _w_matrix(w: torch.Tensor, Cp: int) -> torch.Tensor:
    """[K, C, KH, KW] -> bf16 [K, KH*KW*Cp] in (kh, kw, c) order with channels zero-padded to Cp."""
    K, Cc, KH, KW = w.shape
    wm = w.permute(0, 2, 3, 1)
    if Cp != Cc:
        wm = F.pad(wm, (0, Cp - Cc))
    return wm.reshape(K, KH * KW * Cp).to(_BF16).contiguous()


def _im2col(xh: torch.Tensor, KH, KW, stride, pad, Ho, Wo, Cp) -> torch.Tensor:
    Nn, H, W, Cc = xh.shape
    col = torch.empty((Nn * Ho * Wo, KH * KW * Cp), dtype=xh.dtype, device=xh.device)
    L.call("im2col", L.ptr(xh), L.ptr(col), I(Nn), I(H), I(W), I(Cc), I(Cp), I(KH), I(KW), I(stride), I(pad), I(Ho), I(Wo),
           I(L.dtype_code(xh.dtype)), _s())
    return col


_KEEP_COL_BYTES = int(os.environ.get("B200DIST_CONV_KEEP_COL_MB", "2048")) << 20   # per conv; 0 disables
# implicit-GEMM path (stride-1 K x K convs with C % 64 == 0): forward and input-gradient read the activation through a 4-D
# tensor map, no patch matrix; the weight gradient still builds one (in backward).  B200DIST_CONV_IGEMM=0 disables.
_IGEMM = os.environ.get("B200DIST_CONV_IGEMM", "1") not in ("", "0")
_IGEMM_WGRAD = os.environ.get("B200DIST_CONV_IGEMM_WGRAD", "1") not in ("", "0")   # implicit-GEMM weight gradient


_STEM = os.environ.get("B200DIST_CONV_STEM_IGEMM", "1") not in ("", "0")


def _stem_pad(xh: torch.Tensor, pad: int, Ho: int, Wo: int, KH: int) -> torch.Tensor:
    """[N, H, W, C<=8] -> zero-padded [N, Hp, Wp, 8] bf16: `pad` border rows / columns plus what the 8-wide tap window and
    the last filter row reach (see csrc/gemm_tcgen05.cu, launch_conv_stem)."""
    Nn, H, W, Cc = xh.shape
    Hp = max(H + 2 * pad, 2 * (Ho - 1) + KH + 1)
    Wp = max(W + 2 * pad, 2 * (Wo - 1) + 8)
    xp = torch.zeros((Nn, Hp, Wp, 8), dtype=_BF16, device=xh.device)
    xp[:, pad:pad + H, pad:pad + W, :Cc] = xh
    return xp


def _stem_wmat(w: torch.Tensor) -> torch.Tensor:
    """[K, C, KH, KW<=8] -> bf16 [K, KH * 64] in (kh, kw 0..7, c 0..7) order, zero where kw >= KW or c >= C."""
    K, Cc, KH, KW = w.shape
    wm = torch.zeros((K, KH, 8, 8), dtype=_BF16, device=w.device)
    wm[:, :, :KW, :Cc] = w.permute(0, 2, 3, 1)
    return wm.reshape(K, KH * 64)


class _StemConvFn(Function):
    """Stride-2 KxK convolution on <= 8 input channels (ResNet conv1) as an implicit GEMM; no input gradient."""

    @staticmethod
    def forward(ctx, x, w, pad):
        xh = _nhwc(x if x.dtype == _BF16 else x.to(_BF16))
        Nn, H, W, Cc = xh.shape
        K, _, KH, KW = w.shape
        Ho, Wo = _conv_geom(H, W, KH, KW, 2, pad)
        xp = _stem_pad(xh, pad, Ho, Wo, KH)
        y = N.conv2d_stem(xp, _stem_wmat(w), Ho, Wo, KH)
        ctx.save_for_backward(xp, w)
        ctx.x_dtype = x.dtype
        return _nchw_view(y)

    @staticmethod
    def backward(ctx, dy):
        xp, w = ctx.saved_tensors
        K, Cc, KH, KW = w.shape
        dyh = _nhwc(dy if dy.dtype == _BF16 else dy.to(_BF16)).contiguous()
        dw = None
        if ctx.needs_input_grad[1]:
            dwm = N.conv2d_stem_wgrad(xp, dyh, KH)                               # [K, KH * 64] fp32
            dw = dwm.view(K, KH, 8, 8)[:, :, :KW, :Cc].permute(0, 3, 1, 2).to(w.dtype)
        if ctx.needs_input_grad[0]:
            raise NotImplementedError("stem convolution: input gradients are not implemented (images need none)")
        return None, dw, None


class _Conv2dFn(Function):
    @staticmethod
    def forward(ctx, x, w, bias, stride, pad):
        xh = _nhwc(x if x.dtype == _BF16 else x.to(_BF16))
        Nn, H, W, Cc = xh.shape
        K, _, KH, KW = w.shape
        Ho, Wo = _conv_geom(H, W, KH, KW, stride, pad)
        Cp = (Cc + 7) // 8 * 8
        direct = KH == 1 and KW == 1 and stride == 1 and pad == 0 and Cp == Cc
        out_dtype = torch.float32 if x.dtype == torch.float32 else _BF16
        igemm = (_IGEMM and not direct and stride in (1, 2) and Cc % 64 == 0 and K % 8 == 0 and out_dtype == _BF16
                 and Ho > 0 and Wo > 0)
        if igemm:
            # stride 2 (3x3 s2 and the 1x1 s2 shortcut of a ResNet stage transition): same kernel, the tensor map steps
            # over every second pixel
            y = N.conv2d_igemm(xh.contiguous(), _w_matrix(w, Cc), KH, KW, pad,
                               bias=bias.to(_BF16) if bias is not None else None, stride=stride)
            ctx.save_for_backward(xh, w, None)
            ctx.geom = (stride, pad, Ho, Wo, Cp, direct, x.dtype, bias is not None, K)
            ctx.igemm = True
            return _nchw_view(y)
        ctx.igemm = False
        a = xh.reshape(Nn * H * W, Cc) if direct else _im2col(xh, KH, KW, stride, pad, Ho, Wo, Cp)
        wm = _w_matrix(w, Cp)
        Kp = (K + 7) // 8 * 8
        if Kp != K:  # GEMM N must be a multiple of 8 (e.g. 65 output channels): pad, then slice
            wm = F.pad(wm, (0, 0, 0, Kp - K))
        y = N.gemm(a, wm, bias=(F.pad(bias, (0, Kp - K)) if (bias is not None and Kp != K) else bias), out_dtype=out_dtype)
        if Kp != K:
            y = y[:, :K].contiguous()
        # keep the patch matrix for the weight gradient instead of rebuilding it in backward (im2col is the single most
        # expensive kernel of a ResNet step: 22 %): ~7 GB extra for ResNet-50 at batch 256, which 180 GB of HBM affords
        keep_col = (not direct) and w.requires_grad and a.numel() * 2 <= _KEEP_COL_BYTES
        ctx.save_for_backward(xh, w, a if keep_col else None)
        ctx.geom = (stride, pad, Ho, Wo, Cp, direct, x.dtype, bias is not None, Kp)
        return _nchw_view(y.view(Nn, Ho, Wo, K))

    @staticmethod
    def backward(ctx, dy):
        xh, w, col = ctx.saved_tensors
        stride, pad, Ho, Wo, Cp, direct, x_dtype, has_bias, Kp = ctx.geom
        Nn, H, W, Cc = xh.shape
        K, _, KH, KW = w.shape
        dyh = _nhwc(dy if dy.dtype == _BF16 else dy.to(_BF16)).reshape(Nn * Ho * Wo, K)
        if Kp != K:
            dyh = F.pad(dyh, (0, Kp - K))
        dx = dw = db = None
        wm = _w_matrix(w, Cp)
        if Kp != K:
            wm = F.pad(wm, (0, 0, 0, Kp - K))
        if ctx.needs_input_grad[0] and ctx.igemm and stride == 1 and K % 64 == 0 and Cc % 8 == 0:
            # input gradient = stride-1 conv of dY with the spatially flipped, channel-transposed filter, pad' = K - 1 - pad
            wflip = w.flip(2, 3).permute(1, 2, 3, 0).reshape(Cc, KH * KW * K).to(_BF16).contiguous()
            dxh = N.conv2d_igemm(dyh.view(Nn, Ho, Wo, K), wflip, KH, KW, KH - 1 - pad)
            dx = _nchw_view(dxh).to(x_dtype)
        elif ctx.needs_input_grad[0]:
            dcol = N.gemm(dyh, wm, b_mn=True, out_dtype=_BF16)  # [rows, KH*KW*Cp]
            if direct:
                dxh = dcol.view(Nn, H, W, Cc)
            else:
                dxh = torch.empty((Nn, H, W, Cc), dtype=_BF16, device=dy.device)
                L.call("col2im", L.ptr(dcol), L.ptr(dxh), I(Nn), I(H), I(W), I(Cc), I(Cp), I(KH), I(KW), I(stride), I(pad),
                       I(Ho), I(Wo), I(0), _s())
            dx = _nchw_view(dxh).to(x_dtype)
        if ctx.needs_input_grad[1] and ctx.igemm and _IGEMM_WGRAD and Nn * Ho * Wo >= 128:
            dwm = N.conv2d_igemm_wgrad(xh.contiguous(), dyh.view(Nn, Ho, Wo, K), KH, KW, pad, stride=stride)  # [K, KH*KW*C] fp32
            dw = dwm.view(K, KH, KW, Cc).permute(0, 3, 1, 2).to(w.dtype)
        elif ctx.needs_input_grad[1]:
            a = xh.reshape(Nn * H * W, Cc) if direct else (col if col is not None else
                                                            _im2col(xh, KH, KW, stride, pad, Ho, Wo, Cp))
            dwm = N.gemm(dyh, a, a_mn=True, b_mn=True, out_dtype=torch.float32)[:K]  # [K, KH*KW*Cp]
            dw = dwm.view(K, KH, KW, Cp)[..., :Cc].permute(0, 3, 1, 2).to(w.dtype)
        if has_bias and ctx.needs_input_grad[2]:
            db = dy.float().sum((0, 2, 3))
        return dx, dw, db, None, None


def conv2d(x, weight, bias=None, stride: int = 1, padding: int = 0):
    if not x.is_cuda:
        return F.conv2d(x, weight, bias, stride, padding)
    K, Cc, KH, KW = weight.shape
    if (_STEM and int(stride) == 2 and Cc <= 8 and KH == KW and 3 <= KW <= 8 and K <= 64 and K % 8 == 0 and bias is None
            and not x.requires_grad and x.dtype in (_BF16, torch.float32) and x.shape[-1] // 2 <= 256
            and x.shape[0] * (x.shape[-2] // 2) * (x.shape[-1] // 2) >= 128):
        y = _StemConvFn.apply(x, weight, int(padding))
        return y if x.dtype == _BF16 else y.float()
    return _Conv2dFn.apply(x, weight, bias, int(stride), int(padding))


# ------------------------------------------------------------------------------ batch norm
class _BatchNormFn(Function):
    @staticmethod
    def forward(ctx, x, gamma, beta, running_mean, running_var, residual, training, momentum, eps, relu):
        xh = _nhwc(x)
        Nn, H, W, Cc = xh.shape
        rows = Nn * H * W
        dt = L.dtype_code(xh.dtype)
        g32, b32 = gamma.float().contiguous(), beta.float().contiguous()
        mean = torch.empty(Cc, dtype=torch.float32, device=x.device)
        rstd = torch.empty_like(mean)
        if training:
            sums = torch.zeros(2 * Cc, dtype=torch.float32, device=x.device)
            L.call("col_reduce", L.ptr(xh), L.ptr(None), L.ptr(None), L.ptr(None), L.ptr(None), L.ptr(sums), LL(rows), I(Cc),
                   I(0), I(0), I(dt), _s())
            L.call("bn_finalize", L.ptr(sums), L.ptr(mean), L.ptr(rstd), L.ptr(running_mean), L.ptr(running_var), LL(rows),
                   I(Cc), Fl(eps), Fl(momentum), _s())
        else:
            mean.copy_(running_mean)
            rstd.copy_(torch.rsqrt(running_var.float() + eps))
        rh = _nhwc(residual.to(xh.dtype)) if residual is not None else None
        yh = torch.empty_like(xh)
        L.call("bn_apply", L.ptr(xh), L.ptr(mean), L.ptr(rstd), L.ptr(g32), L.ptr(b32), L.ptr(rh), L.ptr(yh), LL(rows), I(Cc),
               I(int(relu)), I(dt), _s())
        ctx.save_for_backward(xh, yh if relu else None, mean, rstd, g32)
        ctx.meta = (relu, residual is not None, gamma.dtype, training)
        return _nchw_view(yh)

    @staticmethod
    def backward(ctx, dy):
        xh, yh, mean, rstd, g32 = ctx.saved_tensors
        relu, has_res, gdt, training = ctx.meta
        Nn, H, W, Cc = xh.shape
        rows = Nn * H * W
        dt = L.dtype_code(xh.dtype)
        dyh = _nhwc(dy.to(xh.dtype))
        sums = torch.zeros(2 * Cc, dtype=torch.float32, device=dy.device)
        L.call("col_reduce", L.ptr(dyh), L.ptr(xh), L.ptr(yh), L.ptr(mean), L.ptr(rstd), L.ptr(sums), LL(rows), I(Cc), I(1),
               I(int(relu)), I(dt), _s())
        dgamma, dbeta = sums[Cc:].clone(), sums[:Cc].clone()
        if not training:  # eval mode: statistics are constants
            sums = torch.zeros_like(sums)
        dxh = torch.empty_like(xh)
        dres = torch.empty_like(xh) if has_res else None
        L.call("bn_bwd_apply", L.ptr(dyh), L.ptr(xh), L.ptr(yh), L.ptr(mean), L.ptr(rstd), L.ptr(g32), L.ptr(sums), L.ptr(dxh),
               L.ptr(dres), LL(rows), I(Cc), I(int(relu)), I(dt), _s())
        return (_nchw_view(dxh), dgamma.to(gdt), dbeta.to(gdt), None, None,
                _nchw_view(dres) if has_res else None, None, None, None, None)


def batch_norm(x, weight, bias, running_mean, running_var, training: bool, momentum: float = 0.1, eps: float = 1e-5,
               residual=None, relu: bool = False):
    """BatchNorm2d (+ optional residual add and ReLU fused into the normalisation pass)."""
    if not x.is_cuda or x.shape[1] % 8:
        y = F.batch_norm(x, running_mean, running_var, weight, bias, training, momentum, eps)
        if residual is not None:
            y = y + residual
        return F.relu(y) if relu else y
    return _BatchNormFn.apply(x, weight, bias, running_mean, running_var, residual, training, momentum, eps, relu)


# ------------------------------------------------------------------------------ pooling
class _MaxPoolFn(Function):
    @staticmethod
    def forward(ctx, x, k, stride, pad):
        xh = _nhwc(x)
        Nn, H, W, Cc = xh.shape
        Ho, Wo = _conv_geom(H, W, k, k, stride, pad)
        yh = torch.empty((Nn, Ho, Wo, Cc), dtype=xh.dtype, device=x.device)
        idx = torch.empty((Nn, Ho, Wo, Cc), dtype=torch.uint8, device=x.device)
        L.call("maxpool", L.ptr(None), L.ptr(xh), L.ptr(idx), L.ptr(yh), I(Nn), I(H), I(W), I(Cc), I(k), I(stride), I(pad),
               I(Ho), I(Wo), I(0), I(L.dtype_code(xh.dtype)), _s())
        ctx.save_for_backward(idx)
        ctx.g = (k, stride, pad, Ho, Wo, Nn, H, W, Cc, xh.dtype)
        return _nchw_view(yh)

    @staticmethod
    def backward(ctx, dy):
        (idx,) = ctx.saved_tensors
        k, stride, pad, Ho, Wo, Nn, H, W, Cc, dt = ctx.g
        dyh = _nhwc(dy.to(dt))
        dxh = torch.empty((Nn, H, W, Cc), dtype=dt, device=dy.device)
        L.call("maxpool", L.ptr(dyh), L.ptr(None), L.ptr(idx), L.ptr(dxh), I(Nn), I(H), I(W), I(Cc), I(k), I(stride), I(pad),
               I(Ho), I(Wo), I(1), I(L.dtype_code(dt)), _s())
        return _nchw_view(dxh), None, None, None


def max_pool2d(x, kernel_size: int, stride: Optional[int] = None, padding: int = 0):
    stride = kernel_size if stride is None else stride
    if not x.is_cuda or x.shape[1] % 8:
        return F.max_pool2d(x, kernel_size, stride, padding)
    return _MaxPoolFn.apply(x, kernel_size, stride, padding)


class _AvgPoolFn(Function):
    @staticmethod
    def forward(ctx, x):
        xh = _nhwc(x)
        Nn, H, W, Cc = xh.shape
        y = torch.empty((Nn, Cc), dtype=xh.dtype, device=x.device)
        L.call("avgpool", L.ptr(xh), L.ptr(y), I(Nn), I(H * W), I(Cc), I(0), I(L.dtype_code(xh.dtype)), _s())
        ctx.shape = (Nn, H, W, Cc)
        return y

    @staticmethod
    def backward(ctx, dy):
        Nn, H, W, Cc = ctx.shape
        dy = dy.contiguous()
        dxh = torch.empty((Nn, H, W, Cc), dtype=dy.dtype, device=dy.device)
        L.call("avgpool", L.ptr(dy), L.ptr(dxh), I(Nn), I(H * W), I(Cc), I(1), I(L.dtype_code(dy.dtype)), _s())
        return _nchw_view(dxh)


def global_avg_pool(x):
    """AdaptiveAvgPool2d(1) + flatten: [N,C,H,W] -> [N,C]."""
    if not x.is_cuda:
        return x.mean((2, 3))
    return _AvgPoolFn.apply(x)


# ------------------------------------------------------------------------------ transposed conv 2x2 / stride 2
class _ConvT2x2Fn(Function):
    @staticmethod
    def forward(ctx, x, w, bias):
        # w: [Cin, Cout, 2, 2] (nn.ConvTranspose2d layout); y[n, 2h+dh, 2w+dw, k] = sum_c x[n,h,w,c] w[c,k,dh,dw]
        xh = _nhwc(x if x.dtype == _BF16 else x.to(_BF16))
        Nn, H, W, Cin = xh.shape
        Cout = w.shape[1]
        wm = w.permute(2, 3, 1, 0).reshape(4 * Cout, Cin).to(_BF16).contiguous()   # [(dh,dw,k), c]
        b4 = bias.to(_BF16).repeat(4) if bias is not None else None
        g = N.gemm(xh.reshape(-1, Cin), wm, bias=b4, out_dtype=_BF16)               # [NHW, 4*Cout]
        yh = torch.empty((Nn, 2 * H, 2 * W, Cout), dtype=_BF16, device=x.device)
        L.call("pixel_shuffle2", L.ptr(g), L.ptr(yh), I(Nn), I(H), I(W), I(Cout), I(0), I(0), _s())
        ctx.save_for_backward(xh, w)
        ctx.meta = (x.dtype, bias is not None)
        return _nchw_view(yh).to(x.dtype)

    @staticmethod
    def backward(ctx, dy):
        xh, w = ctx.saved_tensors
        x_dtype, has_bias = ctx.meta
        Nn, H, W, Cin = xh.shape
        Cout = w.shape[1]
        dyh = _nhwc(dy.to(_BF16))
        dg = torch.empty((Nn * H * W, 4 * Cout), dtype=_BF16, device=dy.device)
        L.call("pixel_shuffle2", L.ptr(dyh), L.ptr(dg), I(Nn), I(H), I(W), I(Cout), I(1), I(0), _s())
        wm = w.permute(2, 3, 1, 0).reshape(4 * Cout, Cin).to(_BF16).contiguous()
        dx = dw = db = None
        if ctx.needs_input_grad[0]:
            dx = _nchw_view(N.gemm(dg, wm, b_mn=True, out_dtype=_BF16).view(Nn, H, W, Cin)).to(x_dtype)
        if ctx.needs_input_grad[1]:
            dwm = N.gemm(dg, xh.reshape(-1, Cin), a_mn=True, b_mn=True, out_dtype=torch.float32)  # [4*Cout, Cin]
            dw = dwm.view(2, 2, Cout, Cin).permute(3, 2, 0, 1).to(w.dtype)
        if has_bias and ctx.needs_input_grad[2]:
            db = dy.float().sum((0, 2, 3))
        return dx, dw, db


def conv_transpose2x2(x, weight, bias=None):
    """nn.ConvTranspose2d(kernel_size=2, stride=2): GEMM + pixel shuffle."""
    if not x.is_cuda or x.shape[1] % 8 or weight.shape[1] % 2:
        return F.conv_transpose2d(x, weight, bias, stride=2)
    return _ConvT2x2Fn.apply(x, weight, bias)


# ------------------------------------------------------------------------------ bilinear resize
class _BilinearFn(Function):
    @staticmethod
    def forward(ctx, x, Ho, Wo):
        xh = _nhwc(x)
        Nn, Hi, Wi, Cc = xh.shape
        yh = torch.empty((Nn, Ho, Wo, Cc), dtype=xh.dtype, device=x.device)
        L.call("bilinear", L.ptr(xh), L.ptr(yh), I(Nn), I(Hi), I(Wi), I(Ho), I(Wo), I(Cc), I(0), I(L.dtype_code(xh.dtype)), _s())
        ctx.g = (Nn, Hi, Wi, Ho, Wo, Cc, x.dtype)
        return _nchw_view(yh)

    @staticmethod
    def backward(ctx, dy):
        Nn, Hi, Wi, Ho, Wo, Cc, dt = ctx.g
        dyh = _nhwc(dy)
        dx32 = torch.zeros((Nn, Hi, Wi, Cc), dtype=torch.float32, device=dy.device)
        L.call("bilinear", L.ptr(dyh), L.ptr(dx32), I(Nn), I(Hi), I(Wi), I(Ho), I(Wo), I(Cc), I(1),
               I(L.dtype_code(dyh.dtype)), _s())
        return _nchw_view(dx32.to(dt)), None, None


def bilinear_resize(x, size):
    """F.interpolate(x, size, mode="bilinear", align_corners=False)."""
    Ho, Wo = int(size[0]), int(size[1])
    if tuple(x.shape[2:]) == (Ho, Wo):
        return x
    if not x.is_cuda:
        return F.interpolate(x, size=(Ho, Wo), mode="bilinear", align_corners=False)
    return _BilinearFn.apply(x, Ho, Wo)


# ------------------------------------------------------------------------------ latitude-weighted MSE
def lat_weights(H: int, device) -> torch.Tensor:
    lat = torch.linspace(90, -90, H, device=device)
    w = torch.cos(torch.deg2rad(lat))
    return (w / w.mean()).float().contiguous()


class _LatMSEFn(Function):
    @staticmethod
    def forward(ctx, pred, target):
        cl = pred.is_contiguous(memory_format=torch.channels_last) and not pred.is_contiguous()
        p = pred if (pred.is_contiguous() or cl) else pred.contiguous()
        cl = p.is_contiguous(memory_format=torch.channels_last) and not p.is_contiguous()
        t = target.to(p.dtype).contiguous(memory_format=torch.channels_last if cl else torch.contiguous_format)
        Nn, Cc, H, W = p.shape
        wl = lat_weights(H, p.device)
        loss = torch.zeros(1, dtype=torch.float32, device=p.device)
        h_div = W * Cc if cl else W
        L.call("lat_mse", L.ptr(p), L.ptr(t), L.ptr(wl), L.ptr(loss), L.ptr(None), L.ptr(None), LL(p.numel()), I(H), LL(h_div),
               I(0), I(L.dtype_code(p.dtype)), _s())
        ctx.save_for_backward(p, t, wl)
        ctx.h_div = h_div
        return loss[0]

    @staticmethod
    def backward(ctx, g):
        p, t, wl = ctx.saved_tensors
        grad = torch.empty_like(p)
        gs = g.to(torch.float32).reshape(1).contiguous()
        L.call("lat_mse", L.ptr(p), L.ptr(t), L.ptr(wl), L.ptr(None), L.ptr(grad), L.ptr(gs), LL(p.numel()), I(p.shape[2]),
               LL(ctx.h_div), I(1), I(L.dtype_code(p.dtype)), _s())
        return grad, None


def latitude_weighted_mse(pred, target):
    """mean(cos(lat)-normalised weights * (pred - target)^2)  (reference multinode_ddp_unet.py:221-229)."""
    if not pred.is_cuda:
        w = lat_weights(pred.shape[-2], pred.device).view(1, 1, -1, 1)
        return (w * (pred - target) ** 2).mean()
    return _LatMSEFn.apply(pred, target)
