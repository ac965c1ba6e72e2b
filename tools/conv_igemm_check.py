"""Numerics + timing of the implicit-GEMM convolution (fprop and dgrad) against F.conv2d and the im2col path.
    python tools/conv_igemm_check.py"""
import sys
from pathlib import Path

import torch
import torch.nn.functional as F

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from b200dist.ops import conv as cops  # noqa: E402
from b200dist.ops import native as N  # noqa: E402

dev = torch.device("cuda", 0)
torch.manual_seed(0)
fails = 0


def ev(fn, it=10):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    s.record()
    for _ in range(it):
        fn()
    e.record(); torch.cuda.synchronize()
    return s.elapsed_time(e) / it * 1e3


for (B, C, H, W, K, ks, pad) in [(2, 64, 56, 56, 64, 3, 1), (3, 128, 28, 28, 128, 3, 1), (2, 256, 14, 14, 256, 3, 1),
                                 (5, 512, 7, 7, 512, 3, 1), (2, 256, 56, 56, 64, 1, 0), (2, 64, 45, 90, 128, 3, 1),
                                 (1, 64, 20, 181, 72, 3, 1), (2, 64, 9, 9, 64, 3, 0), (256, 64, 56, 56, 64, 3, 1)]:
    x = torch.randn(B, C, H, W, device=dev).bfloat16()
    w = (torch.randn(K, C, ks, ks, device=dev) * (1.0 / (C * ks * ks) ** 0.5)).bfloat16()
    xh = x.permute(0, 2, 3, 1).contiguous()
    wm = w.permute(0, 2, 3, 1).reshape(K, ks * ks * C).contiguous()
    y = N.conv2d_igemm(xh, wm, ks, ks, pad)
    torch.cuda.synchronize()
    ref = F.conv2d(x.float(), w.float(), padding=pad).permute(0, 2, 3, 1)
    err = float((y.float() - ref).abs().max()); scale = float(ref.abs().max())
    ok = err <= 0.02 * scale + 0.02
    fails += (not ok)
    msg = f"[{'PASS' if ok else 'FAIL'}] igemm fprop B{B} C{C} {H}x{W} -> K{K} k{ks} p{pad}: max_err={err:.4g} (|ref| {scale:.3g})"
    # dgrad through the same kernel: dY [B, Ho, Wo, K] with the flipped filter, pad' = ks - 1 - pad
    if K % 64 == 0 and C % 8 == 0:
        dy = torch.randn_like(ref).bfloat16().contiguous()
        wflip = w.flip(2, 3).permute(1, 2, 3, 0).reshape(C, ks * ks * K).contiguous()
        dx = N.conv2d_igemm(dy, wflip, ks, ks, ks - 1 - pad)
        torch.cuda.synchronize()
        xr = x.float().requires_grad_(True)
        F.conv2d(xr, w.float(), padding=pad).backward(dy.float().permute(0, 3, 1, 2))
        dref = xr.grad.permute(0, 2, 3, 1)
        e2 = float((dx.float() - dref).abs().max()); s2 = float(dref.abs().max())
        ok2 = e2 <= 0.02 * s2 + 0.02 and dx.shape == dref.shape
        fails += (not ok2)
        msg += f" | dgrad max_err={e2:.4g} (|ref| {s2:.3g}) {'PASS' if ok2 else 'FAIL'}"
    if K % 8 == 0:   # weight gradient through the pixel-box split-K kernel
        dyw = torch.randn_like(ref).bfloat16().contiguous()
        dwm = N.conv2d_igemm_wgrad(xh, dyw, ks, ks, pad)
        torch.cuda.synchronize()
        wr = w.float().requires_grad_(True)
        F.conv2d(x.float(), wr, padding=pad).backward(dyw.float().permute(0, 3, 1, 2))
        wref = wr.grad.permute(0, 2, 3, 1).reshape(K, -1)
        e3 = float((dwm - wref).abs().max()); s3 = float(wref.abs().max())
        ok3 = e3 <= 0.02 * s3 + 0.02
        fails += (not ok3)
        msg += f" | wgrad max_err={e3:.4g} (|ref| {s3:.3g}) {'PASS' if ok3 else 'FAIL'}"
        if B >= 128:
            t_wg = ev(lambda: N.conv2d_igemm_wgrad(xh, dyw, ks, ks, pad))
            col = cops._im2col(xh, ks, ks, 1, pad, ref.shape[1], ref.shape[2], C)
            dy2 = dyw.view(-1, K)
            t_g = ev(lambda: N.gemm(dy2, col, a_mn=True, b_mn=True, out_dtype=torch.float32))
            t_i = ev(lambda: cops._im2col(xh, ks, ks, 1, pad, ref.shape[1], ref.shape[2], C))
            msg += f" | wgrad time igemm {t_wg:.0f} us vs GEMM {t_g:.0f} + im2col {t_i:.0f} us"
    if B >= 128:
        t_ig = ev(lambda: N.conv2d_igemm(xh, wm, ks, ks, pad))
        xc = x.contiguous(memory_format=torch.channels_last)
        t_old = ev(lambda: cops.conv2d(xc, w, None, 1, pad))
        t_cudnn = ev(lambda: F.conv2d(xc, w, padding=pad))
        msg += f" | time igemm {t_ig:.0f} us, im2col+GEMM {t_old:.0f} us, cuDNN {t_cudnn:.0f} us"
    print(msg, flush=True)
# ---- stride-2 forward and weight gradient (tensor map with element strides): ResNet stage transitions, odd extents ----
for (B, C, H, W, K, ks, pad) in [(2, 128, 56, 56, 128, 3, 1), (2, 256, 56, 56, 512, 1, 0), (3, 64, 29, 41, 72, 3, 1),
                                 (2, 512, 14, 14, 512, 3, 1), (4, 1024, 14, 14, 2048, 1, 0), (64, 128, 56, 56, 128, 3, 1)]:
    x = torch.randn(B, C, H, W, device=dev).bfloat16()
    w = (torch.randn(K, C, ks, ks, device=dev) * (1.0 / (C * ks * ks) ** 0.5)).bfloat16()
    xh = x.permute(0, 2, 3, 1).contiguous()
    wm = w.permute(0, 2, 3, 1).reshape(K, ks * ks * C).contiguous()
    y = N.conv2d_igemm(xh, wm, ks, ks, pad, stride=2)
    torch.cuda.synchronize()
    ref = F.conv2d(x.float(), w.float(), stride=2, padding=pad).permute(0, 2, 3, 1)
    ok = tuple(y.shape) == tuple(ref.shape)
    err = float((y.float() - ref).abs().max()) if ok else float("inf"); scale = float(ref.abs().max())
    ok = ok and err <= 0.02 * scale + 0.02
    fails += (not ok)
    msg = f"[{'PASS' if ok else 'FAIL'}] igemm fprop STRIDE 2 B{B} C{C} {H}x{W} -> K{K} k{ks} p{pad}: max_err={err:.4g} (|ref| {scale:.3g})"
    dyw = torch.randn_like(ref).bfloat16().contiguous()
    if B * ref.shape[1] * ref.shape[2] >= 128:
        dwm = N.conv2d_igemm_wgrad(xh, dyw, ks, ks, pad, stride=2)
        torch.cuda.synchronize()
        wr = w.float().requires_grad_(True)
        F.conv2d(x.float(), wr, stride=2, padding=pad).backward(dyw.float().permute(0, 3, 1, 2))
        wref = wr.grad.permute(0, 2, 3, 1).reshape(K, -1)
        e3 = float((dwm - wref).abs().max()); s3 = float(wref.abs().max())
        ok3 = e3 <= 0.02 * s3 + 0.02
        fails += (not ok3)
        msg += f" | wgrad max_err={e3:.4g} (|ref| {s3:.3g}) {'PASS' if ok3 else 'FAIL'}"
    # the autograd op end to end (forward igemm, dgrad = GEMM + col2im, wgrad igemm) vs F.conv2d
    xa = x.clone().requires_grad_(True); wa = w.clone().requires_grad_(True)
    ya = cops.conv2d(xa, wa, None, 2, pad)
    ya.backward(dyw.permute(0, 3, 1, 2))
    xr = x.float().requires_grad_(True); wr2 = w.float().requires_grad_(True)
    F.conv2d(xr, wr2, stride=2, padding=pad).backward(dyw.float().permute(0, 3, 1, 2))
    e4 = float((xa.grad.float() - xr.grad).abs().max()); s4 = float(xr.grad.abs().max())
    e5 = float((wa.grad.float() - wr2.grad).abs().max()); s5 = float(wr2.grad.abs().max())
    ok4 = e4 <= 0.03 * s4 + 0.02 and e5 <= 0.03 * s5 + 0.02
    fails += (not ok4)
    msg += f" | autograd dx err {e4:.3g} dw err {e5:.3g} {'PASS' if ok4 else 'FAIL'}"
    if B >= 64:
        t_ig = ev(lambda: N.conv2d_igemm(xh, wm, ks, ks, pad, stride=2))
        xc = x.contiguous(memory_format=torch.channels_last)
        t_cudnn = ev(lambda: F.conv2d(xc, w, stride=2, padding=pad))
        Ho, Wo = ref.shape[1], ref.shape[2]
        t_im = ev(lambda: cops._im2col(xh, ks, ks, 2, pad, Ho, Wo, C))
        msg += f" | time igemm {t_ig:.0f} us, im2col alone {t_im:.0f} us, cuDNN {t_cudnn:.0f} us"
    print(msg, flush=True)
# ---- small-channel stem (7x7 stride 2 pad 3 on RGB, 5-D tensor map): forward + weight gradient through the autograd op ----
for (B, H, W, K, ks, pad) in [(4, 224, 224, 64, 7, 3), (3, 96, 130, 32, 7, 3), (2, 64, 64, 64, 5, 2), (128, 224, 224, 64, 7, 3)]:
    x = torch.randn(B, 3, H, W, device=dev).bfloat16()
    w = (torch.randn(K, 3, ks, ks, device=dev) * (1.0 / (3 * ks * ks) ** 0.5)).bfloat16()
    wa = w.clone().requires_grad_(True)
    y = cops.conv2d(x, wa, None, 2, pad)
    ref = F.conv2d(x.float(), w.float(), stride=2, padding=pad)
    ok = tuple(y.shape) == tuple(ref.shape)
    err = float((y.float() - ref).abs().max()) if ok else float("inf"); scale = float(ref.abs().max())
    ok = ok and err <= 0.02 * scale + 0.02
    dy = torch.randn_like(ref).bfloat16()
    y.backward(dy)
    wr = w.float().requires_grad_(True)
    F.conv2d(x.float(), wr, stride=2, padding=pad).backward(dy.float())
    e2 = float((wa.grad.float() - wr.grad).abs().max()); s2 = float(wr.grad.abs().max())
    ok2 = e2 <= 0.03 * s2 + 0.02
    fails += (not ok) + (not ok2)
    msg = (f"[{'PASS' if ok else 'FAIL'}] stem igemm B{B} 3x{H}x{W} -> K{K} k{ks} s2 p{pad}: max_err={err:.4g} (|ref| {scale:.3g}) | "
           f"wgrad max_err={e2:.4g} (|ref| {s2:.3g}) {'PASS' if ok2 else 'FAIL'}")
    if B >= 128:
        xc = x.contiguous(memory_format=torch.channels_last)
        t_st = ev(lambda: cops.conv2d(xc, w, None, 2, pad))
        import os
        t_cudnn = ev(lambda: F.conv2d(xc, w, stride=2, padding=pad))
        msg += f" | time stem igemm (incl. padding copy) {t_st:.0f} us, cuDNN {t_cudnn:.0f} us"
    print(msg, flush=True)
print(f"== conv igemm: failures {fails}")
sys.exit(1 if fails else 0)
