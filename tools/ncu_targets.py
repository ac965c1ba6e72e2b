"""Tiny launcher for ncu captures: runs ONE kernel family a few times.
    python tools/ncu_targets.py gemm|gemm_wgrad|attn_fwd|attn_bwd|rmsnorm|adamw|conv|conv_s2|conv_wgrad"""
import sys
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import torch

from b200dist.ops import native as N

what = sys.argv[1]
dev = "cuda"
torch.manual_seed(0)
if what == "gemm":
    a = torch.randn(8192, 4096, device=dev).bfloat16(); b = torch.randn(4096, 4096, device=dev).bfloat16()
    for _ in range(4):
        N.gemm(a, b, config=3)
elif what == "gemm_wgrad":
    a = torch.randn(8192, 4096, device=dev).bfloat16(); b = torch.randn(8192, 4096, device=dev).bfloat16()
    for _ in range(4):
        N.gemm(a, b, a_mn=True, b_mn=True, config=3)
elif what in ("attn_fwd", "attn_bwd"):
    q = torch.randn(1, 4096, 32, 128, device=dev).bfloat16(); k = torch.randn_like(q); v = torch.randn_like(q); do = torch.randn_like(q)
    for _ in range(3):
        o, lse = N.flash_attn_fwd(q, k, v, causal=True)
        if what == "attn_bwd":
            N.flash_attn_bwd(q, k, v, o, do, lse, causal=True)
elif what == "rmsnorm":
    x = torch.randn(8192, 4096, device=dev).bfloat16(); w = torch.ones(4096, device=dev).bfloat16()
    for _ in range(4):
        y, r = N.rmsnorm_fwd(x, w, 1e-5)
        N.rmsnorm_bwd(x, x, w, r)
elif what == "adamw":
    n = 64 << 20
    p = torch.randn(n, device=dev); g = torch.randn(n, device=dev).bfloat16(); m = torch.zeros(n, device=dev); v = torch.zeros(n, device=dev)
    lp = torch.empty(n, device=dev, dtype=torch.bfloat16)
    for _ in range(4):
        N.adamw_step(p, g, m, v, lr=1e-3, beta1=0.9, beta2=0.95, eps=1e-8, wd=0.1, step=3, p_lowp=lp)
elif what in ("conv", "conv_wgrad", "conv_s2"):
    stride = 2 if what == "conv_s2" else 1
    C, K, H = (128, 128, 56) if stride == 2 else (64, 64, 56)
    x = torch.randn(256, H, H, C, device=dev).bfloat16()
    wm = (torch.randn(K, 9 * C, device=dev) * 0.05).bfloat16()
    Ho = (H + 2 - 3) // stride + 1
    dy = torch.randn(256, Ho, Ho, K, device=dev).bfloat16()
    for _ in range(3):
        if what == "conv_wgrad":
            N.conv2d_igemm_wgrad(x, dy, 3, 3, 1)
        else:
            N.conv2d_igemm(x, wm, 3, 3, 1, stride=stride)
torch.cuda.synchronize()
print("done", what)
