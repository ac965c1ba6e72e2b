"""SimpleViT for weather grids, same module names as the reference
(/root/reference/scripts/03_tensor_parallel_tp/tensor_parallel_vit.py:82-202): ``patch_embed.proj`` (conv with
kernel = stride = patch, executed as reshape + GEMM), ``pos_embed``, ``blocks.{i}.{norm1,attn.{q,k,v,out}_proj,
norm2,mlp.{fc1,fc2}}``, ``norm``, ``head``.  The reference materialises softmax(QK^T) as a [B,H,N,N] tensor
(:120-122); here it is the non-causal tcgen05 flash-attention kernel.  ``reshape(B, N, -1, hd)`` keeps the head
count implicit so the same code runs on head-sharded (tensor-parallel) projections."""
from __future__ import annotations

import torch
from torch import nn

from .. import ops
from ..ops import conv as cops
from .layers import GELU, LayerNorm, Linear


class PatchEmbed(nn.Module):
    def __init__(self, in_channels, embed_dim, patch_size):
        super().__init__()
        self.proj = nn.Conv2d(in_channels, embed_dim, kernel_size=patch_size, stride=patch_size)
        self.patch_size = patch_size

    def forward(self, x):
        B, C, H, W = x.shape
        p = self.patch_size
        hp, wp = H // p, W // p
        # non-overlapping patches: unfold is a pure reshape, the conv is one GEMM
        patches = x[:, :, :hp * p, :wp * p].reshape(B, C, hp, p, wp, p).permute(0, 2, 4, 1, 3, 5).reshape(B, hp * wp, C * p * p)
        return ops.linear(patches, self.proj.weight.reshape(self.proj.weight.shape[0], -1), self.proj.bias)


class Attention(nn.Module):
    def __init__(self, dim, num_heads):
        super().__init__()
        self.num_heads, self.head_dim = num_heads, dim // num_heads
        self.scale = self.head_dim ** -0.5
        self.q_proj, self.k_proj, self.v_proj, self.out_proj = Linear(dim, dim), Linear(dim, dim), Linear(dim, dim), Linear(dim, dim)

    def forward(self, x):
        B, N, _ = x.shape
        q = self.q_proj(x).reshape(B, N, -1, self.head_dim)
        k = self.k_proj(x).reshape(B, N, -1, self.head_dim)
        v = self.v_proj(x).reshape(B, N, -1, self.head_dim)
        o = ops.attention(q, k, v, causal=False, scale=self.scale)
        return self.out_proj(o.reshape(B, N, -1))


class MLP(nn.Module):
    def __init__(self, dim, hidden_dim):
        super().__init__()
        self.fc1, self.act, self.fc2 = Linear(dim, hidden_dim), GELU(), Linear(hidden_dim, dim)

    def forward(self, x):
        return self.fc2(self.act(self.fc1(x)))


class TransformerBlock(nn.Module):
    def __init__(self, dim, num_heads, mlp_ratio=4):
        super().__init__()
        self.norm1, self.attn = LayerNorm(dim), Attention(dim, num_heads)
        self.norm2, self.mlp = LayerNorm(dim), MLP(dim, int(dim * mlp_ratio))

    def forward(self, x):
        x = x + self.attn(self.norm1(x))
        return x + self.mlp(self.norm2(x))


class SimpleViT(nn.Module):
    def __init__(self, in_channels, out_channels, patch_size, lat, lon, embed_dim=256, depth=6, num_heads=8, mlp_ratio=4):
        super().__init__()
        self.patch_size, self.out_channels = patch_size, out_channels
        self.h_patches, self.w_patches = lat // patch_size, lon // patch_size
        self.patch_embed = PatchEmbed(in_channels, embed_dim, patch_size)
        self.pos_embed = nn.Parameter(torch.randn(1, self.h_patches * self.w_patches, embed_dim) * 0.02)
        self.blocks = nn.ModuleList([TransformerBlock(embed_dim, num_heads, mlp_ratio) for _ in range(depth)])
        self.norm = LayerNorm(embed_dim)
        self.head = Linear(embed_dim, out_channels * patch_size * patch_size)

    def forward(self, x):
        B = x.shape[0]
        x = self.patch_embed(x) + self.pos_embed
        for blk in self.blocks:
            x = blk(x)
        x = self.head(self.norm(x))
        p = self.patch_size
        x = x.reshape(B, self.h_patches, self.w_patches, self.out_channels, p, p)
        return x.permute(0, 3, 1, 4, 2, 5).reshape(B, self.out_channels, self.h_patches * p, self.w_patches * p)


def vit_tp_plan(model: SimpleViT, tp_mesh):
    """q/k/v/fc1 column-wise, out_proj/fc2 row-wise per block (reference tensor_parallel_vit.py:352-361)."""
    from ..parallel.tp import ColwiseParallel, RowwiseParallel, parallelize_module

    for blk in model.blocks:
        parallelize_module(blk, tp_mesh, {
            "attn.q_proj": ColwiseParallel(), "attn.k_proj": ColwiseParallel(), "attn.v_proj": ColwiseParallel(),
            "attn.out_proj": RowwiseParallel(), "mlp.fc1": ColwiseParallel(), "mlp.fc2": RowwiseParallel()})
    return model
