"""PipelineTransformer of the reference's PP training recipe
(/root/reference/scripts/04_pipeline_parallel_pp/03_pipeline_training.py:51-120): token + learned position
embedding (1024 positions), ``num_stages`` stages of pre-norm blocks (multi-head attention with dropout 0.1,
Linear-GELU-Linear-Dropout), final LayerNorm + bias-free vocab projection.  Defaults: vocab 10000, dim 256,
8 heads, 2 layers per stage.  ``build_stage`` constructs ONLY the requested stage (the reference builds the whole
model on every rank and traces it)."""
from __future__ import annotations

import torch
from torch import nn

from .. import ops
from .layers import GELU, LayerNorm, Linear


class MHA(nn.Module):
    """Self-attention with fused in-projection (parameter names as nn.MultiheadAttention: in_proj_weight,
    in_proj_bias, out_proj.*), non-causal like the reference block; dropout on the attention PROBABILITIES, as
    nn.MultiheadAttention(dropout=p) applies it (Philox mask regenerated in backward, ``ops.attention_dropout``)."""

    def __init__(self, dim, heads, dropout=0.1):
        super().__init__()
        self.heads, self.hd = heads, dim // heads
        self.in_proj_weight = nn.Parameter(torch.empty(3 * dim, dim))
        self.in_proj_bias = nn.Parameter(torch.zeros(3 * dim))
        self.out_proj = Linear(dim, dim)
        self.dropout = dropout
        nn.init.xavier_uniform_(self.in_proj_weight)

    def forward(self, x):
        B, S, D = x.shape
        qkv = ops.linear(x, self.in_proj_weight, self.in_proj_bias).view(B, S, 3, self.heads, self.hd)
        o = ops.attention_dropout(qkv[:, :, 0], qkv[:, :, 1], qkv[:, :, 2], self.dropout, causal=False, training=self.training)
        return self.out_proj(o.reshape(B, S, D))


class Block(nn.Module):
    def __init__(self, dim, heads, ff_dim=None, dropout=0.1):
        super().__init__()
        ff_dim = ff_dim or 4 * dim
        self.norm1, self.attn = LayerNorm(dim), MHA(dim, heads, dropout)
        self.norm2 = LayerNorm(dim)
        self.ff = nn.Sequential(Linear(dim, ff_dim), GELU(), Linear(ff_dim, dim), nn.Dropout(dropout))

    def forward(self, x):
        x = x + self.attn(self.norm1(x))
        return x + self.ff(self.norm2(x))


class Stage(nn.Module):
    def __init__(self, index, num_stages, vocab, dim, heads, layers, max_pos=1024, dropout=0.1):
        super().__init__()
        self.first, self.last = index == 0, index == num_stages - 1
        if self.first:
            self.tok_emb, self.pos_emb = nn.Embedding(vocab, dim), nn.Embedding(max_pos, dim)
        self.blocks = nn.Sequential(*[Block(dim, heads, dropout=dropout) for _ in range(layers)])
        if self.last:
            self.norm, self.head = LayerNorm(dim), Linear(dim, vocab, bias=False)

    def forward(self, x):
        if self.first:
            pos = torch.arange(x.shape[1], device=x.device)
            x = ops.embedding(x, self.tok_emb.weight) + ops.embedding(pos, self.pos_emb.weight).unsqueeze(0)
        x = self.blocks(x)
        if self.last:
            x = self.head(self.norm(x))
        return x


class PipelineTransformer(nn.Module):
    def __init__(self, vocab_size=10000, dim=256, num_heads=8, num_stages=4, layers_per_stage=2, dropout=0.1):
        super().__init__()
        self.stages = nn.ModuleList(Stage(i, num_stages, vocab_size, dim, num_heads, layers_per_stage, dropout=dropout)
                                    for i in range(num_stages))

    def forward(self, x):
        for s in self.stages:
            x = s(x)
        return x


def build_stage(index, num_stages, vocab_size=10000, dim=256, num_heads=8, layers_per_stage=2, dropout=0.1) -> Stage:
    return Stage(index, num_stages, vocab_size, dim, num_heads, layers_per_stage, dropout=dropout)
