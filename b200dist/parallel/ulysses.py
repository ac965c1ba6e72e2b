"""Ulysses sequence parallelism: all-to-all swaps the sharded dimension from sequence to heads around attention.

Docs-only in the reference (/root/reference/docs/guide/08_sequence_parallel.md:41-80: [B, S/P, H, hd] -> all-to-all ->
[B, S, H/P, hd] -> flash attention -> all-to-all back; needs H % P == 0, :151); implemented here on the peer-memory
all-to-all kernel (direct peer pulls of [S/P, H/P, hd] tiles) and the tcgen05 flash-attention kernel.
"""
from __future__ import annotations

import torch
from torch.autograd import Function

from .. import ops


def _seq_to_heads(x, comm):
    """[B, S/P, H, hd] (sequence shard, all heads) -> [B, S, H/P, hd] (full sequence, head shard)."""
    P = comm.world
    B, s, H, hd = x.shape
    assert H % P == 0, f"Ulysses needs heads ({H}) divisible by the group size ({P})"
    send = x.reshape(B, s, P, H // P, hd).permute(2, 0, 1, 3, 4).contiguous()      # [P(dst), B, s, H/P, hd]
    recv = torch.empty_like(send)                                                    # [P(src), B, s, H/P, hd]
    comm.all_to_all(recv, send, channel=4)
    return recv.permute(1, 0, 2, 3, 4).reshape(B, P * s, H // P, hd)


def _heads_to_seq(x, comm):
    """Inverse of ``_seq_to_heads``: [B, S, H/P, hd] -> [B, S/P, H, hd]."""
    P = comm.world
    B, S, h, hd = x.shape
    s = S // P
    send = x.reshape(B, P, s, h, hd).permute(1, 0, 2, 3, 4).contiguous()            # [P(dst = seq chunk), B, s, h, hd]
    recv = torch.empty_like(send)                                                    # [P(src = head group), B, s, h, hd]
    comm.all_to_all(recv, send, channel=4)
    return recv.permute(1, 2, 0, 3, 4).reshape(B, s, P * h, hd)


class _SeqToHeads(Function):
    @staticmethod
    def forward(ctx, x, comm):
        ctx.comm = comm
        return _seq_to_heads(x, comm)

    @staticmethod
    def backward(ctx, g):
        return _heads_to_seq(g.contiguous(), ctx.comm), None


class _HeadsToSeq(Function):
    @staticmethod
    def forward(ctx, x, comm):
        ctx.comm = comm
        return _heads_to_seq(x, comm)

    @staticmethod
    def backward(ctx, g):
        return _seq_to_heads(g.contiguous(), ctx.comm), None


def ulysses_attention(q, k, v, comm, causal: bool = True, scale=None):
    """Attention over a sequence sharded across ``comm``: inputs / output are [B, S/P, H, hd] local shards."""
    if comm.world == 1:
        return ops.attention(q, k, v, causal=causal, scale=scale)
    qf, kf, vf = (_SeqToHeads.apply(t.contiguous(), comm) for t in (q, k, v))
    o = ops.attention(qf, kf, vf, causal=causal, scale=scale)
    return _HeadsToSeq.apply(o.contiguous(), comm)
