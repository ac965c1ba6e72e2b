"""Domain parallelism: a spatially sharded convolution with halo exchange.

Docs-only in the reference (/root/reference/docs/guide/10_domain_parallel.md:45-149: splitting the height dimension
breaks a 3x3 convolution at the cut unless each rank first receives a one-row halo from its neighbours; the guide points
to PhysicsNeMo ShardTensor).  Here the input is sharded along H; ``halo_exchange`` attaches ``halo`` boundary rows of
the neighbouring ranks (zeros at the global border) through ONE small peer-memory all-gather of boundary rows, and its
backward returns the halo gradients to their owners.  ``halo_conv2d`` then equals the corresponding rows of the
unsharded convolution exactly.
"""
from __future__ import annotations

import torch
from torch.autograd import Function

from ..ops import conv as cops


def _exchange(x, comm, halo):
    """-> (top, bottom): halo rows received from rank-1 / rank+1 (zeros at the domain border)."""
    P, r = comm.world, comm.rank
    edges = torch.stack([x[:, :, :halo, :], x[:, :, -halo:, :]]).contiguous()       # [2, N, C, halo, W]: my top / bottom rows
    allb = torch.empty((P,) + tuple(edges.shape), dtype=x.dtype, device=x.device)
    comm.all_gather(allb, edges, channel=5)
    zeros = torch.zeros_like(edges[0])
    top = allb[r - 1, 1] if r > 0 else zeros            # neighbour above contributes ITS bottom rows
    bottom = allb[r + 1, 0] if r < P - 1 else zeros     # neighbour below contributes ITS top rows
    return top, bottom


class _HaloExchange(Function):
    @staticmethod
    def forward(ctx, x, comm, halo):
        ctx.comm, ctx.halo = comm, halo
        top, bottom = _exchange(x, comm, halo)
        return torch.cat([top, x, bottom], dim=2)

    @staticmethod
    def backward(ctx, g):
        comm, h = ctx.comm, ctx.halo
        P, r = comm.world, comm.rank
        core = g[:, :, h:-h, :].clone()
        # gradient of my top halo belongs to rank-1's bottom rows; of my bottom halo to rank+1's top rows
        send = torch.stack([g[:, :, :h, :], g[:, :, -h:, :]]).contiguous()          # [2(top, bottom), N, C, h, W]
        allg = torch.empty((P,) + tuple(send.shape), dtype=g.dtype, device=g.device)
        comm.all_gather(allg, send, channel=5)
        if r < P - 1:
            core[:, :, -h:, :] += allg[r + 1, 0]   # rank+1's top-halo gradient lands on my bottom rows
        if r > 0:
            core[:, :, :h, :] += allg[r - 1, 1]    # rank-1's bottom-halo gradient lands on my top rows
        return core, None, None


def halo_exchange(x, comm, halo: int = 1):
    """[N, C, h_local, W] -> [N, C, h_local + 2*halo, W] with neighbour rows attached."""
    if halo == 0:
        return x
    if comm.world == 1:
        z = torch.zeros_like(x[:, :, :halo, :])
        return torch.cat([z, x, z], dim=2)
    return _HaloExchange.apply(x, comm, halo)


def halo_conv2d(x, weight, bias, comm, padding: int = 1):
    """Stride-1 convolution of an H-sharded input; returns this rank's rows of the full convolution's output."""
    k = weight.shape[-1]
    assert weight.shape[-2] == k and padding == k // 2, "halo_conv2d: 'same' square kernels (padding = k // 2)"
    xh = halo_exchange(x, comm, padding)
    y = cops.conv2d(xh, weight, bias, stride=1, padding=padding)   # pads H too: 2*padding surplus rows ...
    return y[:, :, padding:-padding, :] if padding else y            # ... dropped here
