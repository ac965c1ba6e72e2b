"""Domain parallelism: a spatially sharded convolution with halo exchange.

Docs-only in the reference (/root/reference/docs/guide/10_domain_parallel.md:45-149: splitting the height dimension
breaks a 3x3 convolution at the cut unless each rank first receives a one-row halo from its neighbours; the guide points
to PhysicsNeMo ShardTensor).  Here the input is sharded along H; ``halo_exchange`` attaches ``halo`` boundary rows of
the neighbouring ranks (zeros at the global border) with one neighbour exchange - each rank reads two strips from its two
neighbours over peer memory, whatever the number of ranks - and its
backward returns the halo gradients to their owners.  ``halo_conv2d`` then equals the corresponding rows of the
unsharded convolution exactly.
"""
from __future__ import annotations

import torch
from torch.autograd import Function

from ..ops import conv as cops


def _exchange(x, comm, halo):
    """-> (top, bottom): halo rows received from rank-1 / rank+1 (zeros at the domain border).  Two neighbour transfers
    per rank (``Communicator.neighbor_exchange``), independent of the number of ranks."""
    my_top, my_bottom = x[:, :, :halo, :].contiguous(), x[:, :, -halo:, :].contiguous()
    # the rank above needs my top rows as ITS bottom halo; the rank below needs my bottom rows as ITS top halo
    return comm.neighbor_exchange(my_top, my_bottom)


class _HaloExchange(Function):
    @staticmethod
    def forward(ctx, x, comm, halo):
        ctx.comm, ctx.halo = comm, halo
        top, bottom = _exchange(x, comm, halo)
        return torch.cat([top, x, bottom], dim=2)

    @staticmethod
    def backward(ctx, g):
        comm, h = ctx.comm, ctx.halo
        core = g[:, :, h:-h, :].clone()
        # the gradient of my top halo belongs to rank-1's bottom rows, that of my bottom halo to rank+1's top rows:
        # the same neighbour exchange, with the halo gradients as payload
        from_prev, from_next = comm.neighbor_exchange(g[:, :, :h, :].contiguous(), g[:, :, -h:, :].contiguous())
        core[:, :, :h, :] += from_prev      # rank-1's bottom-halo gradient lands on my top rows (zeros on rank 0)
        core[:, :, -h:, :] += from_next     # rank+1's top-halo gradient lands on my bottom rows (zeros on the last rank)
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
