"""Parallel strategies: DDP, FSDP, TP/SP, PP, hybrid meshes, long-context attention."""
from .ddp import DDP, DistributedDataParallel
from .fsdp import FSDP, ShardedOptimizer
from .halo import halo_conv2d, halo_exchange
from .mesh import DeviceMesh, init_device_mesh
from .ring_attention import ring_attention, zigzag_ring_attention, zigzag_shard, zigzag_unshard
from .ulysses import ulysses_attention

__all__ = ["DDP", "DistributedDataParallel", "FSDP", "ShardedOptimizer", "DeviceMesh", "init_device_mesh", "halo_conv2d", "halo_exchange",
           "ring_attention", "zigzag_ring_attention", "zigzag_shard", "zigzag_unshard", "ulysses_attention"]
