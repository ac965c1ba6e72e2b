"""Parallel strategies: DDP, FSDP, TP/SP, PP, hybrid meshes, long-context attention."""
from .ddp import DDP, DistributedDataParallel
from .fsdp import FSDP, ShardedOptimizer
from .mesh import DeviceMesh, init_device_mesh

__all__ = ["DDP", "DistributedDataParallel", "FSDP", "ShardedOptimizer", "DeviceMesh", "init_device_mesh"]
