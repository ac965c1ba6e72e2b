"""Distributed sampler that is aware of the DATA-parallel rank (not the global rank).

``torch.utils.data.DistributedSampler`` semantics as the reference uses it
(/root/reference/scripts/01_data_parallel_ddp/distributed_dataloader.py:206-215,248: ``set_epoch``,
shuffle, pad-or-drop), with one fix: ranks of the same tensor-/pipeline-parallel group must see
the SAME samples, so sharding is by ``dp_rank`` / ``dp_size`` (the reference ViT recipe shards by
global rank, tensor_parallel_vit.py:303-305 — SURVEY.md §9 "latent defects").
"""
from __future__ import annotations

import math
from typing import Iterator, Optional

import torch
import torch.distributed as dist
from torch.utils.data import Sampler


class DistributedSampler(Sampler[int]):
    def __init__(self, dataset, num_replicas: Optional[int] = None, rank: Optional[int] = None, shuffle: bool = True,
                 seed: int = 0, drop_last: bool = False, mesh=None, dp_dim: str = "dp"):
        if mesh is not None:  # shard over the data-parallel mesh dim only
            num_replicas, rank = mesh.size(dp_dim), mesh.get_local_rank(dp_dim)
        if num_replicas is None:
            num_replicas = dist.get_world_size() if dist.is_available() and dist.is_initialized() else 1
        if rank is None:
            rank = dist.get_rank() if dist.is_available() and dist.is_initialized() else 0
        if not 0 <= rank < num_replicas:
            raise ValueError(f"invalid rank {rank} for {num_replicas} replicas")
        self.dataset, self.num_replicas, self.rank = dataset, num_replicas, rank
        self.shuffle, self.seed, self.drop_last, self.epoch = shuffle, seed, drop_last, 0
        n = len(dataset)
        if drop_last and n % num_replicas:
            self.num_samples = math.ceil((n - num_replicas) / num_replicas)
        else:
            self.num_samples = math.ceil(n / num_replicas)
        self.total_size = self.num_samples * num_replicas

    def __iter__(self) -> Iterator[int]:
        n = len(self.dataset)
        if self.shuffle:
            g = torch.Generator()
            g.manual_seed(self.seed + self.epoch)
            idx = torch.randperm(n, generator=g).tolist()
        else:
            idx = list(range(n))
        if not self.drop_last:
            pad = self.total_size - len(idx)
            if pad > 0:
                idx += (idx * math.ceil(pad / max(1, len(idx))))[:pad]
        else:
            idx = idx[: self.total_size]
        return iter(idx[self.rank:self.total_size:self.num_replicas])

    def __len__(self) -> int:
        return self.num_samples

    def set_epoch(self, epoch: int) -> None:
        """Call at the start of every epoch so shuffling differs per epoch and agrees across ranks."""
        self.epoch = epoch
