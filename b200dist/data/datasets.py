"""Synthetic datasets (there is no network for real ones; the reference's are synthetic too).

* ``ERA5Dataset``           — weather-like (C = vars x levels, lat, lon) pairs
                               (reference multinode_ddp_unet.py:145-164; ViT variant 64x128, tensor_parallel_vit.py:56-75)
* ``ToyRegressionDataset``  — 20-feature regression  (MyTrainDataset, multinode_ddp_basic.py:89-105)
* ``SimpleDataset``         — 10-feature binary classification (distributed_dataloader.py:143-156)
* ``SyntheticImageDataset`` — ImageNet-shaped images (main.py:268-271 uses one fixed random batch)
* ``RandomTokenDataset``    — LM tokens (01_fsdp_tp_hybrid.py:173-174, 03_pipeline_training.py:221-230)

Unlike the reference's ``__getitem__``-time ``torch.randn``, samples are a pure function of
(seed, index), so every epoch and every worker sees reproducible data.
"""
from __future__ import annotations

import torch
from torch.utils.data import Dataset


def _gen(seed: int, idx: int) -> torch.Generator:
    g = torch.Generator()
    g.manual_seed((seed * 1_000_003 + idx) % (2 ** 63 - 1))
    return g


class ERA5Dataset(Dataset):
    def __init__(self, num_samples=1000, num_variables=5, num_levels=13, lat=181, lon=360, seed=0):
        self.num_samples, self.channels, self.lat, self.lon, self.seed = num_samples, num_variables * num_levels, lat, lon, seed

    def __len__(self):
        return self.num_samples

    def __getitem__(self, idx):
        g = _gen(self.seed, idx)
        return (torch.randn(self.channels, self.lat, self.lon, generator=g),
                torch.randn(self.channels, self.lat, self.lon, generator=g))


class ToyRegressionDataset(Dataset):
    def __init__(self, size=2048, in_features=20, seed=0):
        g = _gen(seed, 0)
        self.x, self.y = torch.rand(size, in_features, generator=g), torch.rand(size, 1, generator=g)

    def __len__(self):
        return self.x.shape[0]

    def __getitem__(self, idx):
        return self.x[idx], self.y[idx]


class SimpleDataset(Dataset):
    def __init__(self, size=1000, in_features=10, num_classes=2, seed=0):
        g = _gen(seed, 0)
        self.x = torch.randn(size, in_features, generator=g)
        self.y = torch.randint(0, num_classes, (size,), generator=g)

    def __len__(self):
        return self.x.shape[0]

    def __getitem__(self, idx):
        return self.x[idx], self.y[idx]


class SyntheticImageDataset(Dataset):
    def __init__(self, size=1024, image_size=224, num_classes=1000, channels=3, seed=0):
        self.size, self.image_size, self.num_classes, self.channels, self.seed = size, image_size, num_classes, channels, seed

    def __len__(self):
        return self.size

    def __getitem__(self, idx):
        g = _gen(self.seed, idx)
        return (torch.rand(self.channels, self.image_size, self.image_size, generator=g),
                int(torch.randint(0, self.num_classes, (1,), generator=g)))


class RandomTokenDataset(Dataset):
    """``dp_rank``-seeded so every TP rank of a DP group draws identical tokens (reference seeds with i + dp_rank)."""

    def __init__(self, size=1024, seq_len=256, vocab_size=32000, seed=0):
        self.size, self.seq_len, self.vocab_size, self.seed = size, seq_len, vocab_size, seed

    def __len__(self):
        return self.size

    def __getitem__(self, idx):
        t = torch.randint(0, self.vocab_size, (self.seq_len + 1,), generator=_gen(self.seed, idx))
        return t[:-1], t[1:]
