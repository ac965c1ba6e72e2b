"""Toy models of the reference recipes (SURVEY.md §2.1 "Toy models")."""
from __future__ import annotations

from torch import nn

from .layers import Linear, ReLU


class ToyModel(nn.Module):
    """64 -> 64 -> 64  (fsdp_tp/tensor_parallel_example.py:76-84)."""

    def __init__(self, dim=64):
        super().__init__()
        self.in_proj, self.relu, self.out_proj = Linear(dim, dim), ReLU(), Linear(dim, dim)

    def forward(self, x):
        return self.out_proj(self.relu(self.in_proj(x)))


class ToyMLP(nn.Module):
    """16 -> 64 -> 16  (scripts/03_tensor_parallel_tp/02_basic_tensor_parallel.py:46-54)."""

    def __init__(self, d_in=16, d_hidden=64):
        super().__init__()
        self.in_proj, self.relu, self.out_proj = Linear(d_in, d_hidden), ReLU(), Linear(d_hidden, d_in)

    def forward(self, x):
        return self.out_proj(self.relu(self.in_proj(x)))


class FourBlockMLP(nn.Module):
    """Four Linear+ReLU blocks of width 512  (scripts/04_pipeline_parallel_pp/02_pipeline_schedules.py:45-60)."""

    def __init__(self, dim=512):
        super().__init__()
        self.block0 = nn.Sequential(Linear(dim, dim), ReLU())
        self.block1 = nn.Sequential(Linear(dim, dim), ReLU())
        self.block2 = nn.Sequential(Linear(dim, dim), ReLU())
        self.block3 = nn.Sequential(Linear(dim, dim), ReLU())

    def forward(self, x):
        return self.block3(self.block2(self.block1(self.block0(x))))

    def stage(self, index: int, num_stages: int = 4) -> nn.Module:
        blocks = [self.block0, self.block1, self.block2, self.block3]
        per = len(blocks) // num_stages
        return nn.Sequential(*blocks[index * per:(index + 1) * per])


class StageModule(nn.Module):
    """One manual pipeline stage: Linear + ReLU  (01_manual_model_split.py:38-50; 128-256-256-256-64)."""

    def __init__(self, d_in, d_out, last=False):
        super().__init__()
        self.fc, self.act = Linear(d_in, d_out), (nn.Identity() if last else ReLU())

    def forward(self, x):
        return self.act(self.fc(x))


class SimpleModel(nn.Module):
    """10 -> 64 -> 2 classifier  (scripts/01_data_parallel_ddp/distributed_dataloader.py:160-172)."""

    def __init__(self, d_in=10, d_hidden=64, n_cls=2):
        super().__init__()
        self.fc1, self.relu, self.fc2 = Linear(d_in, d_hidden), ReLU(), Linear(d_hidden, n_cls)

    def forward(self, x):
        return self.fc2(self.relu(self.fc1(x)))


def regression_head(d_in=20):
    """Linear(20, 1)  (multinode_ddp_basic.py:213)."""
    return Linear(d_in, 1)
