"""Shared bootstrap for the recipe scripts: import path, device choice, CUDA-event timing."""
from __future__ import annotations

import sys
import time
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
if str(ROOT) not in sys.path:
    sys.path.insert(0, str(ROOT))

import torch  # noqa: E402


def setup(backend: str = "nccl", verbose: bool = True):
    """-> (rank, world, local_rank, device).  Works under torchrun / mpirun / PALS or as a single process, on GPU or CPU."""
    from b200dist.utils import init_distributed

    rank, world, local = init_distributed(backend=backend, verbose=verbose)
    device = torch.device("cuda", torch.cuda.current_device()) if torch.cuda.is_available() else torch.device("cpu")
    return rank, world, local, device


class StepTimer:
    """Device-side timing with CUDA events (host clock on CPU); ``max_over_ranks`` for multi-GPU numbers."""

    def __init__(self, device):
        self.cuda = device.type == "cuda"
        self.times = []

    def __enter__(self):
        if self.cuda:
            self.e0, self.e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            self.e0.record()
        else:
            self.t0 = time.perf_counter()
        return self

    def __exit__(self, *a):
        if self.cuda:
            self.e1.record()
            self.e1.synchronize()
            self.times.append(self.e0.elapsed_time(self.e1) / 1e3)
        else:
            self.times.append(time.perf_counter() - self.t0)
        return False

    def mean(self, skip: int = 0) -> float:
        t = self.times[skip:] or self.times
        return sum(t) / max(1, len(t))


def max_over_ranks(value: float, device) -> float:
    import torch.distributed as dist

    if not (dist.is_available() and dist.is_initialized()):
        return value
    t = torch.tensor([value], dtype=torch.float64, device=device if device.type == "cuda" else "cpu")
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())
