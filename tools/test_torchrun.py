#!/usr/bin/env python
"""Smallest possible launcher check: rank 0 sends a one-element tensor to every other rank (reference tests/test_torchrun.py:34-60).
    scripts/launch.sh 2 tools/test_torchrun.py      (gloo on a CPU-only machine)"""
import sys
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import torch
import torch.distributed as dist

from b200dist.utils import cleanup_distributed, init_distributed

rank, world, local = init_distributed(verbose=False)
dev = torch.device("cuda", torch.cuda.current_device()) if torch.cuda.is_available() else torch.device("cpu")
t = torch.zeros(1, device=dev)
if world == 1:
    print("single process: nothing to send")
elif rank == 0:
    t += 42
    for dst in range(1, world):
        dist.send(t, dst)
    print(f"rank 0 sent {t.item():.0f} to {world - 1} rank(s)")
else:
    dist.recv(t, 0)
    assert t.item() == 42, t
    print(f"rank {rank} has data {t.item():.0f}")
cleanup_distributed()
