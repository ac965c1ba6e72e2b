#!/usr/bin/env python
"""Point-to-point timing: rank 0 sends a 40 MB tensor to every other rank in turn, 2 warm-up + 20 timed runs, device-timed
(reference tests/send_recv_test.py:36-64, tests/test_torchrun.py).  p2p stays on NCCL (pipeline stages use it)."""
import sys
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import torch
import torch.distributed as dist

from b200dist.utils import cleanup_distributed, init_distributed

rank, world, local = init_distributed(verbose=False)
dev = torch.device("cuda", torch.cuda.current_device()) if torch.cuda.is_available() else torch.device("cpu")
x = torch.zeros((10000, 1000), device=dev)
times = []
for it in range(22):
    if dev.type == "cuda":
        torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    import time

    t0 = time.perf_counter()
    if rank == 0:
        for dst in range(1, world):
            dist.send(x, dst)
    elif world > 1:
        dist.recv(x, 0)
    if dev.type == "cuda":
        torch.cuda.synchronize()
    if it >= 2:
        times.append(time.perf_counter() - t0)
if rank == 0 and times:
    m = sum(times) / len(times)
    print(f"send/recv 40 MB rank0 -> {world - 1} peer(s): {m * 1e3:.3f} ms  ({40e6 * max(1, world - 1) / m / 1e9:.1f} GB/s)")
cleanup_distributed()
