#!/usr/bin/env python
"""Single-process smoke test: versions, GPUs, then a world-size-1 collective (reference tests/test_env.py:54-79)."""
import os
import sys
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import torch
import torch.distributed as dist

print(f"torch {torch.__version__} cuda {torch.version.cuda} available={torch.cuda.is_available()} devices={torch.cuda.device_count()}")
for i in range(torch.cuda.device_count()):
    print(f"  cuda:{i} {torch.cuda.get_device_name(i)}")
os.environ.setdefault("MASTER_ADDR", "127.0.0.1"); os.environ.setdefault("MASTER_PORT", "29517")
backend = "nccl" if torch.cuda.is_available() else "gloo"
dist.init_process_group(backend, rank=0, world_size=1)
t = torch.ones(4, device="cuda" if torch.cuda.is_available() else "cpu")
dist.all_reduce(t)
assert float(t.sum()) == 4.0
from b200dist.comm import get_communicator

get_communicator(None).all_reduce(t)
print(f"{backend} world-size-1 all_reduce ok; b200dist communicator ok")
dist.destroy_process_group()
