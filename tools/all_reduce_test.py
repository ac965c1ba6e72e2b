#!/usr/bin/env python
"""The reference's older all-reduce / broadcast timing script (tests/all_reduce_test.py): fp32 tensors of 10^3 ... 10^8
elements, 2 warm-up + 20 timed runs each, one plain-text line per (operation, size) appended to
``benchmark_results.log``.  Here both data planes are timed - the peer-memory kernels and torch.distributed (NCCL / gloo)
- on the device (CUDA events, max over ranks) rather than with the host clock around barriers (:60-100).  The full CSV
harness is bench/comm_bench.py.
    scripts/launch.sh 8 tools/all_reduce_test.py [--backend nccl|gloo] [--output benchmark_results.log]"""
import argparse
import datetime
import sys
import time
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import torch
import torch.distributed as dist

from b200dist.comm import get_communicator
from b200dist.utils import cleanup_distributed, init_distributed

ap = argparse.ArgumentParser()
ap.add_argument("--backend", default="nccl")
ap.add_argument("--output", default="benchmark_results.log")
ap.add_argument("--max-exp", type=int, default=8)
a = ap.parse_args()
rank, world, local = init_distributed(backend=a.backend, verbose=False)
cuda = torch.cuda.is_available()
dev = torch.device("cuda", torch.cuda.current_device()) if cuda else torch.device("cpu")
comm = get_communicator()


def timed(fn, warmup=2, runs=20):
    out = []
    for i in range(warmup + runs):
        if cuda:
            torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        if cuda:
            s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            s.record(); fn(); e.record(); torch.cuda.synchronize()
            dt = s.elapsed_time(e) / 1e3
        else:
            t0 = time.perf_counter(); fn(); dt = time.perf_counter() - t0
        if world > 1:
            tt = torch.tensor([dt], device=dev, dtype=torch.float64)
            dist.all_reduce(tt, op=dist.ReduceOp.MAX)
            dt = float(tt)
        if i >= warmup:
            out.append(dt)
    return sum(out) / len(out)


lines = []
for exp in range(3, a.max_exp + 1):
    n = 10 ** exp
    n8 = (n + 7) // 8 * 8
    x = torch.ones(n8, device=dev)
    for op, own, ref in (("all_reduce", lambda: comm.all_reduce(x), lambda: dist.all_reduce(x) if world > 1 else None),
                         ("broadcast", lambda: comm.broadcast(x, 0), lambda: dist.broadcast(x, 0) if world > 1 else None)):
        x.fill_(1.0)
        t_own = timed(own)
        x.fill_(1.0)
        t_ref = timed(ref)
        lines.append(f"{datetime.datetime.now().isoformat(timespec='seconds')} world={world} {op} elements={n} bytes={4 * n} "
                     f"peer_memory={t_own * 1e3:.4f} ms torch_distributed[{dist.get_backend() if world > 1 else 'none'}]={t_ref * 1e3:.4f} ms")
if rank == 0:
    with open(a.output, "a") as f:
        f.write("\n".join(lines) + "\n")
    print("\n".join(lines))
cleanup_distributed()
