"""Spawn-based multi-process harness (gloo on CPU, nccl on GPUs) for distributed unit tests."""
from __future__ import annotations

import os
import socket
import sys
import traceback
from pathlib import Path

import torch
import torch.multiprocessing as mp

ROOT = Path(__file__).resolve().parent.parent


def free_port() -> int:
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, backend, fn, args, errq, init_pg=True):
    os.environ.update({"RANK": str(rank), "WORLD_SIZE": str(world), "LOCAL_RANK": str(rank),
                       "MASTER_ADDR": "127.0.0.1", "MASTER_PORT": str(port), "B200DIST_DISABLE_MPI4PY": "1"})
    if str(ROOT) not in sys.path:
        sys.path.insert(0, str(ROOT))
    torch.set_num_threads(1)
    import torch.distributed as dist

    try:
        if not init_pg:
            pass  # the test exercises its own process-group initialisation
        elif backend == "nccl":
            torch.cuda.set_device(rank)
            dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", rank))
        else:
            dist.init_process_group("gloo", rank=rank, world_size=world)
        fn(rank, world, *args)
        if dist.is_initialized():
            dist.barrier()
    except Exception:  # noqa: BLE001
        errq.put((rank, traceback.format_exc()))
    finally:
        if dist.is_initialized():
            dist.destroy_process_group()


def run_distributed(fn, world: int = 2, backend: str = "gloo", args=(), timeout: float = 180.0, init_pg: bool = True):
    """Run ``fn(rank, world, *args)`` in ``world`` processes; raise if any rank failed."""
    ctx = mp.get_context("spawn")
    errq = ctx.Queue()
    port = free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, backend, fn, args, errq, init_pg)) for r in range(world)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(timeout)
    hung = [p for p in procs if p.is_alive()]
    for p in hung:
        p.kill()
    errs = []
    while not errq.empty():
        errs.append(errq.get())
    if hung:
        raise AssertionError(f"{len(hung)} rank(s) hung past {timeout}s; errors so far: {errs}")
    bad = [p.exitcode for p in procs if p.exitcode != 0]
    if errs or bad:
        raise AssertionError("distributed test failed:\n" + "\n".join(f"[rank {r}]\n{tb}" for r, tb in errs) + f"\nexit codes {bad}")
