"""Profiler wrapper: torch.profiler context manager with the reference's schedule and a
rank-0 TensorBoard trace, plus NVTX ranges and CUDA-event timers used by the benchmarks
(reference: /root/reference/utils/profiling.py:25-86)."""
from __future__ import annotations

import contextlib
import os
from typing import Callable, Optional

import torch


@contextlib.contextmanager
def training_profiler(output_dir: str = "./profiler_output", wait: int = 1, warmup: int = 1,
                      active: int = 3, repeat: int = 1, rank: Optional[int] = None,
                      enabled: bool = True):
    """``with training_profiler(...) as prof: ...; prof.step()`` once per iteration.

    CPU+CUDA activities, shapes/memory/stacks recorded; only rank 0 writes the trace.
    Yields ``None`` when ``enabled`` is false so call sites need no branching.
    """
    if not enabled:
        yield None
        return
    from torch.profiler import ProfilerActivity, profile, schedule, tensorboard_trace_handler

    if rank is None:
        rank = int(os.environ.get("RANK", "0"))
    acts = [ProfilerActivity.CPU]
    if torch.cuda.is_available():
        acts.append(ProfilerActivity.CUDA)
    handler = None
    if rank == 0:
        os.makedirs(output_dir, exist_ok=True)
        handler = tensorboard_trace_handler(output_dir)
    with profile(activities=acts, schedule=schedule(wait=wait, warmup=warmup, active=active, repeat=repeat),
                 on_trace_ready=handler, record_shapes=True, profile_memory=True, with_stack=True) as prof:
        yield prof


def print_profiler_summary(prof, rank: int = 0, top_n: int = 10) -> Optional[str]:
    """Top-N table sorted by device time on rank 0 (reference utils/profiling.py:69-86)."""
    if prof is None or rank != 0:
        return None
    key = "cuda_time_total" if torch.cuda.is_available() else "cpu_time_total"
    table = prof.key_averages().table(sort_by=key, row_limit=top_n)
    print(table, flush=True)
    return table


@contextlib.contextmanager
def nvtx_range(name: str):
    """NVTX range around a fused kernel / collective (no-op without CUDA)."""
    if torch.cuda.is_available():
        torch.cuda.nvtx.range_push(name)
        try:
            yield
        finally:
            torch.cuda.nvtx.range_pop()
    else:
        yield


class CudaTimer:
    """CUDA-event timer on the current stream; ``elapsed_ms`` synchronises."""

    def __init__(self) -> None:
        self.start_ev = torch.cuda.Event(enable_timing=True)
        self.stop_ev = torch.cuda.Event(enable_timing=True)

    def start(self) -> None:
        self.start_ev.record()

    def stop(self) -> None:
        self.stop_ev.record()

    def elapsed_ms(self) -> float:
        self.stop_ev.synchronize()
        return self.start_ev.elapsed_time(self.stop_ev)


def time_cuda(fn: Callable[[], None], iters: int = 10, warmup: int = 3, flush_l2: bool = True) -> float:
    """Median device time of ``fn`` in ms, with an L2 flush (256 MiB write) between iterations."""
    for _ in range(warmup):
        fn()
    torch.cuda.synchronize()
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda") if flush_l2 else None
    ts = []
    for _ in range(iters):
        if flush is not None:
            flush.zero_()
        t = CudaTimer()
        t.start()
        fn()
        t.stop()
        ts.append(t.elapsed_ms())
    ts.sort()
    return ts[len(ts) // 2]
