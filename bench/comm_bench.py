#!/usr/bin/env python
"""Collective bandwidth benchmark — superset of /root/reference/tests/torch_comm_bench.py.

Same flags (--backend --tensor_sizes --warmup_runs --benchmark_runs --output_file), same CSV columns
(backend,operation,tensor_elements,tensor_bytes,tensor_mb,warmup_iterations,benchmark_iterations,warmup_mean_sec,
warmup_std_sec,benchmark_mean_sec,benchmark_std_sec,benchmark_min_sec,benchmark_max_sec,bandwidth_gbps), same bus-bandwidth
formula for all-reduce (2(n-1)/n * bytes / t), appended to results/benchmark_results.log.  Added: all_gather,
reduce_scatter, send/recv; sizes 1 KB - 1 GB (fp32 elements 256 ... 268435456); DEVICE timing with CUDA events, max over
ranks (the reference times with the host clock around barriers, :40-56); two implementations side by side:
``--impl b200dist`` (peer-memory kernels) and ``--impl nccl`` (torch.distributed, the reference path); and the
``--backend`` flag really selects the process-group backend (in the reference it only moves the tensor, :16 vs :258).
Bandwidth is reported in GB/s = 1e9 bytes/s.

    torchrun --nproc-per-node 8 --master-addr 127.0.0.1 bench/comm_bench.py --impl both
"""
from __future__ import annotations

import argparse
import datetime
import os
import statistics
import sys
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))

import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

COLUMNS = ("backend,operation,tensor_elements,tensor_bytes,tensor_mb,warmup_iterations,benchmark_iterations,"
           "warmup_mean_sec,warmup_std_sec,benchmark_mean_sec,benchmark_std_sec,benchmark_min_sec,benchmark_max_sec,"
           "bandwidth_gbps")
DEFAULT_SIZES = [256 * 4 ** i for i in range(11)]  # 1 KB ... 1 GB of fp32


def timed(fn, n, device, world):
    import gc

    out = []
    gc.collect()
    gc.disable()        # a collector pause between the two events of one iteration would be charged to the collective
    try:
        return _timed(fn, n, device, world, out)
    finally:
        gc.enable()


def _timed(fn, n, device, world, out):
    for _ in range(n):
        if device.type == "cuda":
            torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        if device.type == "cuda":
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(); fn(); e1.record(); e1.synchronize()
            t = e0.elapsed_time(e1) / 1e3
        else:
            import time

            t0 = time.perf_counter(); fn(); t = time.perf_counter() - t0
        if world > 1:
            tt = torch.tensor([t], dtype=torch.float64, device=device)
            dist.all_reduce(tt, op=dist.ReduceOp.MAX)
            t = float(tt)
        out.append(t)
    return out


def bus_bandwidth(op, nbytes, t, n):
    if op == "all_reduce":
        return 2 * (n - 1) / n * nbytes / t / 1e9
    if op in ("all_gather", "reduce_scatter"):
        return (n - 1) / n * nbytes / t / 1e9     # nbytes = full (gathered / pre-scatter) size
    return nbytes / t / 1e9                        # broadcast, send_recv


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--backend", default="nccl", choices=["nccl", "gloo"])
    ap.add_argument("--tensor_sizes", type=int, nargs="+", default=DEFAULT_SIZES, help="fp32 elements per message")
    ap.add_argument("--warmup_runs", type=int, default=5)
    ap.add_argument("--benchmark_runs", type=int, default=20)
    ap.add_argument("--output_file", default="results/benchmark_results.log")
    ap.add_argument("--impl", default="both", choices=["b200dist", "nccl", "both"])
    ap.add_argument("--no-registered", dest="registered", action="store_false",
                    help="skip the b200dist-registered rows (tensors placed inside the NVSwitch multicast buffer)")
    ap.add_argument("--ops", nargs="+", default=["broadcast", "all_reduce", "all_gather", "reduce_scatter", "send_recv"])
    ap.add_argument("--max_bytes", type=int, default=1 << 30)
    a = ap.parse_args()
    from b200dist.utils import cleanup_distributed, init_distributed

    rank, world, local = init_distributed(backend=a.backend, verbose=False)
    device = torch.device("cuda", torch.cuda.current_device()) if (torch.cuda.is_available() and a.backend == "nccl") else torch.device("cpu")
    from b200dist.comm import Communicator

    impls = ["b200dist", "nccl"] if a.impl == "both" else [a.impl]
    if "b200dist" in impls and a.registered and device.type == "cuda":
        # second arm of ours: the tensor a collective works on lives in "registered" memory (the NVSwitch multicast buffer;
        # what ncclMemAlloc + user-buffer registration is to NCCL), so no staging copy in / copy out is needed
        impls.insert(1, "b200dist-reg")
    comm = Communicator(None, staging_bytes=min(a.max_bytes, 1 << 30) + (1 << 20)) if "b200dist" in impls else None
    if rank == 0:
        os.makedirs(os.path.dirname(os.path.abspath(a.output_file)), exist_ok=True)
        with open(a.output_file, "a") as f:
            f.write(f"# b200dist comm_bench {datetime.datetime.now().isoformat(timespec='seconds')} world={world} "
                    f"backend={a.backend} device={torch.cuda.get_device_name() if device.type == 'cuda' else 'cpu'} "
                    f"torch={torch.__version__} timing=cuda-events,max-over-ranks unit=GB/s(1e9)\n# {COLUMNS}\n")
        print(COLUMNS, flush=True)
    for n_el in a.tensor_sizes:
        nbytes = n_el * 4
        if nbytes > a.max_bytes:
            continue
        n_pad = (n_el + 8 * world - 1) // (8 * world) * (8 * world)
        x = torch.ones(n_pad, device=device)
        gathered = torch.empty(n_pad * world, device=device) if nbytes * world <= (4 << 30) else None
        shard = torch.empty(n_pad // world, device=device)
        for impl in impls:
            tag = {"nccl": a.backend, "b200dist": "b200dist-peer", "b200dist-reg": "b200dist-registered"}[impl]
            reg = impl == "b200dist-reg"
            if reg and not (world >= comm.nvls_min_world and nbytes >= comm.nvls_threshold and n_pad % (4 * world) == 0):
                continue            # the switch paths start at nvls_threshold; below it both arms of ours are the same code

            def registered(numel, channel, fill=None):
                t = comm.nvls_tensor(numel, torch.float32, channel=channel)
                if t is not None and fill is not None:
                    t.fill_(fill)
                return t
            for op in a.ops:
                if op == "broadcast":
                    xb = registered(n_pad, 3, 1.0) if reg else x
                    if xb is None:
                        continue
                    fn = (lambda xb=xb: comm.broadcast(xb, 0)) if impl != "nccl" else (lambda: dist.broadcast(x, 0))
                    size = nbytes
                elif op == "all_reduce":
                    xr = registered(n_pad, 0, 1.0) if reg else x
                    if xr is None:
                        continue
                    if reg:
                        fn = (lambda xr=xr: comm.all_reduce(xr, algo=4))
                    else:
                        fn = (lambda: comm.all_reduce(x)) if impl == "b200dist" else (lambda: dist.all_reduce(x))
                    size = nbytes
                elif op == "all_gather":
                    if gathered is None:
                        continue
                    go = registered(n_pad * world, 1) if reg else gathered
                    if go is None:
                        continue
                    fn = (lambda go=go: comm.all_gather(go, x)) if impl != "nccl" else (lambda: dist.all_gather_into_tensor(gathered, x))
                    size = nbytes * world
                elif op == "reduce_scatter":
                    xi = registered(n_pad, 2, 1.0) if reg else x
                    if xi is None:
                        continue
                    fn = (lambda xi=xi: comm.reduce_scatter(shard, xi)) if impl != "nccl" else (lambda: dist.reduce_scatter_tensor(shard, x))
                    size = nbytes
                elif op == "send_recv":
                    if world < 2 or impl != "nccl":
                        continue  # p2p stays on NCCL by design (pipeline stages)

                    def fn():
                        if rank == 0:
                            dist.send(x, 1)
                        elif rank == 1:
                            dist.recv(x, 0)
                    size = nbytes
                else:
                    continue
                if world == 1 and impl == "nccl" and not dist.is_initialized():
                    continue
                wt = timed(fn, a.warmup_runs, device, world)
                bt = timed(fn, a.benchmark_runs, device, world)
                mean = statistics.mean(bt)
                row = (f"{tag},{op},{n_el},{nbytes},{nbytes / 1e6:.3f},{a.warmup_runs},{a.benchmark_runs},"
                       f"{statistics.mean(wt):.9f},{statistics.pstdev(wt):.9f},{mean:.9f},{statistics.pstdev(bt):.9f},"
                       f"{min(bt):.9f},{max(bt):.9f},{bus_bandwidth(op, size, mean, world):.3f}")
                if rank == 0:
                    print(row, flush=True)
                    with open(a.output_file, "a") as f:
                        f.write(row + "\n")
    if world > 1:
        dist.barrier()
    cleanup_distributed()


if __name__ == "__main__":
    main()
