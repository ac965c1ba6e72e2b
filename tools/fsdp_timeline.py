"""One-rank device timeline of the headline training step (Llama-2 7B shape, FSDP when world > 1).

    python tools/fsdp_timeline.py [--n-layers 32]                                  # 1 GPU
    torchrun --nproc-per-node N tools/fsdp_timeline.py [--n-layers 32]             # FSDP over N GPUs

Records 2 steps on rank 0 with torch.profiler (CUDA activities), exports the chrome trace and reduces it to the
numbers that say where a step's time goes when communication is supposed to hide behind compute:

  * wall time of a step on the device (first kernel start -> last kernel end, / steps)
  * busy time per stream, and the IDLE time of the compute stream split by the kernel that follows the gap
    (a gap before the first GEMM of a block = waiting for its parameter gather, a gap at the end = the last
    reduce-scatter + optimizer, ...)
  * total duration per kernel family on the compute stream, to compare with the 1-GPU run: the difference is what
    sharing HBM / L2 / SMs with the overlapped copies and optimizer kernels costs the GEMMs
  * copy-engine (memcpy) busy time and bytes

A profiler adds launch overhead; the numbers locate exposed time, the bench numbers stay the ones from bench.py.
Writes gpurun_out/fsdp_timeline_<N>gpu.txt (copy what should be kept to profiles/).
"""
from __future__ import annotations

import argparse
import json
import os
import sys
import tempfile
from collections import defaultdict
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))

import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402
from torch.profiler import ProfilerActivity, profile  # noqa: E402


def family(name: str) -> str:
    for key, fam in (("gemm_tcgen05", "gemm"), ("flash_attn", "attention"), ("attn_delta", "attention"),
                     ("adamw", "adamw"), ("rs_adamw", "rs+adamw"), ("comm_", "comm kernel"), ("rmsnorm", "rmsnorm"),
                     ("swiglu", "swiglu"), ("rope", "rope"), ("xent", "xent"), ("embedding", "embedding"),
                     ("Memcpy", "memcpy"), ("Memset", "memset"), ("at::", "aten elementwise"), ("nccl", "nccl")):
        if key in name:
            return fam
    return "other"


def summarise(trace_path: str, steps: int, out):
    with open(trace_path) as f:
        tr = json.load(f)
    evs = [e for e in tr["traceEvents"] if e.get("ph") == "X" and e.get("cat") in ("kernel", "gpu_memcpy", "gpu_memset")]
    if not evs:
        print("no device events in the trace", file=out)
        return
    evs.sort(key=lambda e: e["ts"])
    t_first, t_last = evs[0]["ts"], max(e["ts"] + e["dur"] for e in evs)
    wall = (t_last - t_first) / 1e3
    by_stream = defaultdict(list)
    for e in evs:
        by_stream[e["args"].get("stream", -1)].append(e)
    # the compute stream is the one with the most GEMM time
    gemm_time = {s: sum(e["dur"] for e in v if "gemm_tcgen05" in e["name"]) for s, v in by_stream.items()}
    main = max(gemm_time, key=gemm_time.get)
    print(f"device wall time: {wall / steps:.2f} ms/step over {steps} steps ({len(evs)} device events)", file=out)
    print("\nbusy time per stream (ms/step):", file=out)
    for s, v in sorted(by_stream.items(), key=lambda kv: -sum(e["dur"] for e in kv[1])):
        busy = sum(e["dur"] for e in v) / 1e3 / steps
        fams = defaultdict(float)
        for e in v:
            fams[family(e["name"])] += e["dur"] / 1e3 / steps
        top = ", ".join(f"{k} {t:.2f}" for k, t in sorted(fams.items(), key=lambda kv: -kv[1])[:5])
        print(f"  stream {s}{' (compute)' if s == main else ''}: {busy:8.2f}   [{top}]", file=out)
    mv = by_stream[main]
    busy_main = sum(e["dur"] for e in mv) / 1e3 / steps
    gaps = defaultdict(lambda: [0.0, 0])
    big = []
    end = mv[0]["ts"]
    for e in mv:
        g = e["ts"] - end
        if g > 5.0:   # > 5 us of idle compute stream
            fam = family(e["name"])
            gaps[fam][0] += g / 1e3 / steps
            gaps[fam][1] += 1
            big.append((g / 1e3, e["name"][:70]))
        end = max(end, e["ts"] + e["dur"])
    idle = wall / steps - busy_main
    print(f"\ncompute stream: busy {busy_main:.2f} ms/step, idle {idle:.2f} ms/step", file=out)
    print("idle time by the kernel family that follows the gap (gaps > 5 us):", file=out)
    for fam, (t, n) in sorted(gaps.items(), key=lambda kv: -kv[1][0]):
        print(f"  before {fam:18s} {t:8.3f} ms/step  in {n // steps} gaps/step", file=out)
    print("largest single gaps (ms, next kernel):", file=out)
    for g, n in sorted(big, reverse=True)[:12]:
        print(f"  {g:7.3f}  {n}", file=out)
    print("\nkernel families on the compute stream (ms/step, launches/step):", file=out)
    fams = defaultdict(lambda: [0.0, 0])
    for e in mv:
        f = fams[family(e["name"])]
        f[0] += e["dur"] / 1e3 / steps
        f[1] += 1
    for fam, (t, n) in sorted(fams.items(), key=lambda kv: -kv[1][0]):
        print(f"  {fam:18s} {t:8.2f}   x{n // steps}", file=out)
    mc = [e for e in evs if e.get("cat") == "gpu_memcpy"]
    if mc:
        nbytes = sum(e["args"].get("bytes", 0) for e in mc)
        print(f"\ncopy engines: {sum(e['dur'] for e in mc) / 1e3 / steps:.2f} ms/step busy, {nbytes / steps / 1e9:.2f} GB/step", file=out)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--n-layers", type=int, default=32)
    ap.add_argument("--seq", type=int, default=4096)
    ap.add_argument("--steps", type=int, default=2)
    ap.add_argument("--kernel-gather", action="store_true")
    ap.add_argument("--kernel-reduce", action="store_true")
    ap.add_argument("--fsdp-ctas", type=int, default=64)
    ap.add_argument("--nvls-min-world", type=int, default=None)
    ap.add_argument("--no-nvls-reduce", action="store_true")
    a = ap.parse_args()
    rank, world, local = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)
    from b200dist.engine import Trainer
    from b200dist.models.llama2 import Transformer
    from b200dist.ops import FusedAdamW

    torch.manual_seed(0)
    model = Transformer.from_name("llama2-7b", device=dev, dtype=torch.bfloat16, n_layers=a.n_layers, max_seq_len=a.seq)
    if world > 1:
        from b200dist.comm import get_communicator
        from b200dist.parallel.fsdp import FSDP

        if a.nvls_min_world is not None:
            get_communicator(None).nvls_min_world = a.nvls_min_world
        model = FSDP(model, nvls_reduce=not a.no_nvls_reduce, comm_max_ctas=a.fsdp_ctas, sync_module_states=False, copy_engine_gather=not a.kernel_gather,
                     copy_engine_reduce=not a.kernel_reduce)
        opt = model.build_optimizer("adamw", lr=3e-4, betas=(0.9, 0.95), weight_decay=0.1)
    else:
        opt = FusedAdamW(model.parameters(), lr=3e-4, betas=(0.9, 0.95), weight_decay=0.1, overlap_backward=True)
    tr = Trainer(model, opt)
    tok = torch.randint(0, 32000, (1, a.seq + 1), device=dev)
    for _ in range(3):
        tr.train_step(tok[:, :-1], tok[:, 1:])
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier(device_ids=[local])
    if rank == 0:
        with profile(activities=[ProfilerActivity.CUDA]) as prof:
            for _ in range(a.steps):
                tr.train_step(tok[:, :-1], tok[:, 1:])
            torch.cuda.synchronize()
        path = os.path.join(tempfile.gettempdir(), f"b200dist_trace_{os.getpid()}.json")
        prof.export_chrome_trace(path)
        out_dir = ROOT / "gpurun_out"
        out_dir.mkdir(exist_ok=True)
        with open(out_dir / f"fsdp_timeline_{world}gpu.txt", "w") as f:
            print(f"# llama2-7b shape, {a.n_layers} layers, seq {a.seq}, world {world}, rank 0; "
                  f"gather={'kernel' if a.kernel_gather else 'copy-engine'} reduce={'nvls' if getattr(model, 'nvls_grads', False) else ('kernel' if a.kernel_reduce else 'copy-engine')}",
                  file=f)
            summarise(path, a.steps, f)
        print((out_dir / f"fsdp_timeline_{world}gpu.txt").read_text())
        os.unlink(path)
    else:
        for _ in range(a.steps):
            tr.train_step(tok[:, :-1], tok[:, 1:])
        torch.cuda.synchronize()
    if world > 1:
        dist.barrier(device_ids=[local])
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
