#!/usr/bin/env python
"""Headline benchmark: Llama-2 7B FSDP full-shard bf16 training, seq 4096, tokens/s.

    python bench.py --gpus N --steps K --warmup W            # this framework
    python bench.py --impl reference --gpus N ...            # unmodified reference (baseline/_ref)

For N > 1 the driver launches this file under ``python -m torch.distributed.run``; one rank per
GPU.  Rank 0 prints ONE JSON line (see the field list in the task contract).  Synthetic tokens,
random-init weights, weak scaling (1 x 4096 tokens per GPU).

Timing: W warm-up steps, then K steps bracketed by barrier + cuda.synchronize on both sides,
CUDA events on the launching stream, MAX over ranks.  A training step streams > 100 GB of
weights/optimizer state/activations through HBM, far more than the 126 MB L2, so no explicit L2
flush is needed between iterations ("inputs larger than L2").  SM clocks and throttle reasons are
sampled through NVML during the timed region.
"""
from __future__ import annotations

import argparse
import json
import os
import sys
import threading
import time
from pathlib import Path

ROOT = Path(__file__).resolve().parent
sys.path.insert(0, str(ROOT))

MODEL_DEFAULT = "llama2-7b"
SEQ_DEFAULT = 4096
# Llama-2 sizes for the reference arm (its ModelArgs fields), so that arm imports nothing from b200dist
REF_MODEL_DIMS = {
    "llama2-7b": dict(dim=4096, n_layers=32, n_heads=32, n_kv_heads=32, vocab_size=32000, multiple_of=256),
    "llama2-13b": dict(dim=5120, n_layers=40, n_heads=40, n_kv_heads=40, vocab_size=32000, multiple_of=256),
    "llama2-tiny": dict(dim=512, n_layers=4, n_heads=8, vocab_size=32000),
}


class ClockSampler:
    """Samples SM clock / throttle reasons every 200 ms through NVML while the timed region runs."""

    def __init__(self, index: int):
        self.index, self.samples, self.reasons = index, [], set()
        self._stop = threading.Event()
        self._thr = None
        self.max_mhz = None
        try:
            import pynvml

            pynvml.nvmlInit()
            self.nv = pynvml
            # honour CUDA_VISIBLE_DEVICES remapping
            vis = os.environ.get("CUDA_VISIBLE_DEVICES")
            phys = int(vis.split(",")[index]) if vis and vis.split(",")[index].isdigit() else index
            self.h = pynvml.nvmlDeviceGetHandleByIndex(phys)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
        except Exception:  # noqa: BLE001
            self.nv = None

    def _run(self):
        nv = self.nv
        names = {
            nv.nvmlClocksThrottleReasonHwSlowdown: "hw_slowdown",
            nv.nvmlClocksThrottleReasonHwThermalSlowdown: "hw_thermal_slowdown",
            nv.nvmlClocksThrottleReasonSwThermalSlowdown: "sw_thermal_slowdown",
            nv.nvmlClocksThrottleReasonSwPowerCap: "sw_power_cap",
            nv.nvmlClocksThrottleReasonHwPowerBrakeSlowdown: "hw_power_brake",
        }
        while not self._stop.is_set():
            try:
                self.samples.append(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM))
                mask = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                for bit, name in names.items():
                    if mask & bit:
                        self.reasons.add(name)
            except Exception:  # noqa: BLE001
                pass
            self._stop.wait(0.2)

    def start(self):
        if self.nv is not None:
            self._thr = threading.Thread(target=self._run, daemon=True)
            self._thr.start()

    def stop(self):
        self._stop.set()
        if self._thr is not None:
            self._thr.join(timeout=2)
        s = sorted(self.samples)
        return {"sm_mhz": (s[len(s) // 2] if s else None), "sm_max_mhz": self.max_mhz,
                "reasons": sorted(self.reasons), "samples": len(s)}


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=8)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200dist", choices=["b200dist", "reference"])
    ap.add_argument("--model", default=MODEL_DEFAULT)
    ap.add_argument("--seq-len", type=int, default=SEQ_DEFAULT)
    ap.add_argument("--batch-per-gpu", type=int, default=1)
    ap.add_argument("--n-layers", type=int, default=None, help="debug only: a reduced model is NOT the headline config")
    ap.add_argument("--ref-wrap", default="block", choices=["block", "root"],
                    help="reference arm: fully_shard per block (idiomatic) or root-only (as the recipe is written)")
    ap.add_argument("--ref-reduce-fp32", action="store_true", help="reference arm: reduce gradients in fp32 instead of bf16")
    ap.add_argument("--lr", type=float, default=3e-4)
    ap.add_argument("--fsdp-ctas", type=int, default=64, help="CTAs per peer-memory collective kernel (FSDP)")
    ap.add_argument("--kernel-gather", action="store_true", help="FSDP all-gather with the SM pull kernel instead of DMA peer copies")
    ap.add_argument("--kernel-reduce", action="store_true", help="FSDP reduce-scatter+AdamW pulling over NVLink from the SMs instead of DMA staging")
    ap.add_argument("--no-overlap-optimizer", action="store_true", help="1 GPU: run AdamW after backward instead of from the gradient hooks")
    ap.add_argument("--no-nvls-reduce", action="store_true", help="FSDP: keep the gradient staging in plain peer memory (DMA / pull reduce) instead of NVSwitch multicast memory")
    ap.add_argument("--nvls-min-world", type=int, default=None, help="smallest world size that uses the NVSwitch (multimem) paths (default 4)")
    ap.add_argument("--no-reshard", action="store_true", help="keep parameters gathered between forward and backward (ZeRO-2 style)")
    return ap.parse_args()


def setup_dist(n_gpus: int):
    import torch
    import torch.distributed as dist

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    if world > 1 and not dist.is_initialized():
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        os.environ.setdefault("MASTER_PORT", "29500")
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    if world != n_gpus:
        raise SystemExit(f"--gpus {n_gpus} but WORLD_SIZE={world}: launch with torchrun --nproc-per-node {n_gpus}")
    return rank, world, local


def max_over_ranks(value: float, world: int) -> float:
    import torch
    import torch.distributed as dist

    if world == 1:
        return value
    t = torch.tensor([value], device="cuda", dtype=torch.float64)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def barrier_sync(world: int):
    import torch
    import torch.distributed as dist

    if world > 1:
        dist.barrier(device_ids=[torch.cuda.current_device()])
    torch.cuda.synchronize()


def timed_loop(step_fn, steps: int, warmup: int, world: int, local: int):
    """Returns (ms_per_step [max over ranks], clocks dict)."""
    import torch

    for i in range(warmup):
        step_fn(i)
    barrier_sync(world)
    sampler = ClockSampler(local)
    sampler.start()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ev0.record()
    for i in range(steps):
        step_fn(warmup + i)
    ev1.record()
    barrier_sync(world)
    clocks = sampler.stop()
    ms = ev0.elapsed_time(ev1)
    return max_over_ranks(ms, world) / steps, clocks


def make_batches(n: int, batch: int, seq: int, vocab: int, rank: int, pinned: bool):
    import torch

    g = torch.Generator().manual_seed(1234 + rank)
    xs, ys = [], []
    for _ in range(n):
        t = torch.randint(0, vocab, (batch, seq + 1), generator=g, dtype=torch.int64)
        x, y = t[:, :-1].contiguous(), t[:, 1:].contiguous()
        if pinned:
            x, y = x.pin_memory(), y.pin_memory()
        xs.append(x)
        ys.append(y)
    return xs, ys


# ------------------------------------------------------------------------------------------------
def run_b200dist(a, rank, world, local):
    import torch

    from b200dist.engine import Trainer
    from b200dist.models.llama2 import Transformer, model_args_for
    from b200dist.ops import FusedAdamW
    from b200dist.ops import _lib

    overrides = {"max_seq_len": a.seq_len}
    if a.n_layers:
        overrides["n_layers"] = a.n_layers
    dev = torch.device("cuda", local)
    torch.manual_seed(0)
    model = Transformer.from_name(a.model, device=dev, dtype=torch.bfloat16, **overrides)
    vocab = model.vocab_size
    if world > 1:
        from b200dist.comm import get_communicator
        from b200dist.parallel.fsdp import FSDP

        if a.nvls_min_world is not None:
            get_communicator(None).nvls_min_world = a.nvls_min_world
        model = FSDP(model, nvls_reduce=not a.no_nvls_reduce, comm_max_ctas=a.fsdp_ctas, reshard_after_forward=not a.no_reshard, sync_module_states=False,
                     copy_engine_gather=not a.kernel_gather, copy_engine_reduce=not a.kernel_reduce)
        opt = model.build_optimizer("adamw", lr=a.lr, betas=(0.9, 0.95), weight_decay=0.1)
    else:
        # one GPU: every parameter is updated from its gradient hook on a side stream while backward continues
        opt = FusedAdamW(model.parameters(), lr=a.lr, betas=(0.9, 0.95), weight_decay=0.1,
                         overlap_backward=not a.no_overlap_optimizer)
    trainer = Trainer(model, opt)
    n_batches = a.steps + a.warmup
    xs, ys = make_batches(n_batches, a.batch_per_gpu, a.seq_len, vocab, rank, pinned=True)
    xd, yd = [x.to(dev) for x in xs], [y.to(dev) for y in ys]

    # ---- device-resident timing (kernel-only number) ----
    counted = {}

    def dev_step(i):
        if i == a.warmup:  # first timed step: start counting our kernel launches
            counted["c0"] = _lib.launch_count
        trainer.train_step(xd[i % n_batches], yd[i % n_batches])

    ms, clocks = timed_loop(dev_step, a.steps, a.warmup, world, local)
    launches = _lib.launch_count - counted.get("c0", _lib.launch_count)
    # ---- end-to-end timing through the public API: pinned-host batch in, loss float out ----
    losses = []
    ms_e2e, _ = timed_loop(lambda i: losses.append(trainer.train_step_host(xs[i % n_batches], ys[i % n_batches])),
                           a.steps, 1, world, local)
    tokens = a.batch_per_gpu * a.seq_len * world
    return {
        "ms": ms, "ms_e2e": ms_e2e, "clocks": clocks, "tokens": tokens, "launches": launches,
        "h2d": 2 * a.batch_per_gpu * a.seq_len * 8, "d2h": 4, "loss_last": losses[-1] if losses else None,
        "params": sum(p.numel() for p in model.parameters()) if world == 1 else None,
    }


def run_reference(a, rank, world, local):
    """Unmodified reference model + stock torch FSDP2 (fully_shard) + torch AdamW, bf16 mixed precision.

    Mirrors /root/reference/scripts/06_hybrid_parallelism/01_fsdp_tp_hybrid.py:102-180 with the model
    size / sequence length / dtype / loss parameterised as BASELINE.md §4 row 3 prescribes.
    """
    from baseline.install_ref import ensure_reference, why_unavailable

    ref_root = ensure_reference()  # installs baseline/_ref on first use when the reference tree is reachable
    if ref_root is None:
        if rank == 0:
            print(json.dumps({"impl": "reference", "unavailable": why_unavailable()}))
        return None
    sys.path.insert(0, str(ref_root))
    sys.path.insert(0, str(ref_root / "reference_src" / "fsdp_tp"))
    import torch
    import torch.distributed as dist
    import torch.nn.functional as F
    from llama2_model import ModelArgs, Transformer  # the reference's own model file
    from torch.distributed.fsdp import MixedPrecisionPolicy, fully_shard

    if not dist.is_initialized():  # world == 1: fully_shard still needs a process group
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        os.environ.setdefault("MASTER_PORT", "29511")
        dist.init_process_group("nccl", rank=0, world_size=1, device_id=torch.device("cuda", local))
    cfg = dict(REF_MODEL_DIMS[a.model])  # nothing of b200dist is imported on this path
    cfg["max_seq_len"] = a.seq_len
    if a.n_layers:
        cfg["n_layers"] = a.n_layers
    dev = torch.device("cuda", local)
    torch.manual_seed(0)
    with torch.device(dev):
        model = Transformer.from_model_args(ModelArgs(**cfg))
    mesh = dist.device_mesh.init_device_mesh("cuda", (world,), mesh_dim_names=("dp",))
    # bf16 parameters AND bf16 gradient reduction: what the repo arm does, and what the reference's own FSDP
    # mixed-precision recipe sets (scripts/02_fully_sharded_fsdp/resnet_fsdp_training.py:200-203)
    mp = MixedPrecisionPolicy(param_dtype=torch.bfloat16,
                              reduce_dtype=torch.float32 if a.ref_reduce_fp32 else torch.bfloat16)
    if a.ref_wrap == "block":
        for layer in model.layers:
            fully_shard(layer, mesh=mesh, mp_policy=mp)
    fully_shard(model, mesh=mesh, mp_policy=mp)
    opt = torch.optim.AdamW(model.parameters(), lr=a.lr, betas=(0.9, 0.95), weight_decay=0.1, foreach=True)
    vocab = cfg["vocab_size"]
    n_batches = a.steps + a.warmup
    xs, ys = make_batches(n_batches, a.batch_per_gpu, a.seq_len, vocab, rank, pinned=True)
    xd, yd = [x.to(dev) for x in xs], [y.to(dev) for y in ys]

    def step(x, y):
        opt.zero_grad()
        out = model(x)
        loss = F.cross_entropy(out.view(-1, vocab), y.view(-1))
        loss.backward()
        opt.step()
        return loss

    ms, clocks = timed_loop(lambda i: step(xd[i % n_batches], yd[i % n_batches]), a.steps, a.warmup, world, local)
    losses = []

    def step_host(i):
        x = xs[i % n_batches].to(dev, non_blocking=True)
        y = ys[i % n_batches].to(dev, non_blocking=True)
        losses.append(float(step(x, y).item()))

    ms_e2e, _ = timed_loop(step_host, a.steps, 1, world, local)
    tokens = a.batch_per_gpu * a.seq_len * world
    return {"ms": ms, "ms_e2e": ms_e2e, "clocks": clocks, "tokens": tokens, "launches": None,
            "h2d": 2 * a.batch_per_gpu * a.seq_len * 8, "d2h": 4, "loss_last": losses[-1] if losses else None,
            "params": None}


def main():
    a = parse()
    rank, world, local = setup_dist(a.gpus)
    try:
        r = run_reference(a, rank, world, local) if a.impl == "reference" else run_b200dist(a, rank, world, local)
    except Exception as e:  # noqa: BLE001
        if a.impl == "reference":
            if rank == 0:
                print(json.dumps({"impl": "reference", "unavailable": f"{type(e).__name__}: {str(e)[:200]}"}))
            return 0
        raise
    if r is None:
        return 0
    if rank == 0:
        tps = r["tokens"] / (r["ms"] / 1e3)
        tps_e2e = r["tokens"] / (r["ms_e2e"] / 1e3)
        line = {
            "metric": "llama2_7b_fsdp_train_tokens_per_s" if a.model == "llama2-7b" and not a.n_layers
            else f"{a.model}{'-L%d' % a.n_layers if a.n_layers else ''}_train_tokens_per_s",
            "value": round(tps, 1), "unit": "tokens/s", "n_gpus": world, "steps": a.steps, "warmup": a.warmup,
            "ms_per_step": round(r["ms"], 3), "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "bf16", "data": "synthetic tokens, random-init weights", "impl": a.impl,
            "config": {"model": a.model + (f"-L{a.n_layers}" if a.n_layers else ""),
                       "global_batch": a.batch_per_gpu * world, "seq_len": a.seq_len,
                       "parallelism": f"fsdp{world}" if world > 1 else "single-gpu (fsdp1)",
                       "optimizer": "adamw (fp32 master + moments)",
                       "l2": "per-step working set >> 126 MB L2 (inputs larger than L2)",
                       "ref_wrap": a.ref_wrap if a.impl == "reference" else None},
            "clocks": {"sm_mhz": r["clocks"]["sm_mhz"], "sm_max_mhz": r["clocks"]["sm_max_mhz"],
                       "reasons": r["clocks"]["reasons"]},
            "e2e": {"value": round(tps_e2e, 1), "unit": "tokens/s", "h2d_bytes_per_step": r["h2d"],
                    "d2h_bytes_per_step": r["d2h"], "ms_per_step": round(r["ms_e2e"], 3)},
            "gpu_launches": r["launches"], "loss_last": r["loss_last"],
        }
        print(json.dumps(line), flush=True)
    import torch.distributed as dist

    if dist.is_initialized():
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
