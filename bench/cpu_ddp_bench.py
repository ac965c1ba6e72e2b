#!/usr/bin/env python
"""BASELINE config 1: ResNet-18 DDP on CPU over gloo, world 2, synthetic 224 x 224 batches -> images/s.

    torchrun --nproc-per-node 2 --master-addr 127.0.0.1 bench/cpu_ddp_bench.py [--impl reference]

b200dist arm : models.build_resnet + parallel.DDP (bucketed reducer; on CPU tensors the ops run their PyTorch expressions
               and the collectives go through gloo - the device-agnostic path of the framework) + FusedSGD.
reference arm: what /root/reference/scripts/main.py:249-311 does (torchvision resnet18, torch DistributedDataParallel, SGD
               momentum 0.9 wd 1e-5, cross-entropy, one fixed synthetic batch) behind a CPU-safe init - the script itself
               calls torch.cuda.set_device unconditionally (:246) and cannot start on a CPU box.
Host-timed (no device), max over ranks, one JSON line from rank 0."""
from __future__ import annotations

import argparse
import json
import os
import sys
import time
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--impl", default="b200dist", choices=["b200dist", "reference"])
    ap.add_argument("--arch", default="resnet18")
    ap.add_argument("--batch-per-rank", type=int, default=8)
    ap.add_argument("--image-size", type=int, default=224)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=2)
    ap.add_argument("--threads", type=int, default=0, help="intra-op threads per rank (0: cores / world)")
    a = ap.parse_args()
    rank, world = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1"); os.environ.setdefault("MASTER_PORT", "29533")
    torch.set_num_threads(a.threads or max(1, (os.cpu_count() or 2) // world))
    if world > 1:
        dist.init_process_group("gloo")
    torch.manual_seed(0)
    B = a.batch_per_rank
    x = torch.rand(B, 3, a.image_size, a.image_size)
    y = torch.zeros(B, dtype=torch.long)
    if a.impl == "reference":
        import torchvision

        model = getattr(torchvision.models, a.arch)()
        if world > 1:
            model = torch.nn.parallel.DistributedDataParallel(model)
        opt = torch.optim.SGD(model.parameters(), lr=0.1, momentum=0.9, weight_decay=1e-5)
        crit = torch.nn.CrossEntropyLoss()

        def step():
            opt.zero_grad(); loss = crit(model(x), y); loss.backward(); opt.step(); return loss
    else:
        from b200dist import ops
        from b200dist.models import build_resnet
        from b200dist.ops import FusedSGD
        from b200dist.parallel import DDP

        model = build_resnet(a.arch)
        if world > 1:
            model = DDP(model)
        opt = FusedSGD(model.parameters(), lr=0.1, momentum=0.9, weight_decay=1e-5)

        def step():
            opt.zero_grad(); loss = ops.cross_entropy(model(x), y); loss.backward()
            if world > 1:
                model.finish_backward()
            opt.step(); return loss
    for _ in range(a.warmup):
        step()
    if world > 1:
        dist.barrier()
    t0 = time.perf_counter()
    for _ in range(a.steps):
        loss = step()
    dt = time.perf_counter() - t0
    if world > 1:
        t = torch.tensor([dt], dtype=torch.float64); dist.all_reduce(t, op=dist.ReduceOp.MAX); dt = float(t)
    if rank == 0:
        print(json.dumps({"metric": f"{a.arch}_ddp_cpu_gloo_images_per_s", "value": round(B * world * a.steps / dt, 2), "unit": "images/s",
                          "n_procs": world, "steps": a.steps, "warmup": a.warmup, "ms_per_step": round(dt / a.steps * 1e3, 1),
                          "dtype": "fp32", "device": "cpu", "impl": a.impl, "threads_per_rank": torch.get_num_threads(),
                          "config": {"model": a.arch, "global_batch": B * world, "image_size": a.image_size, "backend": "gloo"},
                          "loss_last": round(float(loss), 4)}), flush=True)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
