#!/usr/bin/env python
"""Secondary headline: ResNet-50 DDP bf16 images/s, synthetic ImageNet-shaped batches (BASELINE config 2).

    python bench/resnet_bench.py --gpus 1                       # this framework (own conv/BN kernels + peer-memory DDP)
    torchrun --nproc-per-node N bench/resnet_bench.py --gpus N  [--impl reference]

``--impl reference`` = the reference recipe's code path (/root/reference/scripts/main.py:249-311): torchvision model,
torch DistributedDataParallel (25 MB buckets), SGD(momentum 0.9, wd 1e-5), cross-entropy, one fixed synthetic batch —
with bf16 autocast + channels_last as BASELINE.md row 2 prescribes.  Prints one JSON line (same fields as bench.py)."""
from __future__ import annotations

import argparse
import json
import os
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--batch-per-gpu", type=int, default=256)
    ap.add_argument("--arch", default="resnet50")
    ap.add_argument("--impl", default="b200dist", choices=["b200dist", "reference"])
    a = ap.parse_args()
    rank, world, local = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1"); os.environ.setdefault("MASTER_PORT", "29500")
        dist.init_process_group("nccl", device_id=dev)
    torch.manual_seed(0)
    B = a.batch_per_gpu
    x = torch.rand(B, 3, 224, 224, device=dev)
    y = torch.zeros(B, dtype=torch.long, device=dev)
    if a.impl == "reference":
        import torchvision

        model = getattr(torchvision.models, a.arch)().to(dev).to(memory_format=torch.channels_last)
        if world > 1:
            model = torch.nn.parallel.DistributedDataParallel(model, device_ids=[local])
        opt = torch.optim.SGD(model.parameters(), lr=0.1, momentum=0.9, weight_decay=1e-5)
        xin = x.contiguous(memory_format=torch.channels_last)
        crit = torch.nn.CrossEntropyLoss()

        def step():
            opt.zero_grad()
            with torch.autocast("cuda", dtype=torch.bfloat16):
                out = model(xin)
            loss = crit(out.float(), y)
            loss.backward()
            opt.step()
            return loss
    else:
        from b200dist import ops
        from b200dist.models import build_resnet
        from b200dist.ops import FusedSGD
        from b200dist.parallel import DDP

        model = build_resnet(a.arch).to(dev).to(torch.bfloat16)
        if world > 1:
            model = DDP(model)
        opt = FusedSGD(model.parameters(), lr=0.1, momentum=0.9, weight_decay=1e-5)
        xin = x.bfloat16().contiguous(memory_format=torch.channels_last)

        def step():
            opt.zero_grad()
            loss = ops.cross_entropy(model(xin), y)
            loss.backward()
            if world > 1:
                model.finish_backward()
            opt.step()
            return loss
    for _ in range(a.warmup):
        step()
    if world > 1:
        dist.barrier(device_ids=[local])
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(a.steps):
        loss = step()
    e1.record()
    if world > 1:
        dist.barrier(device_ids=[local])
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1)
    if world > 1:
        t = torch.tensor([ms], device=dev, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t)
    if rank == 0:
        print(json.dumps({"metric": f"{a.arch}_ddp_train_images_per_s", "value": round(B * world * a.steps / (ms / 1e3), 1), "unit": "images/s",
                          "n_gpus": world, "steps": a.steps, "warmup": a.warmup, "ms_per_step": round(ms / a.steps, 3), "higher_is_better": True,
                          "scaling": "weak", "dtype": "bf16", "data": "synthetic 224x224 images, random-init weights", "impl": a.impl,
                          "config": {"model": a.arch, "global_batch": B * world, "parallelism": f"ddp{world}"}, "loss_last": float(loss)}), flush=True)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
