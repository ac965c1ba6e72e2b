#!/usr/bin/env python
"""ResNet data-parallel benchmark — the reference's flagship entrypoint, same flags
(/root/reference/scripts/main.py:128-209): --num_epochs --batch_size --image_size --learning_rate --random_seed
--model_dir --model_filename --resume --backend {nccl,gloo,mpi} --arch {resnet18,50,101,152} --use_syn --steps_syn
--use_fsdp --logfile (+ --local_rank for old launchers, + --dtype).

What differs: one launcher-detection helper instead of an inline copy, CPU-safe (gloo) — the reference calls
torch.cuda.set_device unconditionally (:246) — our DDP/FSDP and sm_100a kernels instead of torch DDP/FSDP + cuDNN,
device-side timing, and throughput reported as images/s (the reference logs seconds/epoch and divides by steps-1, :347-349).

    torchrun --nproc-per-node 8 --master-addr 127.0.0.1 scripts/main.py --use_syn --arch resnet50 --backend nccl
"""
from __future__ import annotations

import argparse
import os
import time

from _common import StepTimer, max_over_ranks, setup, torch


def parse():
    p = argparse.ArgumentParser(description="ResNet DDP/FSDP synthetic benchmark")
    p.add_argument("--local_rank", type=int, default=0)
    p.add_argument("--num_epochs", type=int, default=100)
    p.add_argument("--batch_size", type=int, default=256)
    p.add_argument("--image_size", type=int, default=224)
    p.add_argument("--learning_rate", type=float, default=0.1)
    p.add_argument("--random_seed", type=int, default=0)
    p.add_argument("--model_dir", type=str, default="saved_models")
    p.add_argument("--model_filename", type=str, default="resnet_distributed.pth")
    p.add_argument("--resume", action="store_true")
    p.add_argument("--backend", type=str, default="nccl", choices=["nccl", "gloo", "mpi"])
    p.add_argument("--arch", type=str, default="resnet18", choices=["resnet18", "resnet50", "resnet101", "resnet152"])
    p.add_argument("--use_syn", action="store_true")
    p.add_argument("--steps_syn", type=int, default=20)
    p.add_argument("--use_fsdp", action="store_true")
    p.add_argument("--logfile", type=str, default="resnet_benchmark.log")
    p.add_argument("--dtype", type=str, default="bfloat16", choices=["bfloat16", "float32"])
    p.add_argument("--num_classes", type=int, default=1000)
    return p.parse_args()


def main():
    a = parse()
    rank, world, local, device = setup(a.backend)
    torch.manual_seed(a.random_seed)
    from b200dist import ops
    from b200dist.data import SyntheticImageDataset
    from b200dist.models import build_resnet
    from b200dist.ops import FusedSGD
    from b200dist.parallel import DDP, FSDP
    from b200dist.utils import DistributedSampler, load_checkpoint, save_checkpoint

    dtype = torch.bfloat16 if (a.dtype == "bfloat16" and device.type == "cuda") else torch.float32
    model = build_resnet(a.arch, num_classes=a.num_classes).to(device).to(dtype)
    path = os.path.join(a.model_dir, a.model_filename)
    if a.resume:
        load_checkpoint(model, None, path, device)
    if a.use_fsdp:
        # one flat unit like the reference's policy-free FSDP(model) (:256), sharded over all ranks
        model = FSDP(model, units=[], fused_optimizer=False)
        opt = model.build_optimizer("sgd", lr=a.learning_rate, momentum=0.9, weight_decay=1e-5)
    else:
        model = DDP(model) if world > 1 else model
        opt = FusedSGD(model.parameters(), lr=a.learning_rate, momentum=0.9, weight_decay=1e-5)
    if a.use_syn:
        x = torch.rand(a.batch_size, 3, a.image_size, a.image_size, device=device).to(dtype)
        y = torch.zeros(a.batch_size, dtype=torch.long, device=device)
        loader = None
    else:
        ds = SyntheticImageDataset(size=a.batch_size * a.steps_syn * world, image_size=a.image_size, num_classes=a.num_classes)
        sampler = DistributedSampler(ds, num_replicas=world, rank=rank)
        loader = torch.utils.data.DataLoader(ds, batch_size=a.batch_size, sampler=sampler, num_workers=2, pin_memory=True)
    epoch_times = []
    for epoch in range(a.num_epochs):
        model.train()
        timer = StepTimer(device)
        t0 = time.perf_counter()
        it = range(a.steps_syn) if loader is None else loader
        if loader is not None:
            sampler.set_epoch(epoch)
        n = 0
        for batch in it:
            if loader is not None:
                x, y = batch[0].to(device, non_blocking=True).to(dtype), batch[1].to(device, non_blocking=True)
            with timer:
                opt.zero_grad()
                loss = ops.cross_entropy(model(x), y)
                loss.backward()
                fin = getattr(model, "finish_backward", None)
                if fin:
                    fin()
                opt.step()
            n += 1
        if device.type == "cuda":
            torch.cuda.synchronize()
        elapsed = max_over_ranks(time.perf_counter() - t0, device)
        if epoch > 0 or a.num_epochs == 1:  # epoch 0 is warm-up, as in the reference (:344-345)
            epoch_times.append(elapsed)
        if rank == 0:
            ips = a.batch_size * n * world / elapsed
            print(f"epoch {epoch}: {elapsed:.3f} s, {n} steps, avg step {elapsed / n * 1e3:.2f} ms, {ips:.1f} images/s "
                  f"(loss {float(loss):.4f})", flush=True)
    if rank == 0 and epoch_times:
        avg = sum(epoch_times) / len(epoch_times)
        ips = a.batch_size * a.steps_syn * world / avg
        line = (f"arch={a.arch} world={world} backend={a.backend} fsdp={a.use_fsdp} dtype={a.dtype} batch={a.batch_size} "
                f"steps={a.steps_syn} Average epoch time: {avg:.4f} s images/s: {ips:.1f}")
        print(line, flush=True)
        with open(a.logfile, "a") as f:
            f.write(line + "\n")
    if a.model_dir and a.num_epochs > 1:
        save_checkpoint(model, None, a.num_epochs, path)
    from b200dist.utils import cleanup_distributed

    cleanup_distributed()


if __name__ == "__main__":
    main()
