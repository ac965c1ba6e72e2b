#!/usr/bin/env python
"""DDP U-Net on ERA5-like grids with latitude-weighted MSE and AdamW
(/root/reference/scripts/01_data_parallel_ddp/multinode_ddp_unet.py).  Reports samples/s per batch window, per epoch and in
total (global and per GPU), device-timed, max over ranks.  Runs on one GPU / CPU without a launcher."""
import argparse
import sys
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from _common import StepTimer, max_over_ranks, setup, torch  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--epochs", type=int, default=5)
    ap.add_argument("--batch_size", type=int, default=4)
    ap.add_argument("--lr", type=float, default=1e-3)
    ap.add_argument("--num_samples", type=int, default=256)
    ap.add_argument("--lat", type=int, default=181)
    ap.add_argument("--lon", type=int, default=360)
    ap.add_argument("--base_dim", type=int, default=64)
    ap.add_argument("--backend", default="nccl")
    ap.add_argument("--save_every", type=int, default=1000)
    ap.add_argument("--strategy", default="ddp", choices=["ddp", "fsdp"])
    a = ap.parse_args()
    rank, world, local, device = setup(a.backend)
    from b200dist.data import ERA5Dataset
    from b200dist.models import SimpleUNet, latitude_weighted_mse
    from b200dist.ops import FusedAdamW
    from b200dist.parallel import DDP, FSDP
    from b200dist.utils import DistributedSampler, cleanup_distributed, save_checkpoint

    ds = ERA5Dataset(a.num_samples, lat=a.lat, lon=a.lon)
    sampler = DistributedSampler(ds, num_replicas=world, rank=rank)
    loader = torch.utils.data.DataLoader(ds, batch_size=a.batch_size, sampler=sampler, num_workers=2, pin_memory=device.type == "cuda")
    torch.manual_seed(0)
    dtype = torch.bfloat16 if device.type == "cuda" else torch.float32
    model = SimpleUNet(ds.channels, ds.channels, a.base_dim).to(device).to(dtype)
    if a.strategy == "fsdp":
        # one unit per encoder/decoder block, like size-based auto-wrap >= 1e5 parameters (reference fsdp unet :202-213)
        units = [(n, m) for n, m in model.named_children() if sum(p.numel() for p in m.parameters()) >= 1e5]
        model = FSDP(model, units=units, fused_optimizer=True)
        opt = model.build_optimizer("adamw", lr=a.lr)
    else:
        model = DDP(model) if world > 1 else model
        opt = FusedAdamW(model.parameters(), lr=a.lr)
    total_samples, total_time = 0, 0.0
    for epoch in range(a.epochs):
        sampler.set_epoch(epoch)
        timer = StepTimer(device)
        for i, (x, y) in enumerate(loader):
            x, y = x.to(device, non_blocking=True).to(dtype), y.to(device, non_blocking=True).to(dtype)
            with timer:
                opt.zero_grad()
                loss = latitude_weighted_mse(model(x), y)
                loss.backward()
                if hasattr(model, "finish_backward"):
                    model.finish_backward()
                opt.step()
            if rank == 0 and i % 10 == 0:
                print(f"  epoch {epoch} batch {i}: loss {float(loss):.4f}  {a.batch_size * world / timer.times[-1]:.1f} samples/s", flush=True)
        secs = max_over_ranks(sum(timer.times), device)
        n = len(loader) * a.batch_size * world
        total_samples += n
        total_time += secs
        if rank == 0:
            print(f"epoch {epoch}: {secs:.2f} s  {n / secs:.1f} samples/s  ({n / secs / world:.1f} per GPU)", flush=True)
        if a.save_every and (epoch + 1) % a.save_every == 0:
            save_checkpoint(model, None, epoch + 1, f"checkpoint_epoch_{epoch + 1}.pt")  # all ranks enter (barrier inside)
    if rank == 0 and total_time:
        print(f"total: {total_samples} samples in {total_time:.2f} s = {total_samples / total_time:.1f} samples/s "
              f"({total_samples / total_time / world:.1f} per GPU, {world} GPU(s))", flush=True)
    cleanup_distributed()


if __name__ == "__main__":
    main()
