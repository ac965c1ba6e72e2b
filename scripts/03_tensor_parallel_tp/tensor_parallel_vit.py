#!/usr/bin/env python
"""Tensor-parallel ViT on ERA5-like data over a 2-D (dp, tp) mesh
(/root/reference/scripts/03_tensor_parallel_tp/tensor_parallel_vit.py).  Two defects of the reference are fixed here
(SURVEY.md §9): data is sharded by DP rank (TP peers see the same batch) and gradients ARE averaged across the dp
dimension (DDP over the dp group)."""
import argparse
import sys
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from _common import StepTimer, max_over_ranks, setup, torch  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--epochs", type=int, default=2)
    ap.add_argument("--batch_size", type=int, default=8)
    ap.add_argument("--tp_size", type=int, default=0)
    ap.add_argument("--lat", type=int, default=64)
    ap.add_argument("--lon", type=int, default=128)
    ap.add_argument("--patch", type=int, default=8)
    ap.add_argument("--embed_dim", type=int, default=256)
    ap.add_argument("--depth", type=int, default=6)
    ap.add_argument("--heads", type=int, default=8)
    ap.add_argument("--num_samples", type=int, default=256)
    ap.add_argument("--lr", type=float, default=1e-3)
    a = ap.parse_args()
    rank, world, local, device = setup()
    from b200dist.data import ERA5Dataset
    from b200dist.models import SimpleViT, latitude_weighted_mse
    from b200dist.models.vit import vit_tp_plan
    from b200dist.ops import FusedAdamW
    from b200dist.parallel import DDP, init_device_mesh
    from b200dist.utils import DistributedSampler, cleanup_distributed

    tp = a.tp_size or min(4, world)
    while world % tp or a.heads % tp:
        tp -= 1
    mesh = init_device_mesh(device.type, (world // tp, tp), ("dp", "tp"))
    ds = ERA5Dataset(a.num_samples, lat=a.lat, lon=a.lon)
    sampler = DistributedSampler(ds, mesh=mesh, dp_dim="dp")
    loader = torch.utils.data.DataLoader(ds, batch_size=a.batch_size, sampler=sampler, num_workers=2)
    torch.manual_seed(0)
    dtype = torch.bfloat16 if device.type == "cuda" else torch.float32
    model = SimpleViT(ds.channels, ds.channels, a.patch, a.lat, a.lon, a.embed_dim, a.depth, a.heads).to(device).to(dtype)
    vit_tp_plan(model, mesh["tp"])
    if mesh.size("dp") > 1:
        model = DDP(model, process_group=mesh.get_group("dp"))
    opt = FusedAdamW(model.parameters(), lr=a.lr)
    for epoch in range(a.epochs):
        sampler.set_epoch(epoch)
        timer = StepTimer(device)
        for x, y in loader:
            x, y = x.to(device).to(dtype), y.to(device).to(dtype)
            with timer:
                opt.zero_grad()
                loss = latitude_weighted_mse(model(x), y)
                loss.backward()
                if hasattr(model, "finish_backward"):
                    model.finish_backward()
                opt.step()
        secs = max_over_ranks(sum(timer.times), device)
        if rank == 0:
            n = len(loader) * a.batch_size * mesh.size("dp")
            print(f"epoch {epoch}: loss {float(loss):.4f}  {n / secs:.1f} samples/s (dp={mesh.size('dp')} tp={tp})", flush=True)
    cleanup_distributed()


if __name__ == "__main__":
    main()
