#!/usr/bin/env python
"""DistributedSampler demo: each rank sees a disjoint shard, ``set_epoch`` reshuffles, rank 0 saves and everyone
meets at a barrier (/root/reference/scripts/01_data_parallel_ddp/distributed_dataloader.py)."""
import argparse
import sys
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from _common import setup, torch  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--epochs", type=int, default=3)
    ap.add_argument("--batch_size", type=int, default=32)
    ap.add_argument("--backend", default="nccl")
    a = ap.parse_args()
    rank, world, local, device = setup(a.backend)
    from b200dist import ops
    from b200dist.data import SimpleDataset
    from b200dist.models import SimpleModel
    from b200dist.ops import FusedSGD
    from b200dist.parallel import DDP
    from b200dist.utils import DistributedSampler, cleanup_distributed, save_checkpoint

    ds = SimpleDataset(1000)
    sampler = DistributedSampler(ds, num_replicas=world, rank=rank, shuffle=True)
    loader = torch.utils.data.DataLoader(ds, batch_size=a.batch_size, sampler=sampler)
    print(f"[rank {rank}] {len(sampler)} of {len(ds)} samples, {len(loader)} batches", flush=True)
    torch.manual_seed(0)
    model = SimpleModel().to(device)
    model = DDP(model) if world > 1 else model
    opt = FusedSGD(model.parameters(), lr=0.05)
    for epoch in range(a.epochs):
        sampler.set_epoch(epoch)  # without this every epoch would replay the same order
        first = next(iter(sampler))
        tot = 0.0
        for x, y in loader:
            x, y = x.to(device), y.to(device)
            opt.zero_grad()
            loss = ops.cross_entropy(model(x), y)
            loss.backward()
            if hasattr(model, "finish_backward"):
                model.finish_backward()
            opt.step()
            tot += float(loss)
        print(f"[rank {rank}] epoch {epoch}: first index {first}, mean loss {tot / len(loader):.4f}", flush=True)
    save_checkpoint(model, None, a.epochs, "dataloader_demo.pt")
    cleanup_distributed()


if __name__ == "__main__":
    main()
