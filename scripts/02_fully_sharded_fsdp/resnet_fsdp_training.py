#!/usr/bin/env python
"""FSDP ResNet-18 (CIFAR stem) — /root/reference/scripts/02_fully_sharded_fsdp/resnet_fsdp_training.py: full shard,
units of >= 1e5 parameters, optional bf16 (--use-amp), SGD, train/test epochs.  CIFAR-10 cannot be downloaded here, so
the data is CIFAR-shaped synthetic images (rank 0 "prepares" the dataset, then a barrier, as the reference does)."""
import argparse
import sys
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from _common import StepTimer, setup, torch  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--epochs", type=int, default=2)
    ap.add_argument("--batch-size", "--batch_size", dest="batch_size", type=int, default=128)
    ap.add_argument("--lr", type=float, default=0.05)
    ap.add_argument("--use-amp", "--use_amp", dest="use_amp", action="store_true")
    ap.add_argument("--backend", default="nccl")
    ap.add_argument("--train-size", type=int, default=4096)
    a = ap.parse_args()
    rank, world, local, device = setup(a.backend)
    from b200dist import ops
    from b200dist.data import SyntheticImageDataset
    from b200dist.models import build_resnet
    from b200dist.parallel import FSDP
    from b200dist.utils import DistributedSampler, cleanup_distributed
    from b200dist.utils.distributed import barrier

    if rank == 0:
        print("preparing CIFAR-shaped synthetic data (32x32x3, 10 classes)", flush=True)
    barrier()
    train = SyntheticImageDataset(a.train_size, image_size=32, num_classes=10, seed=1)
    test = SyntheticImageDataset(a.train_size // 4, image_size=32, num_classes=10, seed=2)
    ts, vs = DistributedSampler(train, world, rank), DistributedSampler(test, world, rank, shuffle=False)
    tl = torch.utils.data.DataLoader(train, batch_size=a.batch_size, sampler=ts, num_workers=2)
    vl = torch.utils.data.DataLoader(test, batch_size=a.batch_size, sampler=vs, num_workers=2)
    dtype = torch.bfloat16 if (a.use_amp and device.type == "cuda") else torch.float32
    torch.manual_seed(0)
    model = build_resnet("resnet18", num_classes=10, cifar_stem=True).to(device).to(dtype)
    units = [(n, m) for n, m in model.named_modules() if n.count(".") == 1 and n.startswith("layer")
             and sum(p.numel() for p in m.parameters()) >= 1e5]
    model = FSDP(model, units=units, fused_optimizer=False)
    opt = model.build_optimizer("sgd", lr=a.lr, momentum=0.9, weight_decay=5e-4)
    for epoch in range(a.epochs):
        ts.set_epoch(epoch)
        model.train()
        timer, seen, loss_sum = StepTimer(device), 0, 0.0
        for x, y in tl:
            x, y = x.to(device).to(dtype), y.to(device)
            with timer:
                opt.zero_grad()
                loss = ops.cross_entropy(model(x), y)
                loss.backward()
                opt.step()
            seen += x.shape[0]
            loss_sum += float(loss) * x.shape[0]
        model.eval()
        correct = total = 0
        with torch.no_grad():
            for x, y in vl:
                x, y = x.to(device).to(dtype), y.to(device)
                correct += int((model(x).argmax(-1) == y).sum())
                total += y.numel()
        stats = torch.tensor([loss_sum, seen, correct, total], dtype=torch.float64, device=device)
        if world > 1:
            torch.distributed.all_reduce(stats)
        if rank == 0:
            print(f"epoch {epoch}: train loss {stats[0] / stats[1]:.4f}  test acc {100 * stats[2] / stats[3]:.2f}%  "
                  f"{seen * world / sum(timer.times):.0f} img/s", flush=True)
    cleanup_distributed()


if __name__ == "__main__":
    main()
