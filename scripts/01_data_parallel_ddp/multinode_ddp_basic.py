#!/usr/bin/env python
"""DDP basics: a Trainer with snapshot resume on a 20-feature regression
(/root/reference/scripts/01_data_parallel_ddp/multinode_ddp_basic.py).  Flags: --total_epochs --backend --save_every
--batch_size.  The snapshot ({MODEL_STATE, EPOCHS_RUN}) is loaded before wrapping and written once per JOB by global
rank 0 (the reference writes it once per node, :195)."""
import argparse
import sys
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from _common import StepTimer, setup, torch  # noqa: E402


class Trainer:
    def __init__(self, model, loader, sampler, optimizer, save_every, snapshot_path, device, rank, world):
        from b200dist.parallel import DDP
        from b200dist.utils import load_snapshot

        self.device, self.rank, self.loader, self.sampler, self.opt = device, rank, loader, sampler, optimizer
        self.save_every, self.snapshot_path = save_every, snapshot_path
        self.epochs_run = load_snapshot(model, snapshot_path, device)
        if self.epochs_run and rank == 0:
            print(f"resuming from {snapshot_path} at epoch {self.epochs_run}")
        self.model = DDP(model) if world > 1 else model

    def _run_epoch(self, epoch):
        self.sampler.set_epoch(epoch)
        timer, last = StepTimer(self.device), None
        for x, y in self.loader:
            x, y = x.to(self.device), y.to(self.device)
            with timer:
                self.opt.zero_grad()
                last = torch.nn.functional.mse_loss(self.model(x), y)
                last.backward()
                if hasattr(self.model, "finish_backward"):
                    self.model.finish_backward()
                self.opt.step()
        return float(last), sum(timer.times)

    def train(self, total_epochs):
        from b200dist.utils import save_snapshot

        for epoch in range(self.epochs_run, total_epochs):
            loss, secs = self._run_epoch(epoch)
            if self.rank == 0:
                print(f"[epoch {epoch}] loss {loss:.5f} | {len(self.loader)} steps | {secs:.3f} s", flush=True)
            if self.save_every and (epoch + 1) % self.save_every == 0:
                save_snapshot(self.model, epoch + 1, self.snapshot_path)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--total_epochs", type=int, default=50)
    ap.add_argument("--backend", default="nccl")
    ap.add_argument("--save_every", type=int, default=100)
    ap.add_argument("--batch_size", type=int, default=32)
    ap.add_argument("--snapshot", default="snapshot.pt")
    a = ap.parse_args()
    rank, world, local, device = setup(a.backend)
    from b200dist.data import ToyRegressionDataset
    from b200dist.models import regression_head
    from b200dist.ops import FusedSGD
    from b200dist.utils import DistributedSampler, cleanup_distributed

    ds = ToyRegressionDataset(2048)
    sampler = DistributedSampler(ds, num_replicas=world, rank=rank)
    loader = torch.utils.data.DataLoader(ds, batch_size=a.batch_size, sampler=sampler, pin_memory=device.type == "cuda")
    torch.manual_seed(0)
    model = regression_head(20).to(device)
    opt = FusedSGD(model.parameters(), lr=1e-3)
    Trainer(model, loader, sampler, opt, a.save_every, a.snapshot, device, rank, world).train(a.total_epochs)
    cleanup_distributed()


if __name__ == "__main__":
    main()
