#!/usr/bin/env python
"""Basic tensor parallelism: ToyMLP 16 -> 64 -> 16 with in_proj column-wise and out_proj row-wise, AdamW, 5 iterations,
the same seed (hence the same input) on every rank (/root/reference/scripts/03_tensor_parallel_tp/02_basic_tensor_parallel.py)."""
import sys
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from _common import setup, torch  # noqa: E402


def main():
    rank, world, local, device = setup()
    from b200dist.models import ToyMLP
    from b200dist.ops import FusedAdamW
    from b200dist.parallel import init_device_mesh
    from b200dist.parallel.tp import ColwiseParallel, RowwiseParallel, parallelize_module
    from b200dist.utils import cleanup_distributed
    from b200dist.utils.distributed import barrier

    mesh = init_device_mesh(device.type, (world,), ("tp",))
    torch.manual_seed(0)
    model = ToyMLP().to(device)
    parallelize_module(model, mesh, {"in_proj": ColwiseParallel(), "out_proj": RowwiseParallel()})
    print(f"[rank {rank}] in_proj.weight {tuple(model.in_proj.weight.shape)} out_proj.weight {tuple(model.out_proj.weight.shape)}", flush=True)
    opt = FusedAdamW(model.parameters(), lr=0.25)
    for i in range(5):
        torch.manual_seed(i)  # identical input on all TP ranks
        x = torch.rand(20, 16, device=device)
        opt.zero_grad()
        loss = model(x).sum()
        loss.backward()
        opt.step()
        if rank == 0:
            print(f"iter {i}: loss {float(loss):.4f}", flush=True)
    barrier()
    cleanup_distributed()


if __name__ == "__main__":
    main()
