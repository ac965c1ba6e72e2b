#!/usr/bin/env python
"""Manual pipeline: one Linear(+ReLU) stage per rank (128-256-256-256-64 for 4 ranks), 4 microbatches of 8, activations
handed over with blocking send/recv, forward only (/root/reference/scripts/04_pipeline_parallel_pp/01_manual_model_split.py)."""
import sys
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from _common import setup, torch  # noqa: E402


def main():
    import torch.distributed as dist

    rank, world, local, device = setup()
    from b200dist.models import StageModule
    from b200dist.utils import cleanup_distributed
    from b200dist.utils.distributed import barrier

    dims = [128] + [256] * (world - 1) + [64] if world > 1 else [128, 64]
    torch.manual_seed(rank)
    stage = StageModule(dims[rank], dims[rank + 1], last=(rank == world - 1)).to(device)
    n_mb, mb = 4, 8
    outs = []
    with torch.no_grad():
        for i in range(n_mb):
            if rank == 0:
                x = torch.randn(mb, dims[0], device=device)
            else:
                x = torch.empty(mb, dims[rank], device=device)
                dist.recv(x, src=rank - 1)
            y = stage(x)
            if rank < world - 1:
                dist.send(y, dst=rank + 1)
            else:
                outs.append(y)
            print(f"[rank {rank}] microbatch {i}: {tuple(x.shape)} -> {tuple(y.shape)}", flush=True)
    if rank == world - 1:
        print(f"[rank {rank}] pipeline output {tuple(torch.cat(outs).shape)}", flush=True)
    barrier()
    cleanup_distributed()


if __name__ == "__main__":
    main()
