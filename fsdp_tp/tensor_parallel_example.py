#!/usr/bin/env python
"""Tensor-parallel toy example (reference fsdp_tp/tensor_parallel_example.py): ToyModel 64 -> 64 -> 64 with the first
linear column-parallel and the second row-parallel, on a 1-D mesh or - with --two-d - on the "tp" dimension of a 2-D
(dp, tp) mesh (:87-122).  Identical inputs on every TP rank; AdamW; a few iterations; prints the loss.
    scripts/launch.sh 4 fsdp_tp/tensor_parallel_example.py [--two-d]"""
import argparse
import sys
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import torch  # noqa: E402

from b200dist.models import ToyModel  # noqa: E402
from b200dist.parallel import init_device_mesh  # noqa: E402
from b200dist.parallel.tp import ColwiseParallel, RowwiseParallel, parallelize_module  # noqa: E402
from b200dist.utils import cleanup_distributed, get_logger, init_distributed, rank_log  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--two-d", action="store_true", help="2-D (dp, tp) mesh, TP over the inner dimension")
    ap.add_argument("--iters", type=int, default=10)
    a = ap.parse_args()
    rank, world, local = init_distributed(verbose=False)
    log = get_logger("tp_example")
    dev = torch.device("cuda", torch.cuda.current_device()) if torch.cuda.is_available() else torch.device("cpu")
    if a.two_d and world >= 4 and world % 2 == 0:
        mesh = init_device_mesh(dev.type, (2, world // 2), ("dp", "tp"))
        tp_mesh = mesh["tp"]
        seed_rank = mesh.get_local_rank("dp")
    else:
        tp_mesh = init_device_mesh(dev.type, (world,), ("tp",))
        seed_rank = 0
    rank_log(rank, log, f"mesh: {'2-D (dp, tp)' if a.two_d else '1-D'}; tp size {tp_mesh.size()}")
    torch.manual_seed(0)
    model = ToyModel().to(dev)
    model = parallelize_module(model, tp_mesh, {"in_proj": ColwiseParallel(), "out_proj": RowwiseParallel()})
    opt = torch.optim.AdamW(model.parameters(), lr=0.25)
    for i in range(a.iters):
        torch.manual_seed(i + seed_rank)            # TP peers must see the same input
        x = torch.rand(20, 64, device=dev)
        loss = model(x).sum()
        opt.zero_grad(); loss.backward(); opt.step()
        rank_log(rank, log, f"iter {i}: loss {float(loss):.4f}")
    rank_log(rank, log, "tensor-parallel example done")
    cleanup_distributed()


if __name__ == "__main__":
    main()
