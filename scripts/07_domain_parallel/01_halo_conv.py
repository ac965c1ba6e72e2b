#!/usr/bin/env python
"""Domain parallelism with halo exchange (reference: docs/guide/10_domain_parallel.md:45-149; its
scripts/07_domain_parallel_shardtensor/ directory does not exist).  A lat x lon grid is split along latitude; a naive
per-shard 3x3 convolution is wrong at the cuts, the halo-exchange convolution matches the unsharded result."""
import sys
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from _common import setup, torch  # noqa: E402


def main():
    rank, world, local, device = setup()
    from b200dist.comm import get_communicator
    from b200dist.ops import conv as cops
    from b200dist.parallel import halo_conv2d
    from b200dist.utils import cleanup_distributed

    comm = get_communicator(None)
    dtype = torch.bfloat16 if device.type == "cuda" else torch.float32
    g = torch.Generator().manual_seed(0)
    H = 24 * world
    x = torch.randn(2, 64, H, 96, generator=g).to(device).to(dtype)
    w = (torch.randn(64, 64, 3, 3, generator=g) * 0.05).to(device).to(dtype)
    full = cops.conv2d(x, w, None, 1, 1)
    rows = slice(rank * 24, (rank + 1) * 24)
    naive = cops.conv2d(x[:, :, rows].contiguous(), w, None, 1, 1)
    halo = halo_conv2d(x[:, :, rows].contiguous(), w, None, comm, padding=1)
    e_naive = float((naive.float() - full[:, :, rows].float()).abs().max())
    e_halo = float((halo.float() - full[:, :, rows].float()).abs().max())
    print(f"[rank {rank}] rows {rows.start}-{rows.stop}: naive split error {e_naive:.4g} | with halo exchange {e_halo:.4g}", flush=True)
    cleanup_distributed()


if __name__ == "__main__":
    main()
