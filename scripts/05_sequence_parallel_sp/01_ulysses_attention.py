#!/usr/bin/env python
"""Ulysses sequence parallelism demo (described in the reference's docs/guide/08_sequence_parallel.md:41-80, whose
scripts/05_sequence_parallel_sp/ directory does not exist).  Checks the sharded result against single-device attention and
reports tokens/s."""
import argparse
import sys
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from _common import StepTimer, max_over_ranks, setup, torch  # noqa: E402


def main(kind="ulysses"):
    ap = argparse.ArgumentParser()
    ap.add_argument("--seq-len", type=int, default=8192)
    ap.add_argument("--heads", type=int, default=32)
    ap.add_argument("--head-dim", type=int, default=128)
    ap.add_argument("--batch", type=int, default=1)
    ap.add_argument("--iters", type=int, default=10)
    ap.add_argument("--check", action="store_true")
    a = ap.parse_args()
    rank, world, local, device = setup()
    from b200dist import ops
    from b200dist.comm import get_communicator
    from b200dist.parallel import ring_attention, ulysses_attention, zigzag_ring_attention, zigzag_shard
    from b200dist.utils import cleanup_distributed

    zig = kind == "zigzag"
    fn = {"ulysses": ulysses_attention, "ring": ring_attention,
          "zigzag": lambda q, k, v, comm, causal=True: zigzag_ring_attention(q, k, v, comm)}[kind]
    comm = get_communicator(None)
    dtype = torch.bfloat16 if device.type == "cuda" else torch.float32
    s = a.seq_len // world
    g = torch.Generator().manual_seed(0)
    full = [torch.randn(a.batch, a.seq_len, a.heads, a.head_dim, generator=g).to(device).to(dtype) for _ in range(3)]
    shard = (lambda t: zigzag_shard(t, world, rank)) if zig else (lambda t: t[:, rank * s:(rank + 1) * s].contiguous())
    q, k, v = [shard(t).requires_grad_(True) for t in full]
    out = fn(q, k, v, comm, causal=True)
    if a.check:
        ref = shard(ops.attention(*full, causal=True))
        err = float((out.float() - ref.float()).abs().max())
        print(f"[rank {rank}] max |{kind} - single-device| = {err:.4g}", flush=True)
    timer = StepTimer(device)
    for _ in range(a.iters):
        with timer:
            o = fn(q, k, v, comm, causal=True)
            o.sum().backward()
    t = max_over_ranks(timer.mean(skip=2), device)
    if rank == 0:
        print(f"{kind} attention fwd+bwd: seq {a.seq_len} over {world} rank(s): {t * 1e3:.2f} ms  "
              f"{a.batch * a.seq_len / t:.0f} tokens/s", flush=True)
    cleanup_distributed()


if __name__ == "__main__":
    main("ulysses")
