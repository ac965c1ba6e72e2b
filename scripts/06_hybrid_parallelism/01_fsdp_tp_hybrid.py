#!/usr/bin/env python
"""Hybrid TP + SP x FSDP Llama on a 2-D (dp, tp) mesh — /root/reference/scripts/06_hybrid_parallelism/01_fsdp_tp_hybrid.py
(and its mpi4py twin fsdp_tp/fsdp_tp_example.py).  TP plan first, FSDP over the dp dimension last ("order matters").
The flags the reference's README advertises but the script never parses (--tp_size / --fsdp_size) exist here, as do
--model / --seq-len / --dtype; the loss is cross-entropy (the reference uses output.sum(), :178).  Inputs are seeded with
step + dp_rank so all TP ranks of a DP group see identical tokens (:172-174)."""
import argparse
import sys
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from _common import StepTimer, max_over_ranks, setup, torch  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--tp_size", "--tp", type=int, default=0)
    ap.add_argument("--fsdp_size", "--dp", type=int, default=0)
    ap.add_argument("--model", default="llama2-toy")
    ap.add_argument("--seq-len", "--seq_len", dest="seq_len", type=int, default=256)
    ap.add_argument("--batch-size", type=int, default=2)
    ap.add_argument("--iters", type=int, default=10)
    ap.add_argument("--lr", type=float, default=3e-3)
    ap.add_argument("--no-sp", action="store_true")
    ap.add_argument("--no-fused", action="store_true",
                    help="keep collectives and GEMMs separate (default on CUDA with SP: fused all-gather->GEMM / GEMM->reduce-scatter blocks)")
    ap.add_argument("--activation-checkpointing", action="store_true")
    ap.add_argument("--dtype", default="bfloat16", choices=["bfloat16", "float32"])
    a = ap.parse_args()
    rank, world, local, device = setup()
    from b200dist import ops
    from b200dist.models import Transformer
    from b200dist.parallel import FSDP, init_device_mesh
    from b200dist.parallel.tp import llama_tp_plan
    from b200dist.utils import cleanup_distributed, get_logger, rank_log

    log = get_logger("hybrid")
    tp = a.tp_size or min(4, world)
    while world % tp:
        tp -= 1
    dp = a.fsdp_size or world // tp
    assert dp * tp == world, f"dp {dp} x tp {tp} != world {world}"
    mesh = init_device_mesh(device.type, (dp, tp), ("dp", "tp"))
    dp_rank = mesh.get_local_rank("dp")
    rank_log(rank, log, f"mesh dp={dp} tp={tp}; model {a.model}; seq {a.seq_len}; sequence parallel {not a.no_sp}")
    dtype = torch.bfloat16 if (a.dtype == "bfloat16" and device.type == "cuda") else torch.float32
    torch.manual_seed(0)
    model = Transformer.from_name(a.model, device=device, dtype=dtype, max_seq_len=max(a.seq_len, 256))
    llama_tp_plan(model, mesh["tp"], sequence_parallel=not a.no_sp)
    if device.type == "cuda" and not a.no_sp and not a.no_fused:
        # one kernel per (collective, GEMM) pair of the plan; inputs that do not qualify (batch > 1, local sequence not a
        # multiple of 256) take the unfused path block by block.  13B, TP 4 x FSDP 2, seq 8192: 220.9 vs 251.7 ms per step
        from b200dist.parallel.tp import enable_fused_tp
        enable_fused_tp(model, mesh["tp"])
    if a.activation_checkpointing:
        model.enable_activation_checkpointing()
    model = FSDP(model, process_group=mesh.get_group("dp"))
    opt = model.build_optimizer("adamw", lr=a.lr)
    timer = StepTimer(device)
    for i in range(a.iters):
        torch.manual_seed(i + dp_rank)
        tok = torch.randint(model.module.vocab_size, (a.batch_size, a.seq_len + 1)).to(device)
        with timer:
            opt.zero_grad()
            loss = ops.cross_entropy(model(tok[:, :-1]), tok[:, 1:])
            loss.backward()
            opt.step()
        rank_log(rank, log, f"iter {i}: loss {float(loss):.4f}  {timer.times[-1] * 1e3:.1f} ms")
    avg = max_over_ranks(timer.mean(skip=2), device)
    rank_log(rank, log, f"{a.batch_size * a.seq_len * dp / avg:.0f} tokens/s (dp={dp} x tp={tp})")
    if device.type == "cuda":
        rank_log(rank, log, f"peak memory {torch.cuda.max_memory_allocated() / 2**30:.2f} GiB")
    cleanup_distributed()


if __name__ == "__main__":
    main()
