#!/usr/bin/env python
"""BASELINE config 4: Llama-2 13B, TP 4 (+ sequence parallel) x FSDP 2, seq 8192, bf16, AdamW -> tokens/s.

    torchrun --nproc-per-node 8 --master-addr 127.0.0.1 bench/hybrid_bench.py --tp 4 [--impl reference] [--no-fused]

b200dist arm : llama_tp_plan (the plan of 01_fsdp_tp_hybrid.py:110-152 on local shards, peer-memory collectives; by default
               with the fused all-gather->GEMM / GEMM->reduce-scatter kernels, --no-fused keeps them apart), then FSDP over the dp mesh dimension.
reference arm: the reference's own Llama model with the SAME plan expressed in torch DTensor styles exactly as the recipe
               does (parallelize_module + ColwiseParallel/RowwiseParallel/SequenceParallel/PrepareModuleInput), then
               fully_shard over dp (per block + root, bf16 params / bf16 reduce) and torch AdamW(foreach) — BASELINE.md §4
               row 4.  Both arms: cross-entropy loss, identical tokens on the TP peers of a dp replica."""
from __future__ import annotations

import argparse
import json
import os
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--tp", type=int, default=4)
    ap.add_argument("--model", default="llama2-13b")
    ap.add_argument("--n-layers", type=int, default=0)
    ap.add_argument("--seq-len", type=int, default=8192)
    ap.add_argument("--batch-per-replica", type=int, default=1)
    ap.add_argument("--steps", type=int, default=6)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--fused", action="store_true", default=True, help="fused TP kernels (default)")
    ap.add_argument("--no-fused", dest="fused", action="store_false", help="separate collective kernels + GEMMs")
    ap.add_argument("--no-sp", action="store_true")
    ap.add_argument("--activation-checkpointing", action="store_true", help="recompute each block in backward (b200dist arm)")
    ap.add_argument("--vocab-parallel-loss", action="store_true", help="keep the logits vocabulary-sharded (no [M, V] all-gather)")
    ap.add_argument("--impl", default="b200dist", choices=["b200dist", "reference"])
    ap.add_argument("--verbose", action="store_true", help="print the loss of every step (adds a sync per step)")
    a = ap.parse_args()
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    tp = a.tp; dp = world // tp
    assert dp * tp == world
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    from b200dist.models.llama2 import PRESETS

    cfg = dict(PRESETS[a.model]); cfg["max_seq_len"] = a.seq_len
    if a.n_layers:
        cfg["n_layers"] = a.n_layers
    vocab, S, B = cfg["vocab_size"], a.seq_len, a.batch_per_replica
    torch.manual_seed(0)
    if a.impl == "b200dist":
        from b200dist import ops
        from b200dist.models import ModelArgs, Transformer
        from b200dist.parallel import FSDP, init_device_mesh
        from b200dist.parallel.tp import enable_fused_tp, llama_tp_plan, vocab_parallel_cross_entropy

        mesh = init_device_mesh("cuda", (dp, tp), ("dp", "tp"))
        dp_rank = mesh.get_local_rank("dp")
        prev = torch.get_default_dtype(); torch.set_default_dtype(torch.bfloat16)
        with torch.device(dev):
            model = Transformer(ModelArgs(**cfg))
        torch.set_default_dtype(prev)
        llama_tp_plan(model, mesh["tp"], sequence_parallel=not a.no_sp, gather_logits=not a.vocab_parallel_loss)
        tp_comm = mesh["tp"].communicator()
        if a.fused:
            enable_fused_tp(model, mesh["tp"])
        if a.activation_checkpointing:
            model.enable_activation_checkpointing()
        torch.cuda.empty_cache()
        model = FSDP(model, process_group=mesh.get_group("dp"))
        opt = model.build_optimizer("adamw", lr=3e-4, betas=(0.9, 0.95), weight_decay=0.1)

        def step(x, y):
            opt.zero_grad()
            if a.vocab_parallel_loss:
                loss = vocab_parallel_cross_entropy(model(x), y, tp_comm)
            else:
                loss = ops.cross_entropy(model(x).view(-1, vocab), y.reshape(-1))
            loss.backward()
            opt.step()
            return loss
    else:
        from baseline.install_ref import ensure_reference, why_unavailable
        ref_root = ensure_reference()
        if ref_root is None:
            if rank == 0:
                print(json.dumps({"impl": "reference", "unavailable": why_unavailable()}))
            dist.destroy_process_group(); return
        sys.path.insert(0, str(ref_root / "reference_src" / "fsdp_tp"))
        import torch.nn.functional as F
        from llama2_model import ModelArgs, Transformer
        from torch.distributed.device_mesh import init_device_mesh
        from torch.distributed.fsdp import MixedPrecisionPolicy, fully_shard
        from torch.distributed.tensor import Replicate, Shard
        from torch.distributed.tensor.parallel import (ColwiseParallel, PrepareModuleInput, RowwiseParallel,
                                                       SequenceParallel, parallelize_module)

        mesh = init_device_mesh("cuda", (dp, tp), mesh_dim_names=("dp", "tp"))
        tp_mesh, dp_mesh = mesh["tp"], mesh["dp"]
        dp_rank = dp_mesh.get_local_rank()
        with torch.device(dev):
            model = Transformer.from_model_args(ModelArgs(**cfg))
        # the recipe's plan, 01_fsdp_tp_hybrid.py:110-152
        model = parallelize_module(model, tp_mesh, {
            "tok_embeddings": RowwiseParallel(input_layouts=Replicate(), output_layouts=Shard(1)),
            "norm": SequenceParallel(),
            "output": ColwiseParallel(input_layouts=Shard(1), output_layouts=Replicate()),
        })
        for block in model.layers:
            parallelize_module(block, tp_mesh, {
                "attention_norm": SequenceParallel(),
                "attention": PrepareModuleInput(input_layouts=(Shard(1), Replicate()),
                                                desired_input_layouts=(Replicate(), Replicate())),
                "attention.wq": ColwiseParallel(use_local_output=False),
                "attention.wk": ColwiseParallel(use_local_output=False),
                "attention.wv": ColwiseParallel(use_local_output=False),
                "attention.wo": RowwiseParallel(output_layouts=Shard(1)),
                "ffn_norm": SequenceParallel(),
                "feed_forward": PrepareModuleInput(input_layouts=(Shard(1),), desired_input_layouts=(Replicate(),)),
                "feed_forward.w1": ColwiseParallel(),
                "feed_forward.w2": RowwiseParallel(output_layouts=Shard(1)),
                "feed_forward.w3": ColwiseParallel(),
            })
        mp = MixedPrecisionPolicy(param_dtype=torch.bfloat16, reduce_dtype=torch.bfloat16)  # same reduce precision as the repo arm (and as the reference FSDP recipe, resnet_fsdp_training.py:200-203)
        for block in model.layers:
            fully_shard(block, mesh=dp_mesh, mp_policy=mp)
        fully_shard(model, mesh=dp_mesh, mp_policy=mp)
        opt = torch.optim.AdamW(model.parameters(), lr=3e-4, betas=(0.9, 0.95), weight_decay=0.1, foreach=True)

        def step(x, y):
            opt.zero_grad()
            loss = F.cross_entropy(model(x).view(-1, vocab), y.reshape(-1))
            loss.backward()
            opt.step()
            return loss

    n_b = 4
    g = torch.Generator().manual_seed(1234 + dp_rank)   # TP peers of one dp replica see identical tokens
    toks = [torch.randint(0, vocab, (B, S + 1), generator=g).to(dev) for _ in range(n_b)]
    for i in range(a.warmup):
        l = step(toks[i % n_b][:, :-1], toks[i % n_b][:, 1:])
        if a.verbose and rank == 0:
            print(f"warmup {i}: loss {float(l.detach()):.4f}", flush=True)
    dist.barrier(device_ids=[local]); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(a.steps):
        k = (a.warmup + i) % n_b
        loss = step(toks[k][:, :-1], toks[k][:, 1:])
        if a.verbose and rank == 0:
            print(f"step {i}: loss {float(loss.detach()):.4f}", flush=True)
    e1.record()
    dist.barrier(device_ids=[local]); torch.cuda.synchronize()
    t = torch.tensor([e0.elapsed_time(e1)], device=dev, dtype=torch.float64)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms = float(t) / a.steps
    tokens = dp * B * S
    peak = torch.cuda.max_memory_allocated() / 2 ** 30
    if rank == 0:
        print(json.dumps({"metric": "llama2_13b_tp_fsdp_train_tokens_per_s", "value": round(tokens / (ms / 1e3), 1), "unit": "tokens/s",
                          "n_gpus": world, "steps": a.steps, "warmup": a.warmup, "ms_per_step": round(ms, 3), "higher_is_better": True,
                          "dtype": "bf16", "data": "synthetic tokens, random-init weights", "impl": a.impl,
                          "config": {"model": a.model, "n_layers": cfg["n_layers"], "seq_len": S, "global_batch": dp * B,
                                     "parallelism": f"tp{tp}{'' if a.no_sp else '+sp'} x fsdp{dp}", "fused_tp": bool(a.fused), "vocab_parallel_loss": bool(a.vocab_parallel_loss)},
                          "peak_mem_gib": round(peak, 1), "loss_last": round(float(loss.detach()), 4)}), flush=True)
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
