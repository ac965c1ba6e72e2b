#!/usr/bin/env python
"""BASELINE config 5a: Llama-2 7B, pipeline parallel (4 stages) x data parallel (2), 1F1B, bf16, seq 4096 -> tokens/s.

    torchrun --nproc-per-node 8 --master-addr 127.0.0.1 bench/pp_bench.py --pp 4 --dp 2 [--impl reference]

b200dist arm : models.TransformerStage per rank (own kernels), parallel.pp.PipelineStage + Schedule1F1B over NCCL p2p,
               stage gradients averaged across the dp group by the peer-memory bucket reducer, FusedAdamW.
reference arm: the reference's own Llama blocks (baseline/_ref/reference_src/fsdp_tp/llama2_model.py) cut into stages and
               driven by torch.distributed.pipelining (PipelineStage + Schedule1F1B, what 03_pipeline_training.py:183-240
               uses) composed with torch DDP on a (dp, pp) mesh, torch AdamW — BASELINE.md §4 row 5a.  The schedule is
               built once (the recipe rebuilds it every step, :236-240).
Global batch = dp x microbatches x mb_size sequences.  Device-timed, max over ranks; one JSON line from rank 0."""
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
from torch import nn  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--pp", type=int, default=4)
    ap.add_argument("--dp", type=int, default=2)
    ap.add_argument("--model", default="llama2-7b")
    ap.add_argument("--n-layers", type=int, default=0)
    ap.add_argument("--seq-len", type=int, default=4096)
    ap.add_argument("--microbatches", type=int, default=8)
    ap.add_argument("--mb-size", type=int, default=1)
    ap.add_argument("--schedule", default="1f1b", choices=["1f1b", "gpipe"])
    ap.add_argument("--steps", type=int, default=6)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200dist", choices=["b200dist", "reference"])
    ap.add_argument("--verbose", action="store_true", help="print the loss of every step (adds a sync per step)")
    a = ap.parse_args()
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    assert world == a.pp * a.dp, f"world {world} != pp {a.pp} x dp {a.dp}"
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    from b200dist.models.llama2 import PRESETS

    cfg = dict(PRESETS[a.model]); cfg["max_seq_len"] = a.seq_len
    if a.n_layers:
        cfg["n_layers"] = a.n_layers
    vocab = cfg["vocab_size"]
    M, S = a.microbatches, a.seq_len
    torch.manual_seed(0)
    if a.impl == "b200dist":
        from b200dist import ops
        from b200dist.models import ModelArgs, TransformerStage
        from b200dist.ops import FusedAdamW
        from b200dist.parallel import init_device_mesh
        from b200dist.parallel.pp import PipelineStage, Schedule1F1B, ScheduleGPipe

        mesh = init_device_mesh("cuda", (a.dp, a.pp), ("dp", "pp"))
        stage_idx, dp_rank = mesh.get_local_rank("pp"), mesh.get_local_rank("dp")
        prev = torch.get_default_dtype(); torch.set_default_dtype(torch.bfloat16)
        with torch.device(dev):
            mod = TransformerStage(ModelArgs(**cfg), stage_idx, a.pp)
        torch.set_default_dtype(prev)
        stage = PipelineStage(mod, stage_idx, a.pp, device=dev, group=mesh.get_group("pp"),
                              dp_group=mesh.get_group("dp") if a.dp > 1 else None)
        loss_fn = lambda out, tgt: ops.cross_entropy(out.view(-1, vocab), tgt.reshape(-1))  # noqa: E731
        sched = (Schedule1F1B if a.schedule == "1f1b" else ScheduleGPipe)(stage, M, loss_fn=loss_fn)
        opt = FusedAdamW(stage.parameters(), lr=3e-4, betas=(0.9, 0.95), weight_decay=0.1)
        is_first, is_last = stage.is_first, stage.is_last
    else:
        from baseline.install_ref import ensure_reference, why_unavailable
        ref_root = ensure_reference()
        if ref_root is None:
            if rank == 0:
                print(json.dumps({"impl": "reference", "unavailable": why_unavailable()}))
            dist.destroy_process_group(); return
        sys.path.insert(0, str(ref_root / "reference_src" / "fsdp_tp"))
        import torch.nn.functional as F
        from llama2_model import ModelArgs, RMSNorm, TransformerBlock, precompute_freqs_cis  # the reference's own classes
        from torch.distributed.device_mesh import init_device_mesh
        from torch.distributed.pipelining import PipelineStage, Schedule1F1B, ScheduleGPipe

        mesh = init_device_mesh("cuda", (a.dp, a.pp), mesh_dim_names=("dp", "pp"))
        stage_idx, dp_rank = mesh.get_local_rank("pp"), mesh.get_local_rank("dp")
        margs = ModelArgs(**cfg)
        per = (margs.n_layers + a.pp - 1) // a.pp

        class RefStage(nn.Module):   # the reference Transformer.forward (:439-448) restricted to this stage's modules
            def __init__(self):
                super().__init__()
                self.first, self.last = stage_idx == 0, stage_idx == a.pp - 1
                if self.first:
                    self.tok_embeddings = nn.Embedding(margs.vocab_size, margs.dim)
                self.layers = nn.ModuleList(TransformerBlock(i, margs) for i in
                                            range(stage_idx * per, min(margs.n_layers, (stage_idx + 1) * per)))
                for l in self.layers:
                    l.init_weights()
                if self.last:
                    self.norm = RMSNorm(dim=margs.dim, eps=margs.norm_eps)
                    self.output = nn.Linear(margs.dim, margs.vocab_size, bias=False)
                self.freqs = None

            def forward(self, x):
                h = self.tok_embeddings(x) if self.first else x
                if self.freqs is None:
                    self.freqs = precompute_freqs_cis(margs.dim // margs.n_heads, margs.max_seq_len * 2).to(h.device)
                fc = self.freqs[0:h.shape[1]]
                for layer in self.layers:
                    h = layer(h, fc)
                return self.output(self.norm(h)).float() if self.last else h

        with torch.device(dev):
            mod = RefStage()
        mod = mod.to(torch.bfloat16)
        if a.dp > 1:
            mod = nn.parallel.DistributedDataParallel(mod, device_ids=[local], process_group=mesh.get_group("dp"))
        stage = PipelineStage(mod, stage_idx, a.pp, dev, group=mesh.get_group("pp"))
        loss_fn = lambda out, tgt: F.cross_entropy(out.view(-1, vocab), tgt.reshape(-1))  # noqa: E731
        sched = (Schedule1F1B if a.schedule == "1f1b" else ScheduleGPipe)(stage, M, loss_fn=loss_fn)
        opt = torch.optim.AdamW(mod.parameters(), lr=3e-4, betas=(0.9, 0.95), weight_decay=0.1, foreach=True)
        is_first, is_last = stage_idx == 0, stage_idx == a.pp - 1

    n_b = 4
    g = torch.Generator().manual_seed(1234 + dp_rank)   # every stage of a dp replica sees the same stream
    toks = [torch.randint(0, vocab, (M * a.mb_size, S + 1), generator=g).to(dev) for _ in range(n_b)]
    last_loss = [None]

    def step(i):
        t = toks[i % n_b]
        opt.zero_grad()
        if is_first and is_last:
            losses = []; sched.step(t[:, :-1], target=t[:, 1:], losses=losses); last_loss[0] = losses
        elif is_first:
            sched.step(t[:, :-1])
        elif is_last:
            losses = []; sched.step(target=t[:, 1:], losses=losses); last_loss[0] = losses
        else:
            sched.step()
        opt.step()
        if a.verbose and is_last and dp_rank == 0 and last_loss[0]:
            print(f"step {i}: loss {float(torch.stack([l.float() for l in last_loss[0]]).mean()):.4f}", flush=True)

    for i in range(a.warmup):
        step(i)
    dist.barrier(device_ids=[local]); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(a.steps):
        step(a.warmup + i)
    e1.record()
    dist.barrier(device_ids=[local]); torch.cuda.synchronize()
    t = torch.tensor([e0.elapsed_time(e1)], device=dev, dtype=torch.float64)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    lossv = torch.zeros(1, device=dev)
    if is_last and dp_rank == 0 and last_loss[0]:
        lossv += torch.stack([l.float() for l in last_loss[0]]).mean()
    dist.all_reduce(lossv)
    ms = float(t) / a.steps
    tokens = a.dp * M * a.mb_size * S
    if rank == 0:
        print(json.dumps({"metric": "llama2_7b_pp_dp_train_tokens_per_s", "value": round(tokens / (ms / 1e3), 1), "unit": "tokens/s",
                          "n_gpus": world, "steps": a.steps, "warmup": a.warmup, "ms_per_step": round(ms, 3), "higher_is_better": True,
                          "dtype": "bf16", "data": "synthetic tokens, random-init weights", "impl": a.impl,
                          "config": {"model": a.model, "n_layers": cfg["n_layers"], "seq_len": S, "global_batch": a.dp * M * a.mb_size,
                                     "microbatches": M, "schedule": a.schedule, "parallelism": f"pp{a.pp} x dp{a.dp}"},
                          "loss_last": round(float(lossv), 4)}), flush=True)
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
