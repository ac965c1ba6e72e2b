#!/usr/bin/env python
"""Pipeline-parallel training of the PipelineTransformer with GPipe or 1F1B, optionally x data parallel
(/root/reference/scripts/04_pipeline_parallel_pp/03_pipeline_training.py: --num-steps --batch-size --seq-len
--num-microbatches --schedule --dim --layers-per-stage).  Any world size divisible by --pp works (the reference
asserts world == 4, :138-142); each rank builds only its own stage; the schedule is built once; first 3 steps are skipped
in the tokens/s average as in the reference (:260-261)."""
import argparse
import sys
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from _common import StepTimer, max_over_ranks, setup, torch  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--num-steps", type=int, default=20)
    ap.add_argument("--batch-size", type=int, default=16)
    ap.add_argument("--seq-len", type=int, default=128)
    ap.add_argument("--num-microbatches", type=int, default=4)
    ap.add_argument("--schedule", default="1f1b", choices=["gpipe", "1f1b"])
    ap.add_argument("--dim", type=int, default=256)
    ap.add_argument("--layers-per-stage", type=int, default=2)
    ap.add_argument("--pp", type=int, default=0, help="pipeline stages (default: world size, or 4 if divisible)")
    ap.add_argument("--vocab", type=int, default=10000)
    a = ap.parse_args()
    rank, world, local, device = setup()
    from b200dist import ops
    from b200dist.models.pp_transformer import build_stage
    from b200dist.ops import FusedAdamW
    from b200dist.parallel import init_device_mesh
    from b200dist.parallel.pp import PipelineStage, Schedule1F1B, ScheduleGPipe, bubble_fraction
    from b200dist.utils import cleanup_distributed

    pp = a.pp or (4 if world % 4 == 0 else world)
    mesh = init_device_mesh(device.type, (world // pp, pp), ("dp", "pp"))
    stage_idx, dp_rank = mesh.get_local_rank("pp"), mesh.get_local_rank("dp")
    torch.manual_seed(stage_idx)  # replicas of a stage start identical
    dtype = torch.bfloat16 if device.type == "cuda" else torch.float32
    mod = build_stage(stage_idx, pp, a.vocab, a.dim, 8, a.layers_per_stage).to(device).to(dtype)
    stage = PipelineStage(mod, stage_idx, pp, device, group=mesh.get_group("pp"),
                          dp_group=mesh.get_group("dp") if mesh.size("dp") > 1 else None)
    opt = FusedAdamW(stage.parameters(), lr=1e-4)
    loss_fn = lambda logits, tgt: ops.cross_entropy(logits, tgt)
    sched = (Schedule1F1B if a.schedule == "1f1b" else ScheduleGPipe)(stage, a.num_microbatches, loss_fn)
    timer = StepTimer(device)
    for step in range(a.num_steps):
        g = torch.Generator().manual_seed(step * 1000 + dp_rank)
        tok = torch.randint(0, a.vocab, (a.batch_size, a.seq_len + 1), generator=g).to(device)
        opt.zero_grad()
        losses = []
        with timer:
            if stage.is_first:
                sched.step(tok[:, :-1])
            elif stage.is_last:
                sched.step(target=tok[:, 1:], losses=losses)
            else:
                sched.step()
            opt.step()
        if stage.is_last and dp_rank == 0 and step % 5 == 0:
            print(f"step {step}: loss {float(sum(losses) / len(losses)):.4f}  {timer.times[-1] * 1e3:.1f} ms", flush=True)
    avg = max_over_ranks(timer.mean(skip=3), device)
    if rank == 0:
        toks = a.batch_size * a.seq_len * mesh.size("dp")
        print(f"{a.schedule}: {avg * 1e3:.1f} ms/step  {toks / avg:.0f} tokens/s  (pp={pp} dp={mesh.size('dp')}, "
              f"bubble {bubble_fraction(pp, a.num_microbatches):.1%})", flush=True)
    cleanup_distributed()


if __name__ == "__main__":
    main()
