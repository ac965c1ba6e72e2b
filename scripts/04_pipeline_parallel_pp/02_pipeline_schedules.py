#!/usr/bin/env python
"""GPipe vs 1F1B on a four-block MLP (dim 512), one block per rank
(/root/reference/scripts/04_pipeline_parallel_pp/02_pipeline_schedules.py).  The reference builds its schedules without a
loss function, so its comparison times forward only; here both schedules run forward AND backward."""
import argparse
import sys
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from _common import StepTimer, max_over_ranks, setup, torch  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--batch", type=int, default=64)
    ap.add_argument("--microbatches", type=int, default=8)
    ap.add_argument("--iters", type=int, default=10)
    a = ap.parse_args()
    rank, world, local, device = setup()
    from b200dist.models import FourBlockMLP
    from b200dist.parallel.pp import PipelineStage, Schedule1F1B, ScheduleGPipe, bubble_fraction, schedule_actions
    from b200dist.utils import cleanup_distributed
    from b200dist.utils.distributed import barrier

    assert 4 % world == 0, "run with 1, 2 or 4 ranks"
    torch.manual_seed(0)
    stage = PipelineStage(FourBlockMLP(512).stage(rank, world).to(device), rank, world, device)
    loss_fn = lambda out, tgt: ((out - tgt) ** 2).mean()
    x, y = torch.randn(a.batch, 512, device=device), torch.randn(a.batch, 512, device=device)
    for cls in (ScheduleGPipe, Schedule1F1B):
        sched = cls(stage, a.microbatches, loss_fn)
        if rank == 0:
            print(f"{cls.kind}: stage-0 order {''.join(k + str(i) + ' ' for k, i in schedule_actions(cls.kind, 0, world, a.microbatches))}", flush=True)
        timer = StepTimer(device)
        for it in range(a.iters):
            for p in stage.parameters():
                p.grad = None
            barrier()
            with timer:
                if rank == 0:
                    sched.step(x)
                elif rank == world - 1:
                    sched.step(target=y, losses=[])
                else:
                    sched.step()
        ms = max_over_ranks(timer.mean(skip=2), device) * 1e3
        if rank == 0:
            print(f"{cls.kind}: {ms:.2f} ms / step ({a.microbatches} microbatches, analytic bubble {bubble_fraction(world, a.microbatches):.1%})", flush=True)
    cleanup_distributed()


if __name__ == "__main__":
    main()
