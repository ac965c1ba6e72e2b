"""Kernel-time breakdown of one Llama training step (torch.profiler, CUDA activities only).
    python tools/profile_step.py [--n-layers 8] > gpurun_out/step_kernels.txt"""
import argparse
import sys
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import torch
from torch.profiler import ProfilerActivity, profile

from b200dist.engine import Trainer
from b200dist.models.llama2 import Transformer
from b200dist.ops import FusedAdamW

ap = argparse.ArgumentParser()
ap.add_argument("--n-layers", type=int, default=8)
ap.add_argument("--seq", type=int, default=4096)
ap.add_argument("--resnet", default=None, help="profile a ResNet step instead (arch name)")
ap.add_argument("--batch", type=int, default=128)
ap.add_argument("--no-overlap", action="store_true", help="AdamW after backward instead of from the gradient hooks")
a = ap.parse_args()
dev = torch.device("cuda", 0)
torch.manual_seed(0)
if a.resnet:
    from b200dist import ops
    from b200dist.models import build_resnet
    from b200dist.ops import FusedSGD

    m = build_resnet(a.resnet).to(dev).to(torch.bfloat16)
    opt = FusedSGD(m.parameters(), lr=0.1, momentum=0.9, weight_decay=1e-5)
    x = torch.rand(a.batch, 3, 224, 224, device=dev).bfloat16(); y = torch.zeros(a.batch, dtype=torch.long, device=dev)

    def step():
        opt.zero_grad(); ops.cross_entropy(m(x), y).backward(); opt.step()
    for _ in range(3):
        step()
    torch.cuda.synchronize()
    with profile(activities=[ProfilerActivity.CUDA]) as prof:
        for _ in range(2):
            step()
        torch.cuda.synchronize()
    ev = prof.key_averages()
    rows = sorted(((e.key, e.count, e.device_time_total) for e in ev if e.device_time_total > 0), key=lambda r: -r[2])
    tot = sum(r[2] for r in rows)
    print(f"# {a.resnet} batch {a.batch}, 2 steps; total device time {tot / 2e3:.2f} ms/step")
    for k, c, t in rows[:30]:
        print(f"{t / 2e3:9.3f} ms/step {100 * t / tot:5.1f}%  x{c // 2:<4d} {k[:110]}")
    sys.exit(0)
m = Transformer.from_name("llama2-7b", device=dev, dtype=torch.bfloat16, n_layers=a.n_layers, max_seq_len=a.seq)
tr = Trainer(m, FusedAdamW(m.parameters(), lr=1e-4, overlap_backward=not a.no_overlap))
tok = torch.randint(0, 32000, (1, a.seq + 1), device=dev)
for _ in range(3):
    tr.train_step(tok[:, :-1], tok[:, 1:])
torch.cuda.synchronize()
with profile(activities=[ProfilerActivity.CUDA]) as prof:
    for _ in range(2):
        tr.train_step(tok[:, :-1], tok[:, 1:])
    torch.cuda.synchronize()
ev = prof.key_averages()
rows = sorted(((e.key, e.count, e.device_time_total) for e in ev if e.device_time_total > 0), key=lambda r: -r[2])
tot = sum(r[2] for r in rows)
print(f"# llama2-7b shape, {a.n_layers} layers, seq {a.seq}, 2 steps; total device time {tot / 2e3:.2f} ms/step")
for k, c, t in rows[:40]:
    print(f"{t / 2e3:9.3f} ms/step {100 * t / tot:5.1f}%  x{c // 2:<4d} {k[:110]}")
