"""2+ GPUs: the fused all-gather -> GEMM with an output width that is not a multiple of the 256-column tile (the LM head of
the TP plan: N = vocab / T = 8000 at T = 4), against all-gather + fp32 matmul; and the GEMM -> reduce-scatter with that K."""
import os
import sys
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parents[2]))
import torch
import torch.distributed as dist

from b200dist.comm import Communicator
from b200dist.parallel import tp_fused

rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
dev = torch.device("cuda", local)
comm = Communicator(None)
fails = 0
for (m, K, N) in ((512, 5120, 8000), (256, 1024, 8), (256, 1024, 264), (512, 4096, 32000 // world)):
    torch.manual_seed(7)
    xs_all = (torch.randn(world * m, K, device=dev) * 0.5).bfloat16()      # same on every rank
    torch.manual_seed(11 + rank)
    w = (torch.randn(N, K, device=dev) * 0.05).bfloat16()
    y, xg = tp_fused._ag_gemm(comm, xs_all[rank * m:(rank + 1) * m].contiguous(), w)
    ref = xs_all.float() @ w.float().t()
    err = float((y.float() - ref).abs().max()); tol = 0.02 * float(ref.abs().max()) + 0.02
    ok = err <= tol and bool(torch.isfinite(y).all())
    fails += (not ok)
    # backward shape of the same layer: dX = reduce_scatter_rows(dY @ W), K = N of the forward
    dy = (torch.randn(world * m, N, device=dev) * 0.5).bfloat16()
    dx = tp_fused._gemm_rs(comm, dy, w, b_mn=True)
    full = dy.float() @ w.float()
    dist.all_reduce(full)
    r2 = full[rank * m:(rank + 1) * m]
    err2 = float((dx.float() - r2).abs().max()); tol2 = 0.02 * float(r2.abs().max()) + 0.05
    ok2 = err2 <= tol2
    fails += (not ok2)
    if rank == 0:
        print(f"[{'PASS' if ok else 'FAIL'}] ag_gemm m={m} K={K} N={N}: max err {err:.4g} (tol {tol:.3g}) | "
              f"[{'PASS' if ok2 else 'FAIL'}] gemm_rs K={N}: max err {err2:.4g} (tol {tol2:.3g})", flush=True)
t = torch.tensor([fails], device=dev); dist.all_reduce(t)
if rank == 0:
    print(f"== ragged fused GEMM check world={world}: failures {int(t)}", flush=True)
dist.destroy_process_group()
