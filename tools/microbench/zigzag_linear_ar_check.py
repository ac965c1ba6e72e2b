"""2+ GPUs: zig-zag ring attention (forward + backward) against unsharded flash attention, and the backward of the fused
row-parallel linear (GEMM -> all-reduce, ``tp_fused.linear_ar``) against an fp32 reference."""
import os
import sys
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parents[2]))
import torch
import torch.distributed as dist

from b200dist import ops
from b200dist.comm import Communicator
from b200dist.parallel import tp_fused
from b200dist.parallel.ring_attention import zigzag_ring_attention, zigzag_shard

rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
dev = torch.device("cuda", local)
comm = Communicator(None)
fails = 0


def check(name, got, ref, atol, rtol):
    global fails
    err = (got.float() - ref.float()).abs()
    ok = bool((err <= atol + rtol * ref.float().abs()).all()) and bool(torch.isfinite(got.float()).all())
    fails += (not ok)
    if rank == 0 or not ok:
        print(f"[{'PASS' if ok else 'FAIL'}] rank{rank} {name}: max err {float(err.max()):.4g} (|ref| max {float(ref.abs().max()):.3g})", flush=True)


# ---- zig-zag ring attention: same q, k, v on every rank; rank r owns blocks r and 2 world - 1 - r
for (B, S, H, Hkv, hd) in ((1, 256 * 2 * world, 8, 8, 128), (2, 128 * 2 * world, 8, 2, 64)):
    torch.manual_seed(5)
    q = torch.randn(B, S, H, hd, device=dev).bfloat16(); k = torch.randn(B, S, Hkv, hd, device=dev).bfloat16()
    v = torch.randn(B, S, Hkv, hd, device=dev).bfloat16(); do = torch.randn(B, S, H, hd, device=dev).bfloat16()
    qr, kr, vr = [t.clone().requires_grad_(True) for t in (q, k, v)]
    ref = ops.attention(qr, kr, vr, causal=True)
    ref.backward(do)
    ql, kl, vl = [zigzag_shard(t, world, rank).contiguous().requires_grad_(True) for t in (q, k, v)]
    out = zigzag_ring_attention(ql, kl, vl, comm)
    out.backward(zigzag_shard(do, world, rank).contiguous())
    tag = f"B{B} S{S} H{H}/{Hkv} hd{hd}"
    check(f"zig-zag ring attention O {tag}", out, zigzag_shard(ref.detach(), world, rank), 2e-2, 2e-2)
    check(f"zig-zag ring attention dQ {tag}", ql.grad, zigzag_shard(qr.grad, world, rank), 5e-2, 3e-2)
    check(f"zig-zag ring attention dK {tag}", kl.grad, zigzag_shard(kr.grad, world, rank), 5e-2, 3e-2)
    check(f"zig-zag ring attention dV {tag}", vl.grad, zigzag_shard(vr.grad, world, rank), 5e-2, 3e-2)

# ---- fused row-parallel linear with replicated output: forward and backward
M, K, Nn = 512 * world, 1024, 768            # K = this rank's slice of the contraction dimension
torch.manual_seed(21 + rank)
x = (torch.randn(M, K, device=dev) * 0.5).bfloat16().requires_grad_(True)
w = (torch.randn(Nn, K, device=dev) * 0.05).bfloat16().requires_grad_(True)
torch.manual_seed(99)
bias = (torch.randn(Nn, device=dev) * 0.1).bfloat16().requires_grad_(True)    # replicated, like the upstream gradient
dy = torch.randn(M, Nn, device=dev).bfloat16()       # replicated upstream gradient
y = tp_fused.linear_ar(comm, x, w, bias)
y.backward(dy)
yr = x.detach().float() @ w.detach().float().t()
dist.all_reduce(yr)
yr = yr + bias.detach().float()
check("linear_ar forward", y, yr, 5e-2, 2e-2)
check("linear_ar dX", x.grad, dy.float() @ w.detach().float(), 5e-2, 2e-2)
check("linear_ar dW", w.grad, dy.float().t() @ x.detach().float(), 0.5, 3e-2)
check("linear_ar dbias", bias.grad, dy.float().sum(0), 0.5, 3e-2)
t = torch.tensor([fails], device=dev); dist.all_reduce(t)
if rank == 0:
    print(f"== zig-zag ring / linear_ar check world={world}: failures {int(t)}", flush=True)
dist.destroy_process_group()
