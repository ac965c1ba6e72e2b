"""Multi-GPU correctness + bandwidth check of the peer-memory collectives against NCCL.

    torchrun --nproc-per-node 2 --master-addr 127.0.0.1 tools/comm_check.py
"""
from __future__ import annotations

import os
import sys
from pathlib import Path

import torch
import torch.distributed as dist

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))

from b200dist.comm import Communicator  # noqa: E402


def ev_time(fn, iters=10, warmup=3):
    for _ in range(warmup):
        fn()
    torch.cuda.synchronize()
    dist.barrier()
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    s.record()
    for _ in range(iters):
        fn()
    e.record()
    torch.cuda.synchronize()
    t = torch.tensor([s.elapsed_time(e) / iters], device="cuda")
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def main():
    rank = int(os.environ["RANK"]); world = int(os.environ["WORLD_SIZE"]); local = int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    comm = Communicator(None, staging_bytes=1 << 30)
    dev = torch.device("cuda", local)
    fails = 0

    def check(name, got, ref, tol=0.0):
        nonlocal fails
        err = (got.float() - ref.float()).abs().max().item()
        ok = err <= tol
        fails += (not ok)
        if rank == 0 or not ok:
            print(f"[{'PASS' if ok else 'FAIL'}] rank{rank} {name}: max_err={err:.4g}", flush=True)

    torch.manual_seed(100 + rank)
    comm.barrier()
    torch.cuda.synchronize()
    for dt in (torch.float32, torch.bfloat16):
        for n in (8 * world, 1024, 100003, 1 << 20, 5 * (1 << 20) + 8 * world):
            x = torch.randint(-8, 8, (n,), device=dev).to(dt)   # integer payload: sums are exact
            ref = x.clone(); dist.all_reduce(ref)
            for algo in (1, 2, 4):   # 4 = NVLS (multimem through the switch) when the platform offers it, else two-shot
                y = x.clone()
                comm.all_reduce(y, algo=algo)
                check(f"all_reduce {dt} n={n} algo={algo}", y, ref)
            ys = x.clone()
            comm.all_reduce(ys, scale=0.5, algo=4)     # the scale inside the NVLS kernel (DDP averages this way)
            check(f"all_reduce scaled in place {dt} n={n} algo=4", ys, ref * 0.5)
            y32 = torch.empty(n, device=dev, dtype=torch.float32)
            comm.all_reduce(x, scale=0.5, out=y32)
            check(f"all_reduce scale+cast {dt} n={n}", y32, ref.float() * 0.5)
        # push (latency) path: default algo for small messages; back-to-back calls alternate the slot set
        for it in range(5):
            n = 8 * (it + 1) * 37
            x = torch.randint(-8, 8, (n,), device=dev).to(dt)
            ref = x.clone(); dist.all_reduce(ref)
            y = x.clone(); comm.all_reduce(y)
            check(f"all_reduce push {dt} n={n} call {it}", y, ref)
            yo = torch.empty(n, device=dev, dtype=torch.float32 if dt == torch.bfloat16 else torch.bfloat16)
            comm.all_reduce(x, scale=0.25, out=yo)
            check(f"all_reduce push cast {dt} n={n}", yo, (ref.float() * 0.25).to(yo.dtype))
        keep_push = comm.push_threshold
        for push in (0, keep_push):                            # pull kernels, then the push (latency) kernels
            comm.push_threshold = push
            tag = "push" if push else "pull"
            for n in (8 * 33, 8 * 32 * 5 + 8 * 7, 1 << 18):      # reduce-scatter tails around the 32-vector fp32 layout
                xin = torch.randint(-8, 8, (n * world,), device=dev).to(dt)
                o = torch.empty(n, device=dev, dtype=dt); r = torch.empty(n, device=dev, dtype=dt)
                dist.reduce_scatter_tensor(r, xin)
                comm.reduce_scatter(o, xin)
                check(f"reduce_scatter {tag} {dt} n={n}", o, r)
            for it in range(4):                                  # back-to-back small collectives alternate the slot sets
                n = 8 * (it + 1) * 41
                xin = torch.randint(-8, 8, (n * world,), device=dev).to(dt)
                r = torch.empty(n, device=dev, dtype=dt); dist.reduce_scatter_tensor(r, xin)
                o32 = torch.empty(n, device=dev, dtype=torch.float32); comm.reduce_scatter(o32, xin, scale=0.5)
                check(f"reduce_scatter {tag} scale+cast {dt} n={n} call {it}", o32, r.float() * 0.5)
                x = torch.randint(-8, 8, (n,), device=dev).to(dt)
                ref = torch.empty(n * world, device=dev, dtype=dt); dist.all_gather_into_tensor(ref, x)
                out = torch.empty(n * world, device=dev, dtype=dt); comm.all_gather(out, x)
                check(f"all_gather {tag} {dt} n={n} call {it}", out, ref)
                root = it % world
                b = torch.randint(-8, 8, (n,), device=dev).to(dt); rb = b.clone()
                dist.broadcast(rb, src=root); comm.broadcast(b, src=root)
                check(f"broadcast {tag} {dt} n={n} root {root}", b, rb)
        comm.push_threshold = keep_push
        bb = torch.randint(-8, 8, ((4 << 20) // (2 if dt == torch.bfloat16 else 4) * world // world,), device=dev).to(dt)
        bb = bb[: bb.numel() // (8 * world) * (8 * world)].contiguous(); rbb = bb.clone()
        dist.broadcast(rbb, src=0)
        comm.broadcast2_min_world = 2
        comm.broadcast(bb, src=0)
        comm.broadcast2_min_world = 3
        check(f"broadcast two-stage {dt} {bb.numel()} elems", bb, rbb)
        prev = (comm.nvls_broadcast, comm.nvls_min_world)
        comm.nvls_broadcast, comm.nvls_min_world = True, 2      # broadcast through the switch (multimem.st), both roots
        for root in (0, world - 1):
            for rep in range(2):
                bn = torch.randint(-8, 8, (bb.numel(),), device=dev).to(dt); rbn = bn.clone()
                dist.broadcast(rbn, src=root)
                comm.broadcast(bn, src=root)
                check(f"broadcast NVLS {dt} root {root} call {rep}", bn, rbn)
        comm.nvls_broadcast, comm.nvls_min_world = prev
        n = 1 << 18
        x = torch.randint(-8, 8, (n,), device=dev).to(dt)
        out = torch.empty(n * world, device=dev, dtype=dt); ref = torch.empty_like(out)
        dist.all_gather_into_tensor(ref, x)
        comm.all_gather(out, x)
        check(f"all_gather {dt}", out, ref)
        xin = torch.randint(-8, 8, (n * world,), device=dev).to(dt)
        o = torch.empty(n, device=dev, dtype=dt); r = torch.empty(n, device=dev, dtype=dt)
        dist.reduce_scatter_tensor(r, xin)
        comm.reduce_scatter(o, xin)
        check(f"reduce_scatter {dt}", o, r)
        b = torch.randint(-8, 8, (12345,), device=dev).to(dt); rb = b.clone()
        dist.broadcast(rb, src=world - 1)
        comm.broadcast(b, src=world - 1)
        check(f"broadcast {dt}", b, rb)
        a2 = torch.randint(-8, 8, (n * world,), device=dev).to(dt)
        oa = torch.empty_like(a2); ra = torch.empty_like(a2)
        dist.all_to_all_single(ra, a2)
        comm.all_to_all(oa, a2)
        check(f"all_to_all {dt}", oa, ra)
    if rank == 0:
        print(f"[INFO] NVLS multicast path active: {comm._nvls}  {getattr(comm, '_nvls_error', '')}", flush=True)
    xr = comm.nvls_tensor(1 << 20, torch.bfloat16)          # zero-copy path: tensor inside the multicast buffer
    if xr is not None:
        src = torch.randint(-8, 8, (1 << 20,), device=dev).to(torch.bfloat16)
        ref = src.clone(); dist.all_reduce(ref)
        xr.copy_(src); comm.all_reduce(xr, algo=4)
        check("all_reduce NVLS registered tensor (in place, zero copy)", xr, ref)
    # ---- reduce-scatter / all-gather through the switch (multimem.ld_reduce / multimem.st), forced on at any world size ----
    prev_nv = (comm.nvls_min_world, comm.nvls_threshold)
    comm.nvls_min_world, comm.nvls_threshold = 2, 0
    keep_push, comm.push_threshold = comm.push_threshold, 0      # small sizes must reach the switch kernels here
    if comm._nvls_ready():
        for dt in (torch.float32, torch.bfloat16):
            for n in (8 * 33, 1 << 18, (1 << 20) + 8 * 5):
                for rep in range(2):    # back-to-back calls reuse the multicast region: the depart handshake must hold
                    xin = torch.randint(-8, 8, (n * world,), device=dev).to(dt)
                    r = torch.empty(n, device=dev, dtype=dt); dist.reduce_scatter_tensor(r, xin)
                    o = torch.empty(n, device=dev, dtype=dt); comm.reduce_scatter(o, xin)
                    check(f"reduce_scatter NVLS {dt} n={n} call {rep}", o, r)
                    o32 = torch.empty(n, device=dev, dtype=torch.float32); comm.reduce_scatter(o32, xin, scale=0.5)
                    check(f"reduce_scatter NVLS scale+cast {dt} n={n} call {rep}", o32, r.float() * 0.5)
                    x = torch.randint(-8, 8, (n,), device=dev).to(dt)
                    ref = torch.empty(n * world, device=dev, dtype=dt); dist.all_gather_into_tensor(ref, x)
                    out = torch.empty(n * world, device=dev, dtype=dt); comm.all_gather(out, x)
                    check(f"all_gather NVLS {dt} n={n} call {rep}", out, ref)
        # "registered" tensors (Communicator.nvls_tensor): the collective runs where the tensor is, no staging copies
        n = 1 << 16
        xi = comm.nvls_tensor(n * world, torch.float32, channel=2)
        go = comm.nvls_tensor(n * world, torch.float32, channel=1)
        xb = comm.nvls_tensor(n, torch.float32, channel=3)
        if xi is not None and go is not None and xb is not None:
            for rep in range(2):
                srcv = torch.randint(-8, 8, (n * world,), device=dev).float(); xi.copy_(srcv)
                r = torch.empty(n, device=dev); dist.reduce_scatter_tensor(r, srcv)
                o = torch.empty(n, device=dev); comm.reduce_scatter(o, xi)
                check(f"reduce_scatter registered input call {rep}", o, r)
                x = torch.randint(-8, 8, (n,), device=dev).float()
                ref = torch.empty(n * world, device=dev); dist.all_gather_into_tensor(ref, x)
                comm.all_gather(go, x)
                check(f"all_gather registered output call {rep}", go, ref)
            keep_b, comm.nvls_broadcast = comm.nvls_broadcast, True
            for root in (0, world - 1):
                val = torch.randint(-8, 8, (n,), device=dev).float(); xb.copy_(val)
                rb = val.clone(); dist.broadcast(rb, src=root)
                comm.broadcast(xb, src=root)
                check(f"broadcast registered root {root}", xb, rb)
            comm.nvls_broadcast = keep_b
        # zero-copy forms on a dedicated multicast buffer (what FSDP uses for its gradient staging)
        n = 1 << 16
        mcb = comm.nvls_buffer(n * world * 4)
        if mcb is not None:
            gi = torch.randint(-8, 8, (n * world,), device=dev).to(torch.bfloat16)
            r = torch.empty(n, device=dev, dtype=torch.float32); dist.reduce_scatter_tensor(r, gi.float())
            mcb.tensor(torch.bfloat16, 0, n * world).copy_(gi)
            o = torch.empty(n, device=dev, dtype=torch.float32)
            comm.reduce_scatter_nvls_from(o, mcb, 0, torch.bfloat16, 1.0)
            check("reduce_scatter NVLS zero-copy bf16 -> fp32", o, r)
            # split handshakes (what FSDP / DDP use on their side streams): waits in one-warp kernels around the bulk kernel
            for rep in range(3):
                gi2 = torch.randint(-8, 8, (n * world,), device=dev).to(torch.bfloat16)
                r2 = torch.empty(n, device=dev, dtype=torch.float32); dist.reduce_scatter_tensor(r2, gi2.float())
                mcb.tensor(torch.bfloat16, 0, n * world).copy_(gi2)
                o2 = torch.empty(n, device=dev, dtype=torch.float32)
                comm.reduce_scatter_nvls_from(o2, mcb, 0, torch.bfloat16, 1.0, split=True)
                check(f"reduce_scatter NVLS split handshake call {rep}", o2, r2)
                ar = torch.randint(-8, 8, (n * world,), device=dev).to(torch.bfloat16)
                rar = ar.clone(); dist.all_reduce(rar)
                mcb.tensor(torch.bfloat16, 0, n * world).copy_(ar)
                comm.all_reduce_nvls_inplace(mcb, 0, n * world, torch.bfloat16, 1.0, channel=2, split=True)
                check(f"all_reduce NVLS in place, split handshake call {rep}", mcb.tensor(torch.bfloat16, 0, n * world), rar)
            sh = torch.randint(-8, 8, (n,), device=dev).to(torch.bfloat16)
            ref = torch.empty(n * world, device=dev, dtype=torch.bfloat16); dist.all_gather_into_tensor(ref, sh)
            comm.all_gather_nvls_into(mcb, 0, sh)
            torch.cuda.synchronize()
            check("all_gather NVLS zero-copy (replica in place)", mcb.tensor(torch.bfloat16, 0, n * world), ref)
            # fused reduce-scatter (in the switch) + AdamW vs the pull kernel on the same gradients
            g2 = torch.randn(n * world, device=dev).bfloat16()
            p0 = torch.randn(n, device=dev)
            pb = comm.symmetric_buffer(g2.numel() * 2); pb.tensor(torch.bfloat16).copy_(g2)
            ms0, mm0, vv0 = p0.clone(), torch.zeros(n, device=dev), torch.zeros(n, device=dev)
            lp0 = torch.empty(n, device=dev, dtype=torch.bfloat16)
            comm.rs_adamw_from(pb, 0, torch.bfloat16, n, 1.0 / world, ms0, lp0, mm0, vv0, lr=1e-2, beta1=0.9, beta2=0.95,
                               eps=1e-8, wd=0.1, step=1)
            mcb.tensor(torch.bfloat16, 0, n * world).copy_(g2)
            ms1, mm1, vv1 = p0.clone(), torch.zeros(n, device=dev), torch.zeros(n, device=dev)
            lp1 = torch.empty(n, device=dev, dtype=torch.bfloat16)
            comm.rs_adamw_nvls_from(mcb, 0, torch.bfloat16, n, 1.0 / world, ms1, lp1, mm1, vv1, lr=1e-2, beta1=0.9,
                                    beta2=0.95, eps=1e-8, wd=0.1, step=1)
            # the switch rounds the fp32 sum to bf16 once; step 1 of Adam is sign-like, so compare the moments
            check("rs_adamw NVLS exp_avg vs pull kernel", mm1, mm0, 2e-2 * float(mm0.abs().max()))
            check("rs_adamw NVLS master vs pull kernel", ms1, ms0, 2.5e-2)
    elif rank == 0:
        print("[INFO] NVLS unavailable: reduce-scatter / all-gather through the switch not checked", flush=True)
    comm.nvls_min_world, comm.nvls_threshold = prev_nv
    comm.push_threshold = keep_push
    # fused reduce-scatter + AdamW vs (NCCL reduce-scatter, torch AdamW)
    n = 1 << 16
    g = torch.randn(n * world, device=dev).bfloat16()
    p0 = torch.randn(n, device=dev)
    gr = torch.empty(n, device=dev, dtype=torch.float32)
    dist.reduce_scatter_tensor(gr, g.float())
    pr = p0.clone().requires_grad_(True)
    opt = torch.optim.AdamW([pr], lr=1e-2, betas=(0.9, 0.95), weight_decay=0.1)
    pr.grad = gr / world
    opt.step()
    buf = comm.symmetric_buffer(g.numel() * 2)
    buf.tensor(torch.bfloat16).copy_(g)
    master = p0.clone(); m = torch.zeros(n, device=dev); v = torch.zeros(n, device=dev)
    lowp = torch.empty(n, device=dev, dtype=torch.bfloat16)
    comm.rs_adamw_from(buf, 0, torch.bfloat16, n, 1.0 / world, master, lowp, m, v, lr=1e-2, beta1=0.9, beta2=0.95,
                       eps=1e-8, wd=0.1, step=1)
    check("rs_adamw master", master, pr.detach(), 2e-3)
    check("rs_adamw lowp", lowp, pr.detach(), 3e-2)
    # copy-engine staged variant and fp32 gradients must give bit-identical results to the pull kernel
    for in_dt, src_t in ((torch.bfloat16, g), (torch.float32, g.float())):
        b2 = comm.symmetric_buffer(src_t.numel() * src_t.element_size())
        b2.tensor(in_dt).copy_(src_t)
        outs = []
        for ce in (False, True):
            ms, mm, vv = p0.clone(), torch.zeros(n, device=dev), torch.zeros(n, device=dev)
            lp = torch.empty(n, device=dev, dtype=torch.bfloat16)
            comm.rs_adamw_from(b2, 0, in_dt, n, 1.0 / world, ms, lp, mm, vv, lr=1e-2, beta1=0.9, beta2=0.95, eps=1e-8,
                               wd=0.1, step=1, copy_engine=ce)
            outs.append((ms, lp))
        check(f"rs_adamw {in_dt} vs reference", outs[0][0], pr.detach(), 2e-3)
        check(f"rs_adamw {in_dt} copy-engine == pull kernel", outs[1][0], outs[0][0])
        check(f"rs_adamw {in_dt} copy-engine lowp", outs[1][1], outs[0][1])
    torch.cuda.synchronize()
    # bandwidth
    for mb in (1, 16, 256):
        n = mb * (1 << 20) // 2
        x = torch.randn(n, device=dev).bfloat16()
        t_own = ev_time(lambda: comm.all_reduce(x, algo=2))
        t_nvls = ev_time(lambda: comm.all_reduce(x, algo=4))
        t_nccl = ev_time(lambda: dist.all_reduce(x))
        bus = lambda t: 2 * (world - 1) / world * n * 2 / t / 1e6
        if rank == 0:
            print(f"[TIME] all_reduce {mb} MiB bf16 x{world}: two-shot {t_own*1e3:.1f} us busbw {bus(t_own):.0f} GB/s | "
                  f"nvls {t_nvls*1e3:.1f} us busbw {bus(t_nvls):.0f} GB/s | NCCL {t_nccl*1e3:.1f} us busbw {bus(t_nccl):.0f} GB/s", flush=True)
        out = torch.empty(n * world, device=dev, dtype=torch.bfloat16)
        t_own = ev_time(lambda: comm.all_gather(out, x))
        t_nccl = ev_time(lambda: dist.all_gather_into_tensor(out, x))
        busg = lambda t: (world - 1) * n * 2 / t / 1e6
        if rank == 0:
            print(f"[TIME] all_gather {mb} MiB/rank: own {t_own*1e3:.1f} us busbw {busg(t_own):.0f} GB/s | NCCL {t_nccl*1e3:.1f} us "
                  f"busbw {busg(t_nccl):.0f} GB/s", flush=True)
    x32 = torch.randn(64 << 20, device=dev)
    t_own = ev_time(lambda: comm.all_reduce(x32, algo=2)); t_nvls = ev_time(lambda: comm.all_reduce(x32, algo=4))
    t_nccl = ev_time(lambda: dist.all_reduce(x32))
    if rank == 0:
        print(f"[TIME] all_reduce 256 MiB fp32 x{world}: two-shot {t_own*1e3:.1f} us | nvls {t_nvls*1e3:.1f} us | NCCL {t_nccl*1e3:.1f} us", flush=True)
    del x32
    x = torch.zeros(8, device=dev)
    t_own = ev_time(lambda: comm.barrier(), iters=50)
    t_small = ev_time(lambda: comm.all_reduce(x), iters=50)
    t_nccl = ev_time(lambda: dist.all_reduce(x), iters=50)
    if rank == 0:
        print(f"[TIME] barrier {t_own*1e3:.1f} us | 32 B all_reduce own {t_small*1e3:.1f} us NCCL {t_nccl*1e3:.1f} us", flush=True)
    tot = torch.tensor([fails], device=dev)
    dist.all_reduce(tot)
    if rank == 0:
        print(f"== comm_check world={world}: total failures {int(tot.item())}", flush=True)
    dist.destroy_process_group()
    sys.exit(1 if int(tot.item()) else 0)


if __name__ == "__main__":
    main()
