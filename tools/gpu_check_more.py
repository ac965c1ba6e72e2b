"""Additional bring-up groups: memory-bound ops, optimizers, attention."""
from __future__ import annotations

import math


def _h():
    from tools.gpu_check import _check, _time
    return _check, _time


def group_norms(results):
    import torch
    import torch.nn.functional as F
    from b200dist.ops import native as N
    _check, _time = _h()
    dev = "cuda"
    torch.manual_seed(1)
    for dt, tol in ((torch.bfloat16, 3e-2), (torch.float32, 1e-4)):
        for rows, D in ((64, 256), (300, 4096), (33, 5120), (17, 64)):
            x = torch.randn(rows, D, device=dev).to(dt)
            w = (1 + 0.1 * torch.randn(D, device=dev)).to(dt)
            b = (0.1 * torch.randn(D, device=dev)).to(dt)
            dy = torch.randn(rows, D, device=dev).to(dt)
            xr = x.float().requires_grad_(True); wr = w.float().requires_grad_(True); br = b.float().requires_grad_(True)
            # RMSNorm (reference expression)
            yr = (xr * torch.rsqrt(xr.pow(2).mean(-1, keepdim=True) + 1e-5)) * wr
            yr.backward(dy.float())
            y, rstd = N.rmsnorm_fwd(x, w, 1e-5)
            dx, dw = N.rmsnorm_bwd(dy, x, w, rstd)
            _check(f"rmsnorm fwd {dt} {rows}x{D}", y, yr, tol, tol, results)
            _check(f"rmsnorm dx {dt} {rows}x{D}", dx, xr.grad, tol, tol, results)
            _check(f"rmsnorm dw {dt} {rows}x{D}", dw, wr.grad, tol * math.sqrt(rows), tol, results)
            xr.grad = None; wr.grad = None
            yl = F.layer_norm(xr, (D,), wr, br, 1e-5)
            yl.backward(dy.float())
            y, mean, rs = N.layernorm_fwd(x, w, b, 1e-5)
            dx, dw, db = N.layernorm_bwd(dy, x, w, mean, rs)
            _check(f"layernorm fwd {dt} {rows}x{D}", y, yl, tol, tol, results)
            _check(f"layernorm dx {dt} {rows}x{D}", dx, xr.grad, tol, tol, results)
            _check(f"layernorm dw {dt} {rows}x{D}", dw, wr.grad, tol * math.sqrt(rows), tol, results)
            _check(f"layernorm db {dt} {rows}x{D}", db, br.grad, tol * math.sqrt(rows), tol, results)
    x = torch.randn(8192, 4096, device=dev).bfloat16(); w = torch.ones(4096, device=dev).bfloat16()
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    ms = _time(lambda: N.rmsnorm_fwd(x, w, 1e-5), flush=flush)
    print(f"[TIME] rmsnorm_fwd 8192x4096 bf16: {ms*1e3:.1f} us  {2*x.numel()*2/ms/1e6:.0f} GB/s", flush=True)
    y, rstd = N.rmsnorm_fwd(x, w, 1e-5)
    ms = _time(lambda: N.rmsnorm_bwd(x, x, w, rstd), flush=flush)
    print(f"[TIME] rmsnorm_bwd 8192x4096 bf16: {ms*1e3:.1f} us  {3*x.numel()*2/ms/1e6:.0f} GB/s", flush=True)


def group_elementwise(results):
    import torch
    import torch.nn.functional as F
    from b200dist.ops import native as N
    _check, _time = _h()
    dev = "cuda"
    torch.manual_seed(2)
    # RoPE vs the reference complex formulation
    B, S, H, hd = 2, 96, 4, 128
    freqs = 1.0 / (10000.0 ** (torch.arange(0, hd, 2, device=dev)[: hd // 2].float() / hd))
    t = torch.arange(256, device=dev).float()
    fc = torch.polar(torch.ones(256, hd // 2, device=dev), torch.outer(t, freqs))
    fr = torch.view_as_real(fc).contiguous()
    for dt, tol in ((torch.bfloat16, 2e-2), (torch.float32, 1e-5)):
        x = torch.randn(B, S, H, hd, device=dev).to(dt)
        xc = torch.view_as_complex(x.float().reshape(B, S, H, hd // 2, 2))
        ref = torch.view_as_real(xc * fc[5:5 + S].view(1, S, 1, hd // 2)).flatten(3)
        _check(f"rope fwd {dt}", N.rope(x, fr, pos_offset=5), ref, tol, tol, results)
        refb = torch.view_as_real(xc * fc[5:5 + S].conj().view(1, S, 1, hd // 2)).flatten(3)
        _check(f"rope bwd {dt}", N.rope(x, fr, pos_offset=5, backward=True), refb, tol, tol, results)
        a = torch.randn(100, 512, device=dev).to(dt); b = torch.randn(100, 512, device=dev).to(dt)
        g = torch.randn(100, 512, device=dev).to(dt)
        ar = a.float().requires_grad_(True); br = b.float().requires_grad_(True)
        yr = F.silu(ar) * br
        yr.backward(g.float())
        _check(f"swiglu fwd {dt}", N.swiglu_fwd(a, b), yr, tol, tol, results)
        da, db = N.swiglu_bwd(g, a, b)
        _check(f"swiglu da {dt}", da, ar.grad, tol * 2, tol, results)
        _check(f"swiglu db {dt}", db, br.grad, tol * 2, tol, results)
        # fused layout: a|b halves of one [M, 2F] buffer
        ab = torch.cat([a, b], dim=1)
        _check(f"swiglu fwd strided {dt}", N.swiglu_fwd(ab[:, :512], ab[:, 512:]), yr, tol, tol, results)
        x1 = torch.randn(1003, device=dev).to(dt); x2 = torch.randn(1003, device=dev).to(dt)
        xr = x1.float().requires_grad_(True)
        F.gelu(xr).backward(x2.float())
        _check(f"gelu {dt}", N.elementwise("gelu", x1), F.gelu(x1.float()), tol, tol, results)
        _check(f"gelu_bwd {dt}", N.elementwise("gelu_bwd", x1, x2), xr.grad, tol, tol, results)
        _check(f"relu_bwd {dt}", N.elementwise("relu_bwd", x1, x2), (x1 > 0).float() * x2.float(), tol, tol, results)
        _check(f"add {dt}", N.elementwise("add", x1, x2), x1.float() + x2.float(), tol, tol, results)
        _check(f"axpy {dt}", N.elementwise("axpy", x1, x2, alpha=0.5), x1.float() + 0.5 * x2.float(), tol, tol, results)
    xf = torch.randn(4099, device=dev)
    _check("cast f32->bf16 scale", N.cast_scale(xf, torch.bfloat16, 0.25), xf * 0.25, 1e-2, 1e-2, results)
    _check("cast bf16->f32", N.cast_scale(xf.bfloat16(), torch.float32, 2.0), xf.bfloat16().float() * 2, 1e-6, 1e-6, results)
    # embedding
    V, D = 1000, 256
    for dt in (torch.bfloat16, torch.float32):
        w = torch.randn(V, D, device=dev).to(dt)
        idx = torch.randint(0, V, (4, 77), device=dev)
        _check(f"embedding fwd {dt}", N.embedding_fwd(idx, w), w[idx], 0, 0, results)
        half = N.embedding_fwd(idx, w[500:], vocab_start=500)
        ref = torch.where((idx >= 500).unsqueeze(-1), w[idx], torch.zeros_like(w[idx]))
        _check(f"embedding fwd vocab-parallel {dt}", half, ref, 0, 0, results)
        dy = torch.randn(4, 77, D, device=dev).to(dt)
        dw = N.embedding_bwd(idx, dy, V)
        ref = torch.zeros(V, D, device=dev).index_add_(0, idx.reshape(-1), dy.reshape(-1, D).float())
        _check(f"embedding bwd {dt}", dw, ref, 1e-3, 1e-3, results)
    # cross entropy
    for dt, tol in ((torch.float32, 1e-4), (torch.bfloat16, 2e-2)):
        for rows, V in ((64, 32000), (33, 1000), (16, 10)):
            lg = (torch.randn(rows, V, device=dev) * 3).to(dt)
            tg = torch.randint(0, V, (rows,), device=dev)
            tg[0] = -100
            lr = lg.float().requires_grad_(True)
            ref = F.cross_entropy(lr, tg, ignore_index=-100, reduction="none")
            loss, lse = N.xent_fwd(lg, tg)
            _check(f"xent fwd {dt} {rows}x{V}", loss, ref, tol, tol, results)
            ref.sum().backward()
            gs = torch.ones(1, device=dev)
            _check(f"xent bwd {dt} {rows}x{V}", N.xent_bwd(lg, tg, lse, gs), lr.grad, tol, tol, results)
    lg = torch.randn(8192, 32000, device=dev); tg = torch.randint(0, 32000, (8192,), device=dev)
    ms = _time(lambda: N.xent_fwd(lg, tg))
    print(f"[TIME] xent_fwd 8192x32000 fp32: {ms:.3f} ms {lg.numel()*4*2/ms/1e6:.0f} GB/s (2 reads)", flush=True)
    del lg
    # the Llama-7B feed-forward shapes of one 4096-token step (w1|w3 output stacked in one [M, 2F] tensor)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    M_, F_ = 4096, 11008
    h13 = torch.randn(M_, 2 * F_, device=dev).bfloat16(); gy = torch.randn(M_, F_, device=dev).bfloat16()
    ms = _time(lambda: N.swiglu_fwd(h13[:, :F_], h13[:, F_:]), flush=flush)
    print(f"[TIME] swiglu_fwd {M_}x{F_} bf16: {ms*1e3:.1f} us  {3*M_*F_*2/ms/1e6:.0f} GB/s", flush=True)
    ms = _time(lambda: N.swiglu_bwd(gy, h13[:, :F_], h13[:, F_:]), flush=flush)
    print(f"[TIME] swiglu_bwd {M_}x{F_} bf16: {ms*1e3:.1f} us  {5*M_*F_*2/ms/1e6:.0f} GB/s", flush=True)
    # optimizers vs torch
    for pdt, gdt in ((torch.float32, torch.float32), (torch.float32, torch.bfloat16)):
        n = 10007
        p0 = torch.randn(n, device=dev); g = torch.randn(n, device=dev).to(gdt)
        pr = p0.clone().requires_grad_(True)
        opt = torch.optim.AdamW([pr], lr=1e-2, betas=(0.9, 0.95), eps=1e-8, weight_decay=0.1)
        p = p0.clone(); m = torch.zeros(n, device=dev); v = torch.zeros(n, device=dev)
        lowp = torch.empty(n, device=dev, dtype=torch.bfloat16)
        for step in (1, 2, 3):
            pr.grad = g.float().clone(); opt.step()
            N.adamw_step(p, g, m, v, lr=1e-2, beta1=0.9, beta2=0.95, eps=1e-8, wd=0.1, step=step, p_lowp=lowp)
        _check(f"adamw p={pdt} g={gdt}", p, pr.detach(), 1e-5, 1e-5, results)
        _check(f"adamw lowp copy g={gdt}", lowp, pr.detach(), 1e-2, 1e-2, results)
        pr = p0.clone().requires_grad_(True)
        opt = torch.optim.SGD([pr], lr=0.1, momentum=0.9, weight_decay=1e-5)
        p = p0.clone(); buf = torch.zeros(n, device=dev)
        for step in (1, 2, 3):
            pr.grad = g.float().clone(); opt.step()
            N.sgd_step(p, g, buf, lr=0.1, momentum=0.9, wd=1e-5, first_step=(step == 1))
        _check(f"sgd p={pdt} g={gdt}", p, pr.detach(), 1e-5, 1e-5, results)
    n = 64 << 20
    p = torch.randn(n, device=dev); g = torch.randn(n, device=dev).bfloat16(); m = torch.zeros(n, device=dev); v = torch.zeros(n, device=dev)
    lowp = torch.empty(n, device=dev, dtype=torch.bfloat16)
    ms = _time(lambda: N.adamw_step(p, g, m, v, lr=1e-3, beta1=0.9, beta2=0.95, eps=1e-8, wd=0.1, step=5, p_lowp=lowp))
    print(f"[TIME] adamw 64Mi params: {ms:.3f} ms {n*(4*3*2+2+2)/ms/1e6:.0f} GB/s", flush=True)


def _attn_ref(q, k, v, causal, q_off=0, k_off=0):
    import torch
    B, Sq, H, hd = q.shape
    Hkv = k.shape[2]
    qf, kf, vf = q.float().transpose(1, 2), k.float().transpose(1, 2), v.float().transpose(1, 2)
    if Hkv != H:
        kf = kf.repeat_interleave(H // Hkv, dim=1); vf = vf.repeat_interleave(H // Hkv, dim=1)
    s = qf @ kf.transpose(-1, -2) * (hd ** -0.5)
    if causal:
        qi = torch.arange(Sq, device=q.device).view(-1, 1) + q_off
        ki = torch.arange(k.shape[1], device=q.device).view(1, -1) + k_off
        s = s.masked_fill(ki > qi, float("-inf"))
    p = torch.softmax(s, dim=-1)
    p = torch.nan_to_num(p, nan=0.0)
    return (p @ vf).transpose(1, 2), torch.logsumexp(s, dim=-1)


def group_attn_fwd(results):
    import torch
    from b200dist.ops import native as N
    _check, _time = _h()
    dev = "cuda"
    torch.manual_seed(3)
    cases = [(1, 128, 128, 1, 1, 128, False), (1, 128, 128, 1, 1, 128, True), (2, 256, 256, 4, 4, 128, True),
             (2, 200, 200, 4, 2, 128, True), (1, 384, 640, 2, 2, 64, False), (2, 320, 320, 4, 4, 64, True),
             (1, 1024, 1024, 8, 8, 128, True)]
    for (B, Sq, Sk, H, Hkv, hd, causal) in cases:
        q = torch.randn(B, Sq, H, hd, device=dev).bfloat16()
        k = torch.randn(B, Sk, Hkv, hd, device=dev).bfloat16()
        v = torch.randn(B, Sk, Hkv, hd, device=dev).bfloat16()
        o, lse = N.flash_attn_fwd(q, k, v, causal=causal)
        torch.cuda.synchronize()
        ro, rl = _attn_ref(q, k, v, causal)
        tag = f"B{B} Sq{Sq} Sk{Sk} H{H}/{Hkv} hd{hd} causal={causal}"
        _check(f"attn fwd O {tag}", o, ro, 2e-2, 2e-2, results)
        _check(f"attn fwd LSE {tag}", lse, rl, 2e-2, 1e-2, results)
    # fused-QKV strided input and chunk offsets (ring-attention interface)
    B, S, H, hd = 2, 256, 4, 128
    qkv = torch.randn(B, S, 3 * H * hd, device=dev).bfloat16()
    q, k, v = [t.view(B, S, H, hd) for t in qkv.split(H * hd, dim=-1)]
    o, lse = N.flash_attn_fwd(q, k, v, causal=True)
    ro, rl = _attn_ref(q, k, v, True)
    _check("attn fwd strided qkv", o, ro, 2e-2, 2e-2, results)
    o2, lse2 = N.flash_attn_fwd(q[:, 128:], k[:, :128], v[:, :128], causal=True, q_offset=128, k_offset=0)
    ro2, rl2 = _attn_ref(q[:, 128:], k[:, :128], v[:, :128], True, 128, 0)
    _check("attn fwd chunk offsets", o2, ro2, 2e-2, 2e-2, results)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    for (B, S, H, hd) in ((2, 4096, 32, 128), (1, 4096, 32, 128), (1, 8192, 32, 128)):
        q = torch.randn(B, S, H, hd, device=dev).bfloat16(); k = torch.randn_like(q); v = torch.randn_like(q)
        ms = _time(lambda: N.flash_attn_fwd(q, k, v, causal=True), flush=flush)
        fl = 4.0 * B * H * S * S * hd / 2
        qt, kt, vt = q.transpose(1, 2), k.transpose(1, 2), v.transpose(1, 2)
        ms_ref = _time(lambda: torch.nn.functional.scaled_dot_product_attention(qt, kt, vt, is_causal=True), flush=flush)
        print(f"[TIME] attn fwd causal B{B} S{S} H{H} hd{hd}: {ms:.3f} ms {fl/ms/1e9:.0f} TFLOP/s | SDPA {ms_ref:.3f} ms {fl/ms_ref/1e9:.0f} TFLOP/s", flush=True)


def group_attn_bwd(results):
    import torch
    from b200dist.ops import native as N
    _check, _time = _h()
    dev = "cuda"
    torch.manual_seed(4)
    cases = [(1, 128, 128, 1, 1, 128, False), (1, 128, 128, 1, 1, 128, True), (2, 256, 256, 4, 4, 128, True),
             (2, 200, 200, 4, 2, 128, True), (1, 384, 640, 2, 2, 64, False), (2, 320, 320, 4, 4, 64, True)]
    for (B, Sq, Sk, H, Hkv, hd, causal) in cases:
        q = torch.randn(B, Sq, H, hd, device=dev).bfloat16()
        k = torch.randn(B, Sk, Hkv, hd, device=dev).bfloat16()
        v = torch.randn(B, Sk, Hkv, hd, device=dev).bfloat16()
        do = torch.randn(B, Sq, H, hd, device=dev).bfloat16()
        qr, kr, vr = [t.float().requires_grad_(True) for t in (q, k, v)]
        ro, _ = _attn_ref(qr, kr, vr, causal)
        ro.backward(do.float())
        o, lse = N.flash_attn_fwd(q, k, v, causal=causal)
        dq, dk, dv = N.flash_attn_bwd(q, k, v, o, do, lse, causal=causal)
        torch.cuda.synchronize()
        tag = f"B{B} Sq{Sq} Sk{Sk} H{H}/{Hkv} hd{hd} causal={causal}"
        _check(f"attn bwd dQ {tag}", dq, qr.grad, 5e-2, 3e-2, results)
        _check(f"attn bwd dK {tag}", dk, kr.grad, 5e-2, 3e-2, results)
        _check(f"attn bwd dV {tag}", dv, vr.grad, 5e-2, 3e-2, results)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    for (B, S, H, hd) in ((2, 4096, 32, 128), (1, 4096, 32, 128)):     # the second one is what a Llama-7B step runs
        q = torch.randn(B, S, H, hd, device=dev).bfloat16(); k = torch.randn_like(q); v = torch.randn_like(q); do = torch.randn_like(q)
        o, lse = N.flash_attn_fwd(q, k, v, causal=True)
        ms = _time(lambda: N.flash_attn_bwd(q, k, v, o, do, lse, causal=True), flush=flush)
        fl = 2.5 * 4.0 * B * H * S * S * hd / 2
        print(f"[TIME] attn bwd causal B{B} S{S} H{H} hd{hd}: {ms:.3f} ms {fl/ms/1e9:.0f} TFLOP/s (5-GEMM flops)", flush=True)


GROUPS = {"norms": group_norms, "elementwise": group_elementwise, "attn_fwd": group_attn_fwd,
          "attn_bwd": group_attn_bwd}


def group_conv(results):
    import torch
    import torch.nn.functional as F
    from b200dist.ops import conv as C
    _check, _time = _h()
    dev = "cuda"
    torch.manual_seed(5)

    def cmp(name, fn_mine, fn_ref, inputs, tol=5e-2, rtol=3e-2):
        mine_in = [t.detach().clone().requires_grad_(t.is_floating_point()) for t in inputs]
        ref_in = [t.detach().float().clone().requires_grad_(t.is_floating_point()) for t in inputs]
        ym, yr = fn_mine(*mine_in), fn_ref(*ref_in)
        _check(f"{name} fwd", ym, yr, tol, rtol, results)
        g = torch.randn_like(yr)
        ym.backward(g.to(ym.dtype)); yr.backward(g)
        for i, (a, b) in enumerate(zip(mine_in, ref_in)):
            if b.grad is not None:
                scale = max(1.0, float(b.grad.abs().max()))
                _check(f"{name} grad{i}", a.grad, b.grad, tol * scale, rtol, results)

    bf = torch.bfloat16
    for (N, Cin, H, W, K, k, s, p) in [(2, 64, 14, 14, 128, 1, 1, 0), (2, 64, 15, 13, 64, 3, 1, 1), (2, 32, 16, 16, 64, 3, 2, 1),
                                        (2, 3, 32, 32, 64, 7, 2, 3), (2, 64, 14, 14, 128, 1, 2, 0), (1, 64, 9, 12, 65, 1, 1, 0)]:
        x = torch.randn(N, Cin, H, W, device=dev).to(bf)
        w = (torch.randn(K, Cin, k, k, device=dev) * (Cin * k * k) ** -0.5).to(bf)
        b = torch.randn(K, device=dev).to(bf)
        cmp(f"conv2d {Cin}->{K} k{k} s{s} {H}x{W}", lambda x, w, b: C.conv2d(x, w, b, s, p),
            lambda x, w, b: F.conv2d(x, w, b, s, p), [x, w, b])
    x = torch.randn(4, 64, 12, 10, device=dev).to(bf)
    g = (1 + 0.1 * torch.randn(64, device=dev)); b = 0.1 * torch.randn(64, device=dev)
    res = torch.randn(4, 64, 12, 10, device=dev).to(bf)
    rm, rv = torch.zeros(64, device=dev), torch.ones(64, device=dev)
    rm2, rv2 = rm.clone(), rv.clone()
    cmp("batchnorm+res+relu", lambda x, g, b, r: C.batch_norm(x, g, b, rm, rv, True, 0.1, 1e-5, residual=r, relu=True),
        lambda x, g, b, r: F.relu(F.batch_norm(x, rm2, rv2, g, b, True, 0.1, 1e-5) + r), [x, g, b, res])
    _check("bn running_mean", rm, rm2, 1e-2, 1e-2, results)
    _check("bn running_var", rv, rv2, 1e-2, 1e-2, results)
    cmp("batchnorm plain", lambda x, g, b: C.batch_norm(x, g, b, None, None, True), lambda x, g, b: F.batch_norm(x, None, None, g, b, True),
        [x, g, b])
    cmp("maxpool 3/2/1", lambda x: C.max_pool2d(x, 3, 2, 1), lambda x: F.max_pool2d(x, 3, 2, 1), [x])
    cmp("maxpool 2", lambda x: C.max_pool2d(x, 2), lambda x: F.max_pool2d(x, 2), [x[:, :, :, :9]])
    cmp("global avgpool", lambda x: C.global_avg_pool(x), lambda x: x.mean((2, 3)), [x])
    wt = (torch.randn(64, 32, 2, 2, device=dev) * 0.1).to(bf); bt = torch.randn(32, device=dev).to(bf)
    cmp("convT 2x2", lambda x, w, b: C.conv_transpose2x2(x, w, b), lambda x, w, b: F.conv_transpose2d(x, w, b, stride=2), [x, wt, bt])
    cmp("bilinear 12x10->25x21", lambda x: C.bilinear_resize(x, (25, 21)),
        lambda x: F.interpolate(x, size=(25, 21), mode="bilinear", align_corners=False), [x])
    tgt = torch.randn_like(x)
    w_l = C.lat_weights(12, dev).view(1, 1, -1, 1)
    cmp("lat mse", lambda p: C.latitude_weighted_mse(p, tgt), lambda p: (w_l * (p - tgt.float()) ** 2).mean(), [x], tol=1e-2)
    cmp("lat mse channels_last", lambda p: C.latitude_weighted_mse(p.contiguous(memory_format=torch.channels_last), tgt),
        lambda p: (w_l * (p - tgt.float()) ** 2).mean(), [x], tol=1e-2)


def group_models(results):
    """Whole-model parity: loss curves of our models vs the reference / torchvision implementations."""
    import sys
    import torch
    import torchvision
    from b200dist import ops
    from b200dist.models import build_resnet
    from b200dist.models.llama2 import ModelArgs, Transformer
    from b200dist.ops import FusedAdamW, FusedSGD
    _check, _time = _h()
    dev = "cuda"
    from pathlib import Path
    root = Path(__file__).resolve().parent.parent
    sys.path.insert(0, str(root))
    from baseline.install_ref import ensure_reference, reference_path, why_unavailable
    if ensure_reference() is None:
        print(f"[SKIP] llama loss curve vs reference: {why_unavailable()}", flush=True)
        ref = None
    else:
        sys.path.insert(0, str(reference_path("fsdp_tp")))
        import llama2_model as ref
    if ref is not None:
        a = dict(dim=256, n_layers=2, n_heads=4, vocab_size=1024, max_seq_len=256)
        torch.manual_seed(0)
        r = ref.Transformer.from_model_args(ref.ModelArgs(**a)).to(dev)
        m = Transformer(ModelArgs(**a)).to(dev)
        m.load_state_dict(r.state_dict())
        m = m.to(torch.bfloat16)
        ro = torch.optim.AdamW(r.parameters(), lr=1e-3, betas=(0.9, 0.95), weight_decay=0.1)
        mo = FusedAdamW(m.parameters(), lr=1e-3, betas=(0.9, 0.95), weight_decay=0.1)
        g = torch.Generator(device="cpu").manual_seed(1)
        lr_, lm_ = [], []
        for step in range(8):
            tok = torch.randint(0, 1024, (4, 129), generator=g).to(dev)
            x, y = tok[:, :-1], tok[:, 1:]
            ro.zero_grad()
            with torch.autocast("cuda", dtype=torch.bfloat16):
                out = r(x)
            l = torch.nn.functional.cross_entropy(out.float().view(-1, 1024), y.reshape(-1)); l.backward(); ro.step()
            mo.zero_grad()
            l2 = ops.cross_entropy(m(x), y); l2.backward(); mo.step()
            lr_.append(float(l)); lm_.append(float(l2))
        print("ref  losses", [round(v, 4) for v in lr_], flush=True)
        print("ours losses", [round(v, 4) for v in lm_], flush=True)
        _check("llama loss curve vs reference (bf16)", torch.tensor(lm_), torch.tensor(lr_), 0.08, 0.01, results)
        assert lm_[-1] < lm_[0]
    # ResNet-18 vs torchvision, SGD
    torch.manual_seed(0)
    tv = torchvision.models.resnet18(num_classes=100).to(dev)
    mine = build_resnet("resnet18", num_classes=100).to(dev)
    mine.load_state_dict(tv.state_dict())
    mine = mine.to(torch.bfloat16)
    to = torch.optim.SGD(tv.parameters(), lr=0.05, momentum=0.9, weight_decay=1e-5)
    mo = FusedSGD(mine.parameters(), lr=0.05, momentum=0.9, weight_decay=1e-5)
    x = torch.rand(16, 3, 64, 64, device=dev); y = torch.randint(0, 100, (16,), device=dev)
    lt, lm = [], []
    for step in range(6):
        to.zero_grad()
        with torch.autocast("cuda", dtype=torch.bfloat16):
            o = tv(x)
        l = torch.nn.functional.cross_entropy(o.float(), y); l.backward(); to.step()
        mo.zero_grad()
        l2 = ops.cross_entropy(mine(x.bfloat16()), y); l2.backward(); mo.step()
        lt.append(float(l)); lm.append(float(l2))
    print("tv   losses", [round(v, 4) for v in lt], flush=True)
    print("ours losses", [round(v, 4) for v in lm], flush=True)
    _check("resnet18 loss curve vs torchvision (bf16)", torch.tensor(lm), torch.tensor(lt), 0.35, 0.05, results)
    assert lm[-1] < lm[0]
    # ResNet-50 throughput, single GPU (no DDP): ours vs torchvision+autocast
    for arch, B in (("resnet50", 128),):
        mine = build_resnet(arch).to(dev).to(torch.bfloat16)
        mo = FusedSGD(mine.parameters(), lr=0.1, momentum=0.9, weight_decay=1e-5)
        x = torch.rand(B, 3, 224, 224, device=dev).bfloat16(); y = torch.zeros(B, dtype=torch.long, device=dev)

        def step_mine():
            mo.zero_grad(); ops.cross_entropy(mine(x), y).backward(); mo.step()
        ms = _time(step_mine, iters=5, warmup=2)
        tv = getattr(torchvision.models, arch)().to(dev).to(memory_format=torch.channels_last)
        to = torch.optim.SGD(tv.parameters(), lr=0.1, momentum=0.9, weight_decay=1e-5)
        xf = x.float().contiguous(memory_format=torch.channels_last)

        def step_tv():
            to.zero_grad()
            with torch.autocast("cuda", dtype=torch.bfloat16):
                o = tv(xf)
            torch.nn.functional.cross_entropy(o.float(), y).backward(); to.step()
        ms_tv = _time(step_tv, iters=5, warmup=2)
        print(f"[TIME] {arch} B{B} train step: ours {ms:.1f} ms ({B/ms*1e3:.0f} img/s) | torchvision+cuDNN bf16 autocast {ms_tv:.1f} ms "
              f"({B/ms_tv*1e3:.0f} img/s)", flush=True)


def group_optim_overlap(results):
    """FusedAdamW(overlap_backward=True) - updates issued from the gradient hooks on a side stream - must produce the
    weights of the plain optimizer (same kernels, same arithmetic per parameter)."""
    import torch
    from b200dist import ops
    from b200dist.models.llama2 import ModelArgs, Transformer
    from b200dist.ops import FusedAdamW
    _check, _time = _h()
    dev = "cuda"
    a = dict(dim=512, n_layers=3, n_heads=4, vocab_size=2048, max_seq_len=256)
    torch.manual_seed(0)
    m0 = Transformer(ModelArgs(**a)).to(dev).to(torch.bfloat16)
    m1 = Transformer(ModelArgs(**a)).to(dev).to(torch.bfloat16)
    m1.load_state_dict(m0.state_dict())
    o0 = FusedAdamW(m0.parameters(), lr=1e-3, betas=(0.9, 0.95), weight_decay=0.1)
    o1 = FusedAdamW(m1.parameters(), lr=1e-3, betas=(0.9, 0.95), weight_decay=0.1, overlap_backward=True)
    g = torch.Generator(device="cpu").manual_seed(3)
    l0, l1 = [], []
    for step in range(6):
        tok = torch.randint(0, 2048, (4, 257), generator=g).to(dev)
        x, y = tok[:, :-1], tok[:, 1:]
        for m, o, ls in ((m0, o0, l0), (m1, o1, l1)):
            o.zero_grad()
            l = ops.cross_entropy(m(x), y); l.backward(); o.step()
            ls.append(float(l))
    torch.cuda.synchronize()
    print("plain   ", [round(v, 4) for v in l0]); print("overlap ", [round(v, 4) for v in l1], flush=True)
    # split-K weight gradients accumulate with fp32 atomics, so two runs agree to rounding, not bit for bit
    _check("overlap_backward losses vs plain", torch.tensor(l1), torch.tensor(l0), 5e-3, 0.0, results)
    assert all(p.grad is None for p in m1.parameters()), "overlap hooks must consume the gradients"
    _check("overlap_backward weights (all)", torch.cat([p.reshape(-1).float() for p in m1.parameters()]),
           torch.cat([p.reshape(-1).float() for p in m0.parameters()]), 4e-3, 0.0, results)


def group_fused_block(results):
    """models/fused_block.py (one autograd node per Llama block: fused q|k|v and w1|w3 GEMMs, in-place strided RoPE,
    residual gradient folded into the RMSNorm backward) against the module-by-module path on the same weights: loss,
    logits and every parameter gradient, for packed weights, unpacked weights (accumulate epilogues) and GQA."""
    import torch
    from b200dist import ops
    from b200dist.models.llama2 import ModelArgs, Transformer
    _check, _time = _h()
    dev = "cuda"
    for name, cfg, pack in (("packed", dict(dim=512, n_layers=2, n_heads=4, vocab_size=1024, max_seq_len=512), True),
                            ("unpacked", dict(dim=512, n_layers=2, n_heads=4, vocab_size=1024, max_seq_len=512), False),
                            ("gqa hd64", dict(dim=512, n_layers=2, n_heads=8, n_kv_heads=2, vocab_size=1024, max_seq_len=512), True)):
        torch.manual_seed(0)
        ref = Transformer(ModelArgs(**cfg)).to(dev).to(torch.bfloat16).fuse_blocks(False)
        fus = Transformer(ModelArgs(**cfg)).to(dev).to(torch.bfloat16)
        fus.load_state_dict(ref.state_dict())
        fus.fuse_blocks(True, pack=pack)
        tok = torch.randint(0, 1024, (2, 385), device=dev)
        x, y = tok[:, :-1], tok[:, 1:]
        lo_r = ref(x); l_r = ops.cross_entropy(lo_r, y); l_r.backward()
        lo_f = fus(x); l_f = ops.cross_entropy(lo_f, y); l_f.backward()
        torch.cuda.synchronize()
        _check(f"fused block [{name}] logits", lo_f, lo_r, 3e-2, 2e-2, results)
        _check(f"fused block [{name}] loss", l_f.detach().reshape(1), l_r.detach().reshape(1), 2e-3, 0.0, results)
        for (n, pf), (_, pr) in zip(fus.named_parameters(), ref.named_parameters()):
            assert pf.grad is not None, n
            g_abs = float(pr.grad.float().abs().max())
            _check(f"fused block [{name}] grad {n}", pf.grad, pr.grad, 0.02 * g_abs + 1e-6, 3e-2, results)
    # timing of one 7B-shape block, forward + backward, both paths
    cfg = dict(dim=4096, n_layers=1, n_heads=32, vocab_size=1024, max_seq_len=4096)
    for fused in (False, True):
        torch.manual_seed(0)
        m = Transformer(ModelArgs(**cfg)).to(dev).to(torch.bfloat16).fuse_blocks(fused, pack=True)
        h = torch.randn(1, 4096, 4096, device=dev, dtype=torch.bfloat16, requires_grad=True)
        blk = m.layers[0]

        def step():
            out = blk(h, m.freqs_cis)
            out.backward(out)
            for p in blk.parameters():
                p.grad = None
        ms = _time(step, iters=10, warmup=3)
        print(f"[TIME] llama-7b block fwd+bwd S4096 {'fused node' if fused else 'module-by-module'}: {ms:.3f} ms", flush=True)


def group_attn_dropout(results):
    """Attention with Philox dropout on the probabilities vs an fp32 PyTorch expression that uses the SAME mask (rebuilt on
    the host from the seed), forward and backward; the pipeline transformer block in train / eval mode."""
    import torch
    from b200dist import ops
    from b200dist.ops import native as N
    from b200dist.ops.functional import philox_keep_mask
    _check, _time = _h()
    dev = "cuda"
    torch.manual_seed(11)
    # 1. the mask itself: uniform scores make every kept probability 1 / cols / (1 - p), every dropped one 0
    for (rows, cols, p, seed, off) in ((64, 128, 0.1, 1234, 1), (37, 516, 0.5, 2 ** 40 + 17, 99)):
        s = torch.zeros(rows, cols, device=dev)
        probs, dropped = N.softmax_dropout_fwd(s, rows, 1.0, False, p, seed, off)
        ref = philox_keep_mask(rows * cols, p, seed, off).view(rows, cols).to(dev)
        _check(f"philox mask {rows}x{cols} p={p}", (dropped > 0).float(), ref.float(), 0.0, 0.0, results)
        _check(f"softmax probs {rows}x{cols}", probs, torch.full_like(s, 1.0 / cols), 1e-3, 1e-2, results)
    # 2. full attention, forward + backward, same mask on both sides
    for (B, S, H, hd, causal, p) in ((2, 64, 8, 32, False, 0.1), (1, 132, 4, 64, True, 0.25), (2, 256, 2, 128, False, 0.0001)):
        q = torch.randn(B, S, H, hd, device=dev).bfloat16(); k = torch.randn_like(q); v = torch.randn_like(q)
        do = torch.randn_like(q)
        seed = 4242 + S
        ql, kl, vl = [t.clone().requires_grad_(True) for t in (q, k, v)]
        import b200dist.ops.functional as Fn
        Fn._dropout_calls = 6                       # the call below uses offset 7
        o = ops.attention_dropout(ql, kl, vl, p, causal=causal, seed=seed)
        o.backward(do)
        keep = philox_keep_mask(B * H * S * S, p, seed, 7).view(B, H, S, S).to(dev)
        qr, kr, vr = [t.float().requires_grad_(True) for t in (q, k, v)]
        sc = (qr.transpose(1, 2) @ kr.transpose(1, 2).transpose(-1, -2)) * hd ** -0.5
        if causal:
            sc = sc.masked_fill(torch.ones(S, S, device=dev, dtype=torch.bool).triu(1), float("-inf"))
        pr = torch.softmax(sc, -1) * keep / (1 - p)
        orf = (pr @ vr.transpose(1, 2)).transpose(1, 2)
        orf.backward(do.float())
        tag = f"B{B} S{S} H{H} hd{hd} causal={causal} p={p}"
        _check(f"attn dropout O {tag}", o, orf, 3e-2, 3e-2, results)
        _check(f"attn dropout dQ {tag}", ql.grad, qr.grad, 6e-2, 4e-2, results)
        _check(f"attn dropout dK {tag}", kl.grad, kr.grad, 6e-2, 4e-2, results)
        _check(f"attn dropout dV {tag}", vl.grad, vr.grad, 6e-2, 4e-2, results)
    # 3. the recipe model: train mode runs (different masks per call), eval mode is deterministic and equals no-dropout
    from b200dist.models.pp_transformer import Block
    blk = Block(256, 8).to(dev).bfloat16()
    x = torch.randn(2, 128, 256, device=dev).bfloat16()
    blk.train()
    y1, y2 = blk(x), blk(x)
    differ = bool((y1 != y2).any().item())
    results.append({"name": "pp block: two training calls draw different masks", "ok": differ, "max_err": 0.0, "bad": int(not differ),
                    "ref_absmax": 0.0})
    print(f"[{'PASS' if differ else 'FAIL'}] pp block: two training calls draw different masks", flush=True)
    blk.eval()
    _check("pp block eval deterministic", blk(x), blk(x), 0.0, 0.0, results)


GROUPS.update({"conv": group_conv, "models": group_models, "optim_overlap": group_optim_overlap,
               "fused_block": group_fused_block, "attn_dropout": group_attn_dropout})
