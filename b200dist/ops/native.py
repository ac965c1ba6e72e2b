"""Thin typed wrappers over the C ABI (one Python function per ``b200_*`` entry point).

No autograd here — see ``b200dist.ops.functional`` for the differentiable ops.  All
functions take CUDA tensors, launch on the current stream and return their outputs.
"""
from __future__ import annotations

import ctypes as C
from typing import Optional

import torch

from . import _lib as L

I, LL, F = C.c_int, C.c_longlong, C.c_float
ACT = {None: 0, "none": 0, "gelu": 1, "relu": 2, "silu": 3}


def _s():
    return L.stream_ptr()


def _rowmajor_2d(t: torch.Tensor) -> torch.Tensor:
    """Return a 2-D view/copy whose last dim is contiguous (row stride arbitrary)."""
    if t.dim() != 2:
        t = t.reshape(-1, t.shape[-1])
    if t.stride(1) != 1 or t.stride(0) < t.shape[1] or (t.data_ptr() & 15) or (t.stride(0) & 7):
        t = t.contiguous()
    return t


def gemm(a: torch.Tensor, b: torch.Tensor, *, a_mn: bool = False, b_mn: bool = False,
         out: Optional[torch.Tensor] = None, out_dtype: torch.dtype = torch.bfloat16,
         bias: Optional[torch.Tensor] = None, residual: Optional[torch.Tensor] = None,
         alpha: float = 1.0, act: Optional[str] = None, accumulate: bool = False, config: int = 0,
         max_ctas: int = 0) -> torch.Tensor:
    """D[M,N] = act(alpha * A @ B^T + bias) + residual (+ D).

    ``a`` is the stored operand: [M,K] (a_mn=False) or [K,M] (a_mn=True); ``b`` is [N,K]
    (b_mn=False) or [K,N] (b_mn=True).  bf16 inputs, fp32 accumulation on tcgen05.
    """
    assert a.dtype == torch.bfloat16 and b.dtype == torch.bfloat16, "gemm operands must be bf16"
    a = _rowmajor_2d(a)
    b = _rowmajor_2d(b)
    M, K = (a.shape[1], a.shape[0]) if a_mn else (a.shape[0], a.shape[1])
    N, K2 = (b.shape[1], b.shape[0]) if b_mn else (b.shape[0], b.shape[1])
    assert K == K2, f"gemm K mismatch: {K} vs {K2}"
    if out is None:
        out = torch.empty((M, N), device=a.device, dtype=out_dtype)
    else:
        assert out.shape[-1] == N and out.stride(-1) == 1
    if bias is not None:
        bias = bias.to(torch.bfloat16).contiguous()
    if residual is not None:
        residual = _rowmajor_2d(residual.to(torch.bfloat16))
    o2 = out if out.dim() == 2 else out.view(-1, N)
    L.call("gemm_bf16", L.ptr(a), L.ptr(b), L.ptr(o2), I(M), I(N), I(K), LL(a.stride(0)), LL(b.stride(0)),
           LL(o2.stride(0)), I(int(a_mn)), I(int(b_mn)), I(1 if out.dtype == torch.float32 else 0),
           L.ptr(bias), L.ptr(residual), LL(residual.stride(0) if residual is not None else 0), F(alpha),
           I(ACT[act]), I(int(accumulate)), I(config), I(max_ctas), _s())
    return out


# ------------------------------------------------------------------ norms
def norm_bwd_partials(rows: int) -> int:
    return int(L.load().b200_norm_bwd_partials(I(rows)))


def rmsnorm_fwd(x, w, eps):
    D = x.shape[-1]
    x2 = x.reshape(-1, D)
    y = torch.empty_like(x2)
    rstd = torch.empty(x2.shape[0], device=x.device, dtype=torch.float32)
    L.call("rmsnorm_fwd", L.ptr(x2), L.ptr(w), L.ptr(y), L.ptr(rstd), I(x2.shape[0]), I(D), F(eps),
           I(L.dtype_code(x.dtype)), _s())
    return y.view(x.shape), rstd


def rmsnorm_bwd(dy, x, w, rstd):
    D = x.shape[-1]
    x2, dy2 = x.reshape(-1, D), dy.reshape(-1, D)
    dx = torch.empty_like(x2)
    dw = torch.empty_like(w)
    scratch = torch.empty((norm_bwd_partials(x2.shape[0]), D), device=x.device, dtype=torch.float32)
    L.call("rmsnorm_bwd", L.ptr(dy2), L.ptr(x2), L.ptr(w), L.ptr(rstd), L.ptr(dx), L.ptr(dw), L.ptr(scratch),
           I(x2.shape[0]), I(D), I(0), I(0), I(L.dtype_code(x.dtype)), _s())
    return dx.view(x.shape), dw


def layernorm_fwd(x, w, b, eps):
    D = x.shape[-1]
    x2 = x.reshape(-1, D)
    y = torch.empty_like(x2)
    mean = torch.empty(x2.shape[0], device=x.device, dtype=torch.float32)
    rstd = torch.empty_like(mean)
    L.call("layernorm_fwd", L.ptr(x2), L.ptr(w), L.ptr(b), L.ptr(y), L.ptr(mean), L.ptr(rstd), I(x2.shape[0]),
           I(D), F(eps), I(L.dtype_code(x.dtype)), _s())
    return y.view(x.shape), mean, rstd


def layernorm_bwd(dy, x, w, mean, rstd, has_bias=True):
    D = x.shape[-1]
    x2, dy2 = x.reshape(-1, D), dy.reshape(-1, D)
    dx = torch.empty_like(x2)
    dw = torch.empty_like(w)
    db = torch.empty_like(w) if has_bias else None
    scratch = torch.empty((2 * norm_bwd_partials(x2.shape[0]), D), device=x.device, dtype=torch.float32)
    L.call("layernorm_bwd", L.ptr(dy2), L.ptr(x2), L.ptr(w), L.ptr(mean), L.ptr(rstd), L.ptr(dx), L.ptr(dw),
           L.ptr(db), L.ptr(scratch), I(x2.shape[0]), I(D), I(0), I(L.dtype_code(x.dtype)), _s())
    return dx.view(x.shape), dw, db


# ------------------------------------------------------------------ elementwise
def rope(x, freqs, pos_offset=0, backward=False, out=None):
    """x: [B, S, H, hd] contiguous; freqs: fp32 [max_pos, hd/2, 2] (cos, sin)."""
    B, S, H, hd = x.shape
    assert x.is_contiguous()
    y = torch.empty_like(x) if out is None else out
    L.call("rope", L.ptr(x), L.ptr(y), L.ptr(freqs), LL(B * S * H), I(S), I(H), I(hd), I(pos_offset),
           I(int(backward)), I(L.dtype_code(x.dtype)), _s())
    return y


def softmax_dropout_fwd(scores, sq: int, scale: float, causal: bool, p_drop: float, seed: int, offset: int):
    """probs, probs_dropped (bf16, same shape) of ``scores`` [..., sq, sk] (bf16 / fp32, contiguous, sk % 4 == 0)."""
    assert scores.is_contiguous() and scores.dtype in (torch.bfloat16, torch.float32)
    cols = scores.shape[-1]
    rows = scores.numel() // cols
    probs = torch.empty(scores.shape, device=scores.device, dtype=torch.bfloat16)
    dropped = torch.empty_like(probs)
    L.call("softmax_dropout_fwd", L.ptr(scores), L.ptr(probs), L.ptr(dropped), LL(rows), I(cols), I(sq), F(scale),
           I(1 if causal else 0), F(p_drop), C.c_ulonglong(seed), C.c_ulonglong(offset), I(L.dtype_code(scores.dtype)), _s())
    return probs, dropped


def softmax_dropout_bwd(probs, d_dropped, scale: float, p_drop: float, seed: int, offset: int):
    assert probs.is_contiguous() and probs.dtype == torch.bfloat16
    d_dropped = d_dropped.to(torch.bfloat16).contiguous()
    cols = probs.shape[-1]
    ds = torch.empty_like(probs)
    L.call("softmax_dropout_bwd", L.ptr(probs), L.ptr(d_dropped), L.ptr(ds), LL(probs.numel() // cols), I(cols), F(scale),
           F(p_drop), C.c_ulonglong(seed), C.c_ulonglong(offset), _s())
    return ds


def rope_strided(x2d, freqs, S: int, n_heads: int, hd: int, pos_offset=0, backward=False, out=None):
    """Rotate the first ``n_heads`` heads of every row of ``x2d`` [tokens, ld] (fused QKV buffer), in place by default."""
    assert x2d.dim() == 2 and x2d.stride(1) == 1
    y = x2d if out is None else out
    L.call("rope_strided", L.ptr(x2d), L.ptr(y), L.ptr(freqs), LL(x2d.shape[0]), I(S), I(n_heads), I(hd), LL(x2d.stride(0)),
           LL(y.stride(0)), I(pos_offset), I(int(backward)), I(L.dtype_code(x2d.dtype)), _s())
    return y


def rmsnorm_bwd_res(dy, x, w, rstd, dres, dw_out=None, accumulate_dw=False):
    """dx = dres + d rmsnorm / dx, dw: the residual branch of a pre-norm block folded into the norm backward."""
    D = x.shape[-1]
    x2, dy2, r2 = x.reshape(-1, D), dy.reshape(-1, D), dres.reshape(-1, D)
    assert x2.is_contiguous() and dy2.is_contiguous() and r2.is_contiguous()
    dx = torch.empty_like(x2)
    dw = torch.empty_like(w) if dw_out is None else dw_out
    scratch = torch.empty((norm_bwd_partials(x2.shape[0]), D), device=x.device, dtype=torch.float32)
    L.call("rmsnorm_bwd_res", L.ptr(dy2), L.ptr(x2), L.ptr(w), L.ptr(rstd), L.ptr(r2), L.ptr(dx), L.ptr(dw), L.ptr(scratch),
           I(x2.shape[0]), I(D), I(int(accumulate_dw)), I(L.dtype_code(x.dtype)), _s())
    return dx.view(x.shape), dw


def swiglu_fwd(a, b):
    F_ = a.shape[-1]
    a2, b2 = a.reshape(-1, F_) if a.is_contiguous() else a, b.reshape(-1, F_) if b.is_contiguous() else b
    assert a2.dim() == 2 and b2.dim() == 2 and a2.stride(1) == 1 and b2.stride(1) == 1
    y = torch.empty((a2.shape[0], F_), device=a.device, dtype=a.dtype)
    L.call("swiglu_fwd", L.ptr(a2), L.ptr(b2), L.ptr(y), LL(a2.shape[0]), I(F_), LL(a2.stride(0)),
           LL(b2.stride(0)), LL(F_), I(L.dtype_code(a.dtype)), _s())
    return y.view(*a.shape[:-1], F_)


def swiglu_bwd(dy, a, b, da=None, db=None):
    F_ = a.shape[-1]
    a2, b2 = (a.reshape(-1, F_) if a.is_contiguous() else a), (b.reshape(-1, F_) if b.is_contiguous() else b)
    dy2 = dy.reshape(-1, F_)
    if not dy2.is_contiguous():
        dy2 = dy2.contiguous()
    if da is None:
        da = torch.empty((a2.shape[0], F_), device=a.device, dtype=a.dtype)
    if db is None:
        db = torch.empty((a2.shape[0], F_), device=a.device, dtype=a.dtype)
    L.call("swiglu_bwd", L.ptr(dy2), L.ptr(a2), L.ptr(b2), L.ptr(da), L.ptr(db), LL(a2.shape[0]), I(F_), LL(F_),
           LL(a2.stride(0)), LL(b2.stride(0)), LL(da.stride(0)), LL(db.stride(0)), I(L.dtype_code(a.dtype)), _s())
    return da, db


EW = {"gelu": 0, "gelu_bwd": 1, "relu": 2, "relu_bwd": 3, "add": 4, "scale": 5, "axpy": 6, "silu": 7, "silu_bwd": 8}


def elementwise(op: str, x, y2=None, alpha=1.0, out=None):
    x = x.contiguous()
    if y2 is not None:
        y2 = y2.contiguous()
    if out is None:
        out = torch.empty_like(x)
    L.call("elementwise", L.ptr(x), L.ptr(y2), L.ptr(out), LL(x.numel()), I(EW[op]), F(alpha),
           I(L.dtype_code(x.dtype)), _s())
    return out


def cast_scale(src, dst_dtype, alpha=1.0, out=None):
    src = src.contiguous()
    if out is None:
        out = torch.empty(src.shape, device=src.device, dtype=dst_dtype)
    L.call("cast_scale", L.ptr(src), I(L.dtype_code(src.dtype)), L.ptr(out), I(L.dtype_code(out.dtype)),
           LL(src.numel()), F(alpha), _s())
    return out


def embedding_fwd(idx, w, vocab_start=0):
    idx = idx.contiguous().to(torch.int64)
    out = torch.empty((*idx.shape, w.shape[1]), device=w.device, dtype=w.dtype)
    L.call("embedding_fwd", L.ptr(idx), L.ptr(w), L.ptr(out), LL(idx.numel()), I(w.shape[1]), LL(vocab_start),
           LL(w.shape[0]), I(L.dtype_code(w.dtype)), _s())
    return out


def embedding_bwd(idx, dy, vocab_local, vocab_start=0, dw=None):
    idx = idx.contiguous().to(torch.int64)
    D = dy.shape[-1]
    dy = dy.contiguous()
    if dw is None:
        dw = torch.zeros((vocab_local, D), device=dy.device, dtype=torch.float32)
    L.call("embedding_bwd", L.ptr(idx), L.ptr(dy), L.ptr(dw), LL(idx.numel()), I(D), LL(vocab_start),
           LL(vocab_local), I(L.dtype_code(dy.dtype)), _s())
    return dw


def xent_fwd(logits, target, ignore_index=-100):
    V = logits.shape[-1]
    l2 = logits.reshape(-1, V)
    assert l2.stride(1) == 1
    t = target.reshape(-1).contiguous().to(torch.int64)
    loss = torch.empty(l2.shape[0], device=logits.device, dtype=torch.float32)
    lse = torch.empty_like(loss)
    L.call("xent_fwd", L.ptr(l2), L.ptr(t), L.ptr(loss), L.ptr(lse), I(l2.shape[0]), I(V), LL(l2.stride(0)),
           LL(ignore_index), I(L.dtype_code(logits.dtype)), _s())
    return loss, lse


def xent_fwd_vp(logits, target, ignore_index, vocab_start):
    """Local-shard forward: (lse_local - logit[target] where the target column is local else 0, lse_local)."""
    V = logits.shape[-1]
    l2 = logits.reshape(-1, V)
    assert l2.stride(1) == 1
    t = target.reshape(-1).contiguous().to(torch.int64)
    loss = torch.empty(l2.shape[0], device=logits.device, dtype=torch.float32)
    lse = torch.empty_like(loss)
    L.call("xent_fwd_vp", L.ptr(l2), L.ptr(t), L.ptr(loss), L.ptr(lse), I(l2.shape[0]), I(V), LL(l2.stride(0)),
           LL(ignore_index), LL(vocab_start), I(L.dtype_code(logits.dtype)), _s())
    return loss, lse


def xent_bwd_vp(logits, target, lse_global, gscale, ignore_index, vocab_start, v_total):
    V = logits.shape[-1]
    l2 = logits.reshape(-1, V)
    t = target.reshape(-1).contiguous().to(torch.int64)
    out = torch.empty_like(l2)
    L.call("xent_bwd_vp", L.ptr(l2), L.ptr(t), L.ptr(lse_global), L.ptr(gscale), L.ptr(out), I(l2.shape[0]), I(V),
           LL(l2.stride(0)), LL(out.stride(0)), LL(ignore_index), LL(vocab_start), LL(v_total),
           I(L.dtype_code(logits.dtype)), _s())
    return out.view(logits.shape)


def xent_bwd(logits, target, lse, gscale, ignore_index=-100, out=None):
    V = logits.shape[-1]
    l2 = logits.reshape(-1, V)
    t = target.reshape(-1).contiguous().to(torch.int64)
    if out is None:
        out = torch.empty_like(l2)
    L.call("xent_bwd", L.ptr(l2), L.ptr(t), L.ptr(lse), L.ptr(gscale), L.ptr(out), I(l2.shape[0]), I(V),
           LL(l2.stride(0)), LL(out.stride(0)), LL(ignore_index), I(L.dtype_code(logits.dtype)), _s())
    return out.view(logits.shape)


def sum_scale(x, scale):
    out = torch.empty(1, device=x.device, dtype=torch.float32)
    L.call("sum_scale", L.ptr(x), L.ptr(out), I(x.numel()), F(scale), _s())
    return out


# ------------------------------------------------------------------ attention
def _bshd_strides(t):
    assert t.dim() == 4 and t.stride(3) == 1, "attention tensors are [B, S, H, hd] with unit hd stride"
    return t.stride(0), t.stride(1), t.stride(2)


def flash_attn_fwd(q, k, v, causal=True, scale=None, q_offset=0, k_offset=0, out=None):
    B, Sq, H, hd = q.shape
    Sk, Hkv = k.shape[1], k.shape[2]
    scale = float(scale if scale is not None else hd ** -0.5)
    o = torch.empty((B, Sq, H, hd), device=q.device, dtype=q.dtype) if out is None else out
    lse = torch.empty((B, H, Sq), device=q.device, dtype=torch.float32)
    qs, ks, vs, os_ = _bshd_strides(q), _bshd_strides(k), _bshd_strides(v), _bshd_strides(o)
    L.call("flash_attn_fwd", L.ptr(q), L.ptr(k), L.ptr(v), L.ptr(o), L.ptr(lse), I(B), I(H), I(Hkv), I(Sq), I(Sk),
           I(hd), *[LL(s) for s in (*qs, *ks, *vs, *os_)], F(scale), I(int(causal)), I(q_offset), I(k_offset), _s())
    return o, lse


def flash_attn_bwd(q, k, v, o, do, lse, causal=True, scale=None, q_offset=0, k_offset=0,
                   dq=None, dk=None, dv=None):
    B, Sq, H, hd = q.shape
    Sk, Hkv = k.shape[1], k.shape[2]
    scale = float(scale if scale is not None else hd ** -0.5)
    if do.stride(3) != 1:
        do = do.contiguous()
    dq = torch.empty_like(q) if dq is None else dq
    dk = torch.empty_like(k) if dk is None else dk
    dv = torch.empty_like(v) if dv is None else dv
    delta = torch.empty((B, H, Sq), device=q.device, dtype=torch.float32)
    st = []
    for t in (q, k, v, o, do, dq, dk, dv):
        st.extend(_bshd_strides(t))
    arr = (LL * 24)(*st)
    L.call("flash_attn_bwd", L.ptr(q), L.ptr(k), L.ptr(v), L.ptr(o), L.ptr(do), L.ptr(lse), L.ptr(delta), L.ptr(dq),
           L.ptr(dk), L.ptr(dv), I(B), I(H), I(Hkv), I(Sq), I(Sk), I(hd), arr, F(scale), I(int(causal)),
           I(q_offset), I(k_offset), _s())
    return dq, dk, dv


# ------------------------------------------------------------------ optimizers
def adamw_step(p, g, m, v, *, lr, beta1, beta2, eps, wd, step, grad_scale=1.0, p_lowp=None):
    bc1 = 1.0 - beta1 ** step
    bc2 = 1.0 - beta2 ** step
    L.call("adamw", L.ptr(p), I(L.dtype_code(p.dtype)), L.ptr(p_lowp), L.ptr(g), I(L.dtype_code(g.dtype)), L.ptr(m),
           L.ptr(v), LL(p.numel()), F(lr), F(beta1), F(beta2), F(eps), F(wd), F(bc1), F(bc2), F(grad_scale), _s())


def adamw_step_dev(p, g, m, v, scalars, *, beta1, beta2, eps, wd, grad_scale=1.0, p_lowp=None):
    """AdamW reading {lr, 1 - beta1^t, sqrt(1 - beta2^t)} from the device tensor ``scalars`` (CUDA-graph friendly)."""
    L.call("adamw_dev", L.ptr(p), I(L.dtype_code(p.dtype)), L.ptr(p_lowp), L.ptr(g), I(L.dtype_code(g.dtype)), L.ptr(m),
           L.ptr(v), LL(p.numel()), F(beta1), F(beta2), F(eps), F(wd), L.ptr(scalars), F(grad_scale), _s())


def sgd_step(p, g, buf, *, lr, momentum, wd, first_step, grad_scale=1.0, p_lowp=None):
    L.call("sgd", L.ptr(p), I(L.dtype_code(p.dtype)), L.ptr(p_lowp), L.ptr(g), I(L.dtype_code(g.dtype)), L.ptr(buf),
           LL(p.numel()), F(lr), F(momentum), F(wd), F(grad_scale), I(int(first_step)), _s())


def conv2d_igemm(x_nhwc, wmat, KH: int, KW: int, pad: int, bias=None, residual=None, act: int = 0, stride: int = 1):
    """Implicit-GEMM convolution (stride 1 or 2): x [N, H, W, C] bf16 contiguous, wmat [Cout, KH * KW * C] bf16 ->
    [N, Ho, Wo, Cout].  C % 64 == 0 and Cout % 8 == 0 (see csrc/gemm_tcgen05.cu, launch_conv_igemm)."""
    Nn, H, W, Cc = x_nhwc.shape
    Cout = wmat.shape[0]
    assert x_nhwc.is_contiguous() and wmat.is_contiguous() and wmat.shape[1] == KH * KW * Cc
    Ho, Wo = (H + 2 * pad - KH) // stride + 1, (W + 2 * pad - KW) // stride + 1
    out = torch.empty((Nn, Ho, Wo, Cout), dtype=torch.bfloat16, device=x_nhwc.device)
    L.call("conv2d_igemm_strided_bf16", L.ptr(x_nhwc), L.ptr(wmat), L.ptr(out), I(Nn), I(H), I(W), I(Cc), I(Cout), I(KH),
           I(KW), I(pad), I(stride), L.ptr(bias), L.ptr(residual), I(act), _s())
    return out


def conv2d_stem(xp_nhwc, wmat, Ho: int, Wo: int, KH: int):
    """Small-channel stride-2 stem convolution: xp [N, Hp, Wp, 8] bf16 (zero-padded channels + border), wmat [Cout, KH * 64]."""
    Nn, Hp, Wp, Cc = xp_nhwc.shape
    assert Cc == 8 and xp_nhwc.is_contiguous() and wmat.is_contiguous() and wmat.shape[1] == KH * 64
    Cout = wmat.shape[0]
    out = torch.empty((Nn, Ho, Wo, Cout), dtype=torch.bfloat16, device=xp_nhwc.device)
    L.call("conv2d_stem_bf16", L.ptr(xp_nhwc), L.ptr(wmat), L.ptr(out), I(Nn), I(Hp), I(Wp), I(Ho), I(Wo), I(Cout), I(KH), _s())
    return out


def conv2d_stem_wgrad(xp_nhwc, dy_nhwc, KH: int):
    """fp32 [Cout, KH * 64] weight gradient of ``conv2d_stem`` in (kh, kw 0..7, c 0..7) column order."""
    Nn, Hp, Wp, Cc = xp_nhwc.shape
    _, Ho, Wo, Cout = dy_nhwc.shape
    assert Cc == 8 and xp_nhwc.is_contiguous() and dy_nhwc.is_contiguous()
    dw = torch.empty((Cout, KH * 64), dtype=torch.float32, device=xp_nhwc.device)
    L.call("conv2d_stem_wgrad", L.ptr(xp_nhwc), L.ptr(dy_nhwc), L.ptr(dw), I(Nn), I(Hp), I(Wp), I(Ho), I(Wo), I(Cout), I(KH), _s())
    return dw


def conv2d_igemm_wgrad(x_nhwc, dy_nhwc, KH: int, KW: int, pad: int, out=None, accumulate: bool = False, stride: int = 1):
    """Implicit-GEMM weight gradient of a stride-1 convolution: x [N, H, W, C], dy [N, Ho, Wo, Cout] bf16 contiguous ->
    fp32 [Cout, KH * KW * C] in (kh, kw, c) column order.  C % 64 == 0, Cout % 8 == 0."""
    Nn, H, W, Cc = x_nhwc.shape
    Cout = dy_nhwc.shape[-1]
    assert x_nhwc.is_contiguous() and dy_nhwc.is_contiguous()
    if out is None:
        out = torch.empty((Cout, KH * KW * Cc), dtype=torch.float32, device=x_nhwc.device)
    L.call("conv2d_igemm_wgrad_strided", L.ptr(x_nhwc), L.ptr(dy_nhwc), L.ptr(out), I(Nn), I(H), I(W), I(Cc), I(Cout), I(KH),
           I(KW), I(pad), I(stride), I(int(accumulate)), _s())
    return out
