"""Differentiable ops.  CUDA tensors run the hand-written sm_100a kernels (and raise if the
native library is missing — no silent fallback on a GPU box); CPU tensors run the plain
PyTorch expression of the same op (unit tests, gloo runs).

Every op mirrors an expression of the reference model code:
  linear        nn.Linear / F.linear        (llama2_model.py:176-183,207,228,267; vit :130-135)
  rms_norm      RMSNorm                     (llama2_model.py:115-142)
  layer_norm    nn.LayerNorm                (tensor_parallel_vit.py:141-143)
  rope          apply_rotary_emb            (llama2_model.py:73-100)
  swiglu        silu(w1 x) * w3 x           (llama2_model.py:266-267)
  attention     F.scaled_dot_product_attention (llama2_model.py:218-223) / explicit softmax (vit :120-122)
  embedding     nn.Embedding                (llama2_model.py:374,440)
  cross_entropy nn.CrossEntropyLoss         (main.py:308; 03_pipeline_training.py:207)
"""
from __future__ import annotations

import math
from typing import Optional

import torch
import torch.nn.functional as F
from torch.autograd import Function

from . import native as N

_BF16 = torch.bfloat16


def _bf16(t: torch.Tensor) -> torch.Tensor:
    return t if t.dtype == _BF16 else t.to(_BF16)


# --------------------------------------------------------------------------- linear
class _LinearFn(Function):
    @staticmethod
    def forward(ctx, x, w, bias, residual):
        K = x.shape[-1]
        xb = _bf16(x.reshape(-1, K))
        wb = _bf16(w)
        out_dtype = torch.float32 if x.dtype == torch.float32 else _BF16
        y = N.gemm(xb, wb, bias=bias, residual=residual.reshape(-1, w.shape[0]) if residual is not None else None,
                   out_dtype=out_dtype)
        ctx.save_for_backward(xb, wb)
        # a parallel wrapper (FSDP) may want the weight gradient written straight into its communication buffer
        ctx.grad_sink = getattr(w, "_b200_grad_sink", None)
        ctx.x_shape, ctx.x_dtype, ctx.w_dtype = x.shape, x.dtype, w.dtype
        ctx.has_bias, ctx.has_res = bias is not None, residual is not None
        ctx.bias_dtype = bias.dtype if bias is not None else None
        return y.view(*x.shape[:-1], w.shape[0])

    @staticmethod
    def backward(ctx, dy):
        xb, wb = ctx.saved_tensors
        Nf = wb.shape[0]
        dyb = _bf16(dy.reshape(-1, Nf))
        dx = dw = db = dres = None
        if ctx.needs_input_grad[0]:
            dx = N.gemm(dyb, wb, b_mn=True,
                        out_dtype=torch.float32 if ctx.x_dtype == torch.float32 else _BF16).view(ctx.x_shape)
        if ctx.needs_input_grad[1]:
            if ctx.grad_sink is not None and ctx.grad_sink(dyb, xb):
                dw = None  # wgrad GEMM wrote into the sink (no autograd accumulation, no extra copy)
            else:
                dw = N.gemm(dyb, xb, a_mn=True, b_mn=True,
                            out_dtype=torch.float32 if ctx.w_dtype == torch.float32 else _BF16)
        if ctx.has_bias and ctx.needs_input_grad[2]:
            db = dy.reshape(-1, Nf).sum(0, dtype=torch.float32).to(ctx.bias_dtype)
        if ctx.has_res and ctx.needs_input_grad[3]:
            dres = dy
        return dx, dw, db, dres


def linear(x, w, bias=None, residual=None):
    """y = x @ w^T (+ bias) (+ residual); tcgen05 GEMM with bf16 operands, fp32 accumulation."""
    if not x.is_cuda:
        y = F.linear(x, w, bias)
        return y + residual if residual is not None else y
    return _LinearFn.apply(x, w, bias, residual)


# --------------------------------------------------------------------------- norms
class _RMSNormFn(Function):
    @staticmethod
    def forward(ctx, x, w, eps):
        xc = x.contiguous()
        wc = w if w.dtype == x.dtype else w.to(x.dtype)
        y, rstd = N.rmsnorm_fwd(xc, wc, eps)
        ctx.save_for_backward(xc, wc, rstd)
        ctx.w_dtype = w.dtype
        return y

    @staticmethod
    def backward(ctx, dy):
        x, w, rstd = ctx.saved_tensors
        dx, dw = N.rmsnorm_bwd(dy.contiguous().to(x.dtype), x, w, rstd)
        return dx, dw.to(ctx.w_dtype), None


def rms_norm(x, weight, eps: float = 1e-6):
    if not x.is_cuda:
        xf = x.float()
        out = (xf * torch.rsqrt(xf.pow(2).mean(-1, keepdim=True) + eps)).type_as(x)
        return out * weight
    return _RMSNormFn.apply(x, weight, eps)


class _LayerNormFn(Function):
    @staticmethod
    def forward(ctx, x, w, b, eps):
        xc = x.contiguous()
        wc = w.to(x.dtype)
        bc = b.to(x.dtype) if b is not None else None
        y, mean, rstd = N.layernorm_fwd(xc, wc, bc, eps)
        ctx.save_for_backward(xc, wc, mean, rstd)
        ctx.has_bias, ctx.w_dtype = b is not None, w.dtype
        return y

    @staticmethod
    def backward(ctx, dy):
        x, w, mean, rstd = ctx.saved_tensors
        dx, dw, db = N.layernorm_bwd(dy.contiguous().to(x.dtype), x, w, mean, rstd, ctx.has_bias)
        return dx, dw.to(ctx.w_dtype), (db.to(ctx.w_dtype) if db is not None else None), None


def layer_norm(x, weight, bias=None, eps: float = 1e-5):
    if not x.is_cuda or x.shape[-1] % 8 or x.shape[-1] > 8192:
        return F.layer_norm(x, (x.shape[-1],), weight, bias, eps)
    return _LayerNormFn.apply(x, weight, bias, eps)


# --------------------------------------------------------------------------- rope
def precompute_freqs_cis(dim: int, end: int, theta: float = 10000.0) -> torch.Tensor:
    """complex64 [end, dim/2] (reference llama2_model.py:30-50)."""
    freqs = 1.0 / (theta ** (torch.arange(0, dim, 2)[: (dim // 2)].float() / dim))
    t = torch.arange(end, device=freqs.device, dtype=torch.float32)
    ang = torch.outer(t, freqs).float()
    # fp32 regardless of torch.get_default_dtype() (polar rejects bf16)
    return torch.polar(torch.ones(end, dim // 2, dtype=torch.float32, device=ang.device), ang)


class _RopeFn(Function):
    @staticmethod
    def forward(ctx, x, freqs_real, pos_offset):
        ctx.save_for_backward(freqs_real)
        ctx.pos_offset = pos_offset
        return N.rope(x.contiguous(), freqs_real, pos_offset)

    @staticmethod
    def backward(ctx, dy):
        (fr,) = ctx.saved_tensors
        return N.rope(dy.contiguous(), fr, ctx.pos_offset, backward=True), None, None


def rope(x, freqs_cis, pos_offset: int = 0):
    """Rotate adjacent pairs of the last dim of x [B, S, H, hd] by freqs_cis[pos_offset + s]."""
    S = x.shape[1]
    if not x.is_cuda or x.shape[-1] % 8:
        xc = torch.view_as_complex(x.float().reshape(*x.shape[:-1], -1, 2))
        fc = freqs_cis[pos_offset:pos_offset + S].view(1, S, 1, -1)
        return torch.view_as_real(xc * fc).flatten(3).type_as(x)
    fr = torch.view_as_real(freqs_cis) if freqs_cis.is_complex() else freqs_cis
    return _RopeFn.apply(x, fr.contiguous(), pos_offset)


# --------------------------------------------------------------------------- swiglu / activations
class _SwiGLUFn(Function):
    @staticmethod
    def forward(ctx, a, b):
        ctx.save_for_backward(a, b)
        return N.swiglu_fwd(a, b)

    @staticmethod
    def backward(ctx, dy):
        a, b = ctx.saved_tensors
        da, db = N.swiglu_bwd(dy, a, b)
        return da.view(a.shape), db.view(b.shape)


def swiglu(a, b):
    """silu(a) * b."""
    if not a.is_cuda or a.shape[-1] % 8:
        return F.silu(a) * b
    return _SwiGLUFn.apply(a, b)


class _UnaryFn(Function):
    @staticmethod
    def forward(ctx, x, op):
        ctx.save_for_backward(x)
        ctx.op = op
        return N.elementwise(op, x)

    @staticmethod
    def backward(ctx, dy):
        (x,) = ctx.saved_tensors
        return N.elementwise(ctx.op + "_bwd", x, dy.to(x.dtype)), None


def gelu(x):
    return _UnaryFn.apply(x, "gelu") if x.is_cuda else F.gelu(x)


def relu(x):
    return _UnaryFn.apply(x, "relu") if x.is_cuda else F.relu(x)


def silu(x):
    return _UnaryFn.apply(x, "silu") if x.is_cuda else F.silu(x)


# --------------------------------------------------------------------------- embedding
class _EmbeddingFn(Function):
    @staticmethod
    def forward(ctx, idx, w, vocab_start):
        ctx.save_for_backward(idx)
        ctx.vocab_local, ctx.vocab_start, ctx.w_dtype = w.shape[0], vocab_start, w.dtype
        return N.embedding_fwd(idx, w, vocab_start)

    @staticmethod
    def backward(ctx, dy):
        (idx,) = ctx.saved_tensors
        dw = N.embedding_bwd(idx, dy, ctx.vocab_local, ctx.vocab_start)
        return None, dw.to(ctx.w_dtype), None


def embedding(idx, weight, vocab_start: int = 0):
    """weight[idx - vocab_start]; rows outside this shard's vocab range give zeros."""
    if not weight.is_cuda or weight.shape[1] % 8:
        local = idx - vocab_start
        ok = (local >= 0) & (local < weight.shape[0])
        out = F.embedding(local.clamp(0, weight.shape[0] - 1), weight)
        return out if vocab_start == 0 and bool(ok.all()) else out * ok.unsqueeze(-1).to(out.dtype)
    return _EmbeddingFn.apply(idx, weight, vocab_start)


# --------------------------------------------------------------------------- cross entropy
class _XentFn(Function):
    @staticmethod
    def forward(ctx, logits, target, ignore_index):
        lg = logits if logits.stride(-1) == 1 else logits.contiguous()
        loss_rows, lse = N.xent_fwd(lg, target, ignore_index)
        n_valid = (target.reshape(-1) != ignore_index).sum().clamp(min=1).to(torch.float32)
        ctx.save_for_backward(lg, target, lse, n_valid)
        ctx.ignore_index = ignore_index
        return loss_rows.sum() / n_valid

    @staticmethod
    def backward(ctx, g):
        lg, target, lse, n_valid = ctx.saved_tensors
        gscale = (g.to(torch.float32) / n_valid).reshape(1).contiguous()
        return N.xent_bwd(lg, target, lse, gscale, ctx.ignore_index), None, None


def cross_entropy(logits, target, ignore_index: int = -100):
    """Mean softmax cross-entropy over non-ignored rows; logits [..., V] bf16/fp32."""
    if not logits.is_cuda:
        return F.cross_entropy(logits.reshape(-1, logits.shape[-1]).float(), target.reshape(-1),
                               ignore_index=ignore_index)
    return _XentFn.apply(logits, target, ignore_index)


# --------------------------------------------------------------------------- attention
class _AttnFn(Function):
    @staticmethod
    def forward(ctx, q, k, v, causal, scale, q_offset, k_offset):
        o, lse = N.flash_attn_fwd(q, k, v, causal, scale, q_offset, k_offset)
        ctx.save_for_backward(q, k, v, o, lse)
        ctx.args = (causal, scale, q_offset, k_offset)
        return o, lse

    @staticmethod
    def backward(ctx, do, _dlse):
        q, k, v, o, lse = ctx.saved_tensors
        causal, scale, qo, ko = ctx.args
        dq, dk, dv = N.flash_attn_bwd(q, k, v, o, do, lse, causal, scale, qo, ko)
        return dq, dk, dv, None, None, None, None


def _attention_reference(q, k, v, causal, scale, q_offset=0, k_offset=0):
    B, Sq, H, hd = q.shape
    Hkv = k.shape[2]
    qf, kf, vf = q.transpose(1, 2).float(), k.transpose(1, 2).float(), v.transpose(1, 2).float()
    if Hkv != H:
        kf = kf.repeat_interleave(H // Hkv, dim=1)
        vf = vf.repeat_interleave(H // Hkv, dim=1)
    s = (qf @ kf.transpose(-1, -2)) * scale
    if causal:
        qi = torch.arange(Sq, device=q.device).view(-1, 1) + q_offset
        ki = torch.arange(k.shape[1], device=q.device).view(1, -1) + k_offset
        s = s.masked_fill(ki > qi, float("-inf"))
    lse = torch.logsumexp(s, dim=-1)
    p = torch.nan_to_num(torch.softmax(s, dim=-1), nan=0.0)
    return (p @ vf).transpose(1, 2).to(q.dtype), lse


def philox_keep_mask(numel: int, p_drop: float, seed: int, offset: int = 0) -> torch.Tensor:
    """Host restatement of the dropout mask of ``csrc/softmax_dropout.cu`` (Philox4x32-10, key = seed, counter =
    (element index / 4, offset), element kept when its 32-bit draw >= p * 2^32): bool [numel] on the CPU.  The oracle of
    the kernel tests; not used on any training path."""
    import numpy as np

    n4 = (numel + 3) // 4
    g = np.arange(n4, dtype=np.uint64)
    c = [g & 0xFFFFFFFF, g >> np.uint64(32), np.full(n4, offset & 0xFFFFFFFF, np.uint64), np.full(n4, (offset >> 32) & 0xFFFFFFFF, np.uint64)]
    k0, k1 = seed & 0xFFFFFFFF, (seed >> 32) & 0xFFFFFFFF
    M0, M1, W0, W1, mask = np.uint64(0xD2511F53), np.uint64(0xCD9E8D57), 0x9E3779B9, 0xBB67AE85, np.uint64(0xFFFFFFFF)
    for _ in range(10):
        p0, p1 = M0 * c[0], M1 * c[2]
        c = [(p1 >> np.uint64(32)) ^ c[1] ^ np.uint64(k0), p1 & mask, (p0 >> np.uint64(32)) ^ c[3] ^ np.uint64(k1), p0 & mask]
        k0, k1 = (k0 + W0) & 0xFFFFFFFF, (k1 + W1) & 0xFFFFFFFF
    draws = np.stack(c, axis=1).reshape(-1)[:numel]
    thresh = min(int(p_drop * 4294967296.0), 0xFFFFFFFF)
    return torch.from_numpy(draws >= np.uint64(thresh))


class _AttnDropoutFn(Function):
    """Explicit attention with dropout on the probabilities: S = Q K^T (batched GEMM), softmax + Philox dropout (one
    kernel, mask regenerated in backward), O = Pd V.  For small-head recipe models; the flash kernels have no dropout."""

    @staticmethod
    def forward(ctx, q, k, v, causal, scale, p_drop, seed, offset):
        # [B, S, H, hd] -> [B, H, S, hd]
        qh, kh, vh = q.transpose(1, 2), k.transpose(1, 2), v.transpose(1, 2)
        s = torch.matmul(qh, kh.transpose(-1, -2)).contiguous()
        probs, dropped = N.softmax_dropout_fwd(s, q.shape[1], scale, causal, p_drop, seed, offset)
        o = torch.matmul(dropped, vh)
        ctx.save_for_backward(qh, kh, vh, probs, dropped)
        ctx.args = (scale, p_drop, seed, offset)
        return o.transpose(1, 2)

    @staticmethod
    def backward(ctx, do):
        qh, kh, vh, probs, dropped = ctx.saved_tensors
        scale, p_drop, seed, offset = ctx.args
        doh = do.transpose(1, 2).to(torch.bfloat16)
        dv = torch.matmul(dropped.transpose(-1, -2), doh)
        d_dropped = torch.matmul(doh, vh.transpose(-1, -2)).contiguous()
        ds = N.softmax_dropout_bwd(probs, d_dropped, scale, p_drop, seed, offset)   # regenerates the mask from the seed
        dq = torch.matmul(ds, kh)
        dk = torch.matmul(ds.transpose(-1, -2), qh)
        return dq.transpose(1, 2), dk.transpose(1, 2), dv.transpose(1, 2), None, None, None, None, None


_dropout_calls = 0


def attention_dropout(q, k, v, p_drop: float, causal: bool = False, scale: Optional[float] = None, training: bool = True,
                      seed: Optional[int] = None):
    """Attention with dropout on the softmax probabilities (``nn.MultiheadAttention(dropout=p)`` semantics) on
    [B, S, H, hd] tensors, H == Hkv.  Without dropout (eval, p == 0) this is :func:`attention`.  The seed comes from
    torch's host generator (so ``torch.manual_seed`` makes runs repeatable and checkpointed recomputation sees the same
    mask); with an explicit ``seed`` every call advances a Philox offset instead."""
    global _dropout_calls
    if not training or p_drop <= 0.0:
        return attention(q, k, v, causal=causal, scale=scale)
    hd = q.shape[-1]
    scale = float(scale if scale is not None else hd ** -0.5)
    if not q.is_cuda or k.shape[1] % 4 or k.shape[2] != q.shape[2]:
        qf, kf, vf = q.transpose(1, 2).float(), k.transpose(1, 2).float(), v.transpose(1, 2).float()
        s_ = (qf @ kf.transpose(-1, -2)) * scale
        if causal:
            Sq, Sk = q.shape[1], k.shape[1]
            keep = torch.arange(Sk, device=q.device).view(1, -1) <= torch.arange(Sq, device=q.device).view(-1, 1) + (Sk - Sq)
            s_ = s_.masked_fill(~keep, float("-inf"))
        pr = F.dropout(torch.softmax(s_, dim=-1), p_drop, training=True)
        return (pr @ vf).transpose(1, 2).to(q.dtype)
    if seed is None:
        # a fresh seed per call from torch's HOST generator (no device sync).  Activation checkpointing restores that
        # generator before it recomputes a block, so the recomputation draws the same seed and rebuilds the same mask.
        seed, offset = int(torch.randint(0, 2 ** 62, (1,)).item()), 0
    else:
        _dropout_calls += 1          # explicit seed: successive calls (layers) get successive Philox offsets
        offset = _dropout_calls
    in_dtype = q.dtype
    o = _AttnDropoutFn.apply(_bf16(q), _bf16(k), _bf16(v), causal, scale, float(p_drop), seed, offset)
    return o.to(in_dtype)


def attention(q, k, v, causal: bool = True, scale: Optional[float] = None, q_offset: int = 0,
              k_offset: int = 0, return_lse: bool = False):
    """softmax(q k^T * scale [+ causal mask]) v on [B, S, H, hd] tensors (GQA when k/v have fewer heads).

    CUDA: tcgen05 flash attention.  Head dims other than 64/128 are zero-padded up (extra
    dims add nothing to q.k and give zero output columns); fp32 inputs compute in bf16.
    """
    hd = q.shape[-1]
    scale = float(scale if scale is not None else hd ** -0.5)
    if not q.is_cuda:
        o, lse = _attention_reference(q, k, v, causal, scale, q_offset, k_offset)
        return (o, lse) if return_lse else o
    if hd > 128:
        raise NotImplementedError("attention: head_dim > 128 is not supported by the sm_100a kernel")
    in_dtype = q.dtype
    pad = (64 if hd <= 64 else 128) - hd
    qb, kb, vb = _bf16(q), _bf16(k), _bf16(v)
    if pad:
        qb, kb, vb = F.pad(qb, (0, pad)), F.pad(kb, (0, pad)), F.pad(vb, (0, pad))
    fix = lambda t: t if (t.stride(-1) == 1 and not any(s % 8 for s in t.stride()[:-1]) and not (t.data_ptr() & 15)) else t.contiguous()
    o, lse = _AttnFn.apply(fix(qb), fix(kb), fix(vb), causal, scale, q_offset, k_offset)
    if pad:
        o = o[..., :hd]
    o = o.to(in_dtype)
    return (o, lse) if return_lse else o
