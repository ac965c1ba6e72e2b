"""One autograd node per Llama block: hand-scheduled forward and backward on the native kernels.

The module-by-module path (``TransformerBlock.forward``) lets autograd stitch the block together: three separate
projections of the normed input, three input gradients that autograd sums with elementwise adds, two more adds for the
residual branches.  This node runs the same arithmetic (reference block: /root/reference/fsdp_tp/llama2_model.py:275-345)
with the structure a B200 wants:

* **q | k | v as ONE GEMM** when the three weights sit next to each other in memory (an FSDP unit's flat buffer, or a
  model packed with ``Transformer.fuse_blocks()``): N = (Hq + 2 Hkv) * hd instead of 3 launches - 768 instead of
  3 x 256 output tiles of 256 x 256 for Llama-7B, i.e. 10.4 instead of 3 x 3.46 waves over the 74 CTA pairs.  RoPE
  rotates q and k in place inside the projection output in one launch and attention reads q, k, v straight out of it.
  Backward: ``dn1 = [dq|dk|dv] @ [Wq;Wk;Wv]`` is one GEMM with K = 3 * dim (the three partial products accumulate in TMEM
  instead of through memory) and the weight gradient is one [3 * dim, dim] GEMM.
* **w1 | w3 as ONE GEMM** under the same condition; SwiGLU reads both halves of the output in place and its backward
  writes ``[da|db]`` into one buffer that feeds a single K = 2 * ffn input-gradient GEMM.
* **No elementwise adds**: residual adds ride the wo / w2 GEMM epilogues (as before) and, in backward, the residual
  gradient is added inside the RMSNorm backward kernel (``b200_rmsnorm_bwd_res``).
* Weight gradients go to the parallel wrapper's gradient sink when it has one (FSDP writes them straight into its
  communication buffer), for the fused groups through a group sink that covers the adjacent parameters at once.

Where weights are not adjacent the node falls back to per-weight GEMMs with accumulate epilogues - still without adds.
CPU tensors and tensor-parallel blocks keep the module-by-module path.
"""
from __future__ import annotations

from typing import List, Optional, Sequence

import torch
from torch.autograd import Function

from ..ops import native as N

_BF16 = torch.bfloat16


def adjacent_rows(ws: Sequence[torch.Tensor]) -> Optional[torch.Tensor]:
    """If the 2-D row-major weights ``ws`` (same in_features) lie back to back in one allocation, return the stacked
    [sum rows, in_features] view over them, else None."""
    k = ws[0].shape[1]
    ptr = ws[0].data_ptr()
    rows = 0
    for w in ws:
        if w.dim() != 2 or w.shape[1] != k or not w.is_contiguous() or w.dtype != ws[0].dtype or w.data_ptr() != ptr:
            return None
        ptr += w.numel() * w.element_size()
        rows += w.shape[0]
    try:
        return torch.as_strided(ws[0], (rows, k), (k, 1))
    except RuntimeError:   # the views do not share one storage large enough (adjacent by coincidence)
        return None


def _wgrad(dy2d: torch.Tensor, x2d: torch.Tensor, params: Sequence[torch.Tensor]) -> List[Optional[torch.Tensor]]:
    """Weight gradients of the row-stacked ``params`` given dy2d [M, sum rows] and x2d [M, K].  Tries the wrapper's
    group sink (all parameters at once), then per-parameter sinks, else returns the gradients (views of one buffer)."""
    if len(params) > 1:
        gs = getattr(params[0], "_b200_group_sink", None)
        if gs is not None and gs(dy2d, x2d, params):
            return [None] * len(params)
    sinks = [getattr(p, "_b200_grad_sink", None) for p in params]
    if len(params) > 1 and all(s is None for s in sinks):
        dw = N.gemm(dy2d, x2d, a_mn=True, b_mn=True, out_dtype=_BF16)   # [sum rows, K] in one launch
        out, r = [], 0
        for p in params:
            out.append(dw[r:r + p.shape[0]])
            r += p.shape[0]
        return out
    out, r = [], 0
    for p, s in zip(params, sinks):
        d = dy2d[:, r:r + p.shape[0]]
        r += p.shape[0]
        if s is not None and s(d, x2d):
            out.append(None)
        else:
            out.append(N.gemm(d, x2d, a_mn=True, b_mn=True, out_dtype=_BF16))
    return out


class LlamaBlockFn(Function):
    """y = block(x).  Inputs: x [B, S, D] bf16, the real view of the RoPE table, and the nine block parameters."""

    @staticmethod
    def forward(ctx, x, freqs_real, pos_offset, eps, n_heads, n_kv_heads, w_an, wq, wk, wv, wo, w_fn, w1, w2, w3):
        B, S, D = x.shape
        hd = wq.shape[0] // n_heads
        Nq, Nkv = n_heads * hd, n_kv_heads * hd
        Ffn = w1.shape[0]
        M = B * S
        x2 = x.reshape(M, D)
        if not x2.is_contiguous():
            x2 = x2.contiguous()
        wan = w_an if w_an.dtype == x.dtype else w_an.to(x.dtype)
        wfn = w_fn if w_fn.dtype == x.dtype else w_fn.to(x.dtype)
        # ---- attention half ----
        n1, rstd1 = N.rmsnorm_fwd(x2, wan, eps)
        qkv = torch.empty((M, Nq + 2 * Nkv), device=x.device, dtype=_BF16)
        wqkv = adjacent_rows((wq, wk, wv))
        if wqkv is not None:
            N.gemm(n1, wqkv, out=qkv)
        else:
            N.gemm(n1, wq, out=qkv[:, :Nq])
            N.gemm(n1, wk, out=qkv[:, Nq:Nq + Nkv])
            N.gemm(n1, wv, out=qkv[:, Nq + Nkv:])
        N.rope_strided(qkv, freqs_real, S, n_heads + n_kv_heads, hd, pos_offset)        # q and k, in place
        q4 = qkv[:, :Nq].view(B, S, n_heads, hd)
        k4 = qkv[:, Nq:Nq + Nkv].view(B, S, n_kv_heads, hd)
        v4 = qkv[:, Nq + Nkv:].view(B, S, n_kv_heads, hd)
        scale = float(hd) ** -0.5
        o4, lse = N.flash_attn_fwd(q4, k4, v4, True, scale, 0, 0)
        o2 = o4.view(M, Nq)
        h = N.gemm(o2, wo, residual=x2)                                                  # x + attn(x)
        # ---- feed-forward half ----
        n2, rstd2 = N.rmsnorm_fwd(h, wfn, eps)
        ab = torch.empty((M, 2 * Ffn), device=x.device, dtype=_BF16)
        w13 = adjacent_rows((w1, w3))
        if w13 is not None:
            N.gemm(n2, w13, out=ab)
        else:
            N.gemm(n2, w1, out=ab[:, :Ffn])
            N.gemm(n2, w3, out=ab[:, Ffn:])
        s = N.swiglu_fwd(ab[:, :Ffn], ab[:, Ffn:])
        y = N.gemm(s, w2, residual=h)
        ctx.save_for_backward(x2, rstd1, n1, qkv, o4, lse, h, rstd2, n2, ab, s, freqs_real,
                              wan, wq, wk, wv, wo, wfn, w1, w2, w3)
        ctx.meta = (B, S, D, hd, n_heads, n_kv_heads, Ffn, pos_offset, scale, w_an.dtype, w_fn.dtype)
        ctx.params = (w_an, wq, wk, wv, wo, w_fn, w1, w2, w3)   # the Parameter objects (sinks hang off them)
        return y.view(B, S, D)

    @staticmethod
    def backward(ctx, dy):
        (x2, rstd1, n1, qkv, o4, lse, h, rstd2, n2, ab, s, freqs_real,
         wan, wq, wk, wv, wo, wfn, w1, w2, w3) = ctx.saved_tensors
        B, S, D, hd, n_heads, n_kv, Ffn, pos_offset, scale, an_dtype, fn_dtype = ctx.meta
        p_an, p_wq, p_wk, p_wv, p_wo, p_fn, p_w1, p_w2, p_w3 = ctx.params
        M = B * S
        Nq, Nkv = n_heads * hd, n_kv * hd
        dy2 = dy.reshape(M, D)
        if dy2.dtype != _BF16 or not dy2.is_contiguous():
            dy2 = dy2.to(_BF16).contiguous()
        # ---- feed-forward half ----
        ds = N.gemm(dy2, w2, b_mn=True)                                                  # [M, F]
        (dw2,) = _wgrad(dy2, s, (p_w2,))
        dab = torch.empty_like(ab)
        N.swiglu_bwd(ds, ab[:, :Ffn], ab[:, Ffn:], da=dab[:, :Ffn], db=dab[:, Ffn:])
        del ds
        w13 = adjacent_rows((w1, w3))
        if w13 is not None:
            dn2 = N.gemm(dab, w13, b_mn=True)                                            # K = 2 * ffn, one accumulation
        else:
            dn2 = N.gemm(dab[:, :Ffn], w1, b_mn=True)
            N.gemm(dab[:, Ffn:], w3, b_mn=True, out=dn2, accumulate=True)
        dw1, dw3 = _wgrad(dab, n2, (p_w1, p_w3))
        del dab
        dh, dw_fn = N.rmsnorm_bwd_res(dn2, h, wfn, rstd2, dy2)                           # dh = dy + d ffn_norm
        del dn2
        # ---- attention half ----
        do2 = N.gemm(dh, wo, b_mn=True)                                                  # [M, Nq]
        o2 = o4.view(M, Nq)
        (dwo,) = _wgrad(dh, o2, (p_wo,))
        q4 = qkv[:, :Nq].view(B, S, n_heads, hd)
        k4 = qkv[:, Nq:Nq + Nkv].view(B, S, n_kv, hd)
        v4 = qkv[:, Nq + Nkv:].view(B, S, n_kv, hd)
        dqkv = torch.empty_like(qkv)
        N.flash_attn_bwd(q4, k4, v4, o4, do2.view(B, S, n_heads, hd), lse, True, scale, 0, 0,
                         dq=dqkv[:, :Nq].view(B, S, n_heads, hd), dk=dqkv[:, Nq:Nq + Nkv].view(B, S, n_kv, hd),
                         dv=dqkv[:, Nq + Nkv:].view(B, S, n_kv, hd))
        del do2
        N.rope_strided(dqkv, freqs_real, S, n_heads + n_kv, hd, pos_offset, backward=True)
        wqkv = adjacent_rows((wq, wk, wv))
        if wqkv is not None:
            dn1 = N.gemm(dqkv, wqkv, b_mn=True)                                          # K = Nq + 2 Nkv
        else:
            dn1 = N.gemm(dqkv[:, :Nq], wq, b_mn=True)
            N.gemm(dqkv[:, Nq:Nq + Nkv], wk, b_mn=True, out=dn1, accumulate=True)
            N.gemm(dqkv[:, Nq + Nkv:], wv, b_mn=True, out=dn1, accumulate=True)
        dwq, dwk, dwv = _wgrad(dqkv, n1, (p_wq, p_wk, p_wv))
        del dqkv
        dx, dw_an = N.rmsnorm_bwd_res(dn1, x2, wan, rstd1, dh)                           # dx = dh + d attention_norm
        return (dx.view(B, S, D), None, None, None, None, None,
                dw_an.to(an_dtype), dwq, dwk, dwv, dwo, dw_fn.to(fn_dtype), dw1, dw2, dw3)


def _untouched(m) -> bool:
    """True when nothing has re-routed this module's forward (tensor-parallel styles patch ``forward`` on the instance or
    hang pre-hooks on it; such blocks keep the module-by-module path)."""
    return "forward" not in m.__dict__ and not m._forward_pre_hooks and not m._forward_hooks


def block_supported(block, x: torch.Tensor) -> bool:
    """The fused node handles plain (non tensor-parallel) bf16 CUDA blocks with head_dim 64 / 128."""
    att, ff = block.attention, block.feed_forward
    if not (x.is_cuda and x.dtype == _BF16 and x.dim() == 3 and x.shape[-1] % 8 == 0):
        return False
    lins = (att.wq, att.wk, att.wv, att.wo, ff.w1, ff.w2, ff.w3)
    mods = lins + (att, ff, block.attention_norm, block.ffn_norm)
    if not all(_untouched(m) for m in mods):
        return False
    if any(l.bias is not None or l.weight.dtype != _BF16 or not l.weight.is_contiguous() for l in lins):
        return False
    hd = att.head_dim
    return (hd in (64, 128) and att.wq.weight.shape[0] == att.n_heads * hd
            and att.wk.weight.shape[0] == att.n_kv_heads * hd and ff.w1.weight.shape[0] % 8 == 0)


def pack_block_weights(block) -> None:
    """Re-home wq|wk|wv and w1|w3 of a plain block into two contiguous buffers (values preserved) so the fused node can
    run each group as one GEMM.  FSDP does not need this: its unit layout already places the groups back to back."""
    att, ff = block.attention, block.feed_forward
    for group in ((att.wq, att.wk, att.wv), (ff.w1, ff.w3)):
        ws = [l.weight for l in group]
        if adjacent_rows([w.data for w in ws]) is not None or len({w.shape[1] for w in ws}) != 1:
            continue
        buf = torch.empty((sum(w.shape[0] for w in ws), ws[0].shape[1]), dtype=ws[0].dtype, device=ws[0].device)
        r = 0
        with torch.no_grad():
            for w in ws:
                buf[r:r + w.shape[0]].copy_(w.data)
                w.data = buf[r:r + w.shape[0]]
                r += w.shape[0]


def run_block(block, x, freqs_cis, pos_offset: int = 0):
    att, ff = block.attention, block.feed_forward
    fr = torch.view_as_real(freqs_cis) if freqs_cis.is_complex() else freqs_cis
    if not fr.is_contiguous():
        fr = fr.contiguous()
    return LlamaBlockFn.apply(x, fr, pos_offset, block.attention_norm.eps, att.n_heads, att.n_kv_heads,
                              block.attention_norm.weight, att.wq.weight, att.wk.weight, att.wv.weight, att.wo.weight,
                              block.ffn_norm.weight, ff.w1.weight, ff.w2.weight, ff.w3.weight)
