"""Fused compute+collective tensor-parallel linears (the north-star TP/SP kernels).

``ag_linear``  : sequence-sharded input -> ONE kernel that all-gathers the peers' chunks over NVLink (communication
                 CTAs inside the GEMM grid, TMA bulk copies, per-sub-chunk ready flags) while the tcgen05 tiles start
                 on the local chunk; further projections of the same input (wk, wv / w3) reuse the gathered buffer.
``linear_rs``  : GEMM whose epilogue stores every output row into the owner rank's staging slab over NVLink, then a
                 slab reduction (+ residual) at the owner: GEMM -> reduce-scatter without a separate collective pass.
Backward is the dual: ag_linear's input gradient is a GEMM->reduce-scatter (with the other projections' partials fused
in as the epilogue residual), linear_rs's is an all-gather->GEMM.

These replace the DTensor redistribute + cuBLAS pairs of the reference's hybrid plan
(/root/reference/scripts/06_hybrid_parallelism/01_fsdp_tp_hybrid.py:126-152; SURVEY.md §3.3 "hot comm #1-#4").
Used when tensors are CUDA bf16, batch == 1 and (S / T) % 256 == 0; otherwise ``tp.py`` falls back to the unfused
collective + GEMM path (same math).
"""
from __future__ import annotations

import ctypes as C
from typing import List, Optional

import torch
from torch.autograd import Function

from ..ops import _lib as L
from ..ops import native as N

I, LL, U = C.c_int, C.c_longlong, C.c_uint
ROWS_PER_FLAG = 256


class FusedWorkspace:
    """Per-communicator symmetric buffers for the fused kernels (grown on demand, collectively)."""

    def __init__(self, comm):
        self.comm = comm
        self.gather = {}     # (M, K) -> (PeerBuffer, flags tensor)
        self.stage = {}      # (M, N) -> PeerBuffer
        self.epoch = 0

    def gather_buf(self, M, K):
        key = (M, K)
        if key not in self.gather:
            buf = self.comm.symmetric_buffer(M * K * 2)
            flags = torch.zeros(M // ROWS_PER_FLAG + 8, dtype=torch.int32, device=self.comm.device)
            self.gather[key] = (buf, flags)
        return self.gather[key]

    def stage_buf(self, M, Nn):
        key = (M, Nn)
        if key not in self.stage:
            self.stage[key] = self.comm.symmetric_buffer(M * Nn * 2)
        return self.stage[key]


_ws = {}


def workspace(comm) -> FusedWorkspace:
    if id(comm) not in _ws:
        _ws[id(comm)] = FusedWorkspace(comm)
    return _ws[id(comm)]


def fusable(x: torch.Tensor, comm) -> bool:
    return (comm.native and x.is_cuda and x.dtype == torch.bfloat16 and x.dim() == 3 and x.shape[0] == 1
            and x.shape[1] % 256 == 0 and comm.world > 1)


def _ag_gemm(comm, x_shard2d, w, b_mn=False, bias=None, n_comm=4):
    """x_shard2d [m, K] -> (y [m*T, N], gathered X view [m*T, K])."""
    ws = workspace(comm)
    T, r = comm.world, comm.rank
    m, K = x_shard2d.shape
    M = m * T
    Nn = w.shape[1] if b_mn else w.shape[0]
    buf, flags = ws.gather_buf(M, K)
    xg = buf.tensor(torch.bfloat16, 0, M * K).view(M, K)
    xg[r * m:(r + 1) * m].copy_(x_shard2d)
    y = torch.empty((M, Nn), dtype=torch.bfloat16, device=x_shard2d.device)
    ws.epoch += 1
    ch = 8 + 0  # dedicated flag channel for the fused all-gather
    comm._seq[ch] += 1
    L.call("ag_gemm_bf16", buf.ptr_array, L.ptr(w), L.ptr(y), I(M), I(Nn), I(K), LL(w.stride(0)), LL(Nn), I(int(b_mn)),
           L.ptr(bias), L.ptr(None), LL(0), I(0), L.ptr(flags), I(ROWS_PER_FLAG), U(ws.epoch), I(n_comm),
           comm.flags.ptr_array, L.ptr(comm.counter), I(ch), U(comm._seq[ch]), I(T), I(r), L.stream_ptr())
    return y, xg


def _gemm_rs(comm, a, w, b_mn=False, residual=None, partial_residual=None):
    """reduce_scatter_rows(a @ w^T [+ partial_residual]) [+ residual]: a [M, K], result [M/T, N]."""
    ws = workspace(comm)
    T, r = comm.world, comm.rank
    M, K = a.shape
    Nn = w.shape[1] if b_mn else w.shape[0]
    stage = ws.stage_buf(M, Nn)
    L.call("gemm_rs_bf16", L.ptr(a), L.ptr(w), stage.ptr_array, I(M), I(Nn), I(K), LL(a.stride(0)), LL(w.stride(0)),
           I(int(b_mn)), L.ptr(partial_residual), LL(partial_residual.stride(0) if partial_residual is not None else 0),
           I(T), I(r), L.stream_ptr())
    out = torch.empty((M // T, Nn), dtype=torch.bfloat16, device=a.device)
    L.call("comm_slab_reduce", *comm._ctx(stage, 9), L.ptr(out), L.ptr(residual), LL(out.numel()), I(comm.max_ctas),
           L.stream_ptr())
    return out


class _AGLinear(Function):
    """ys[i] = all_gather_seq(x) @ ws[i]^T for several column-parallel weights sharing one gathered input."""

    @staticmethod
    def forward(ctx, comm, x, *weights):
        S_loc, K = x.shape[1], x.shape[2]
        y0, xg = _ag_gemm(comm, x.view(S_loc, K), weights[0])
        xg_saved = xg.clone()   # the workspace is reused by the next fused call
        ys = [y0] + [N.gemm(xg_saved, w) for w in weights[1:]]
        ctx.comm = comm
        ctx.save_for_backward(xg_saved, *weights)
        S = S_loc * comm.world
        return tuple(y.view(1, S, -1) for y in ys)

    @staticmethod
    def backward(ctx, *dys):
        comm = ctx.comm
        xg, *weights = ctx.saved_tensors
        dys = [d.reshape(-1, d.shape[-1]) for d in dys]
        dys = [d if d.is_contiguous() else d.contiguous() for d in dys]
        # dX_partial = sum_i dY_i @ W_i: all but the last accumulate locally, the last GEMM adds that partial in its
        # epilogue and scatters rows to their owners (GEMM -> reduce-scatter)
        part = None
        for d, w in zip(dys[:-1], weights[:-1]):
            if part is None:
                part = N.gemm(d, w, b_mn=True)
            else:
                N.gemm(d, w, b_mn=True, out=part, residual=part)
        dx = _gemm_rs(comm, dys[-1], weights[-1], b_mn=True, partial_residual=part)
        dws = [N.gemm(d, xg, a_mn=True, b_mn=True, out_dtype=w.dtype) for d, w in zip(dys, weights)]
        return (None, dx.view(1, dx.shape[0], dx.shape[1]), *dws)


class _LinearRS(Function):
    """reduce_scatter_seq(x @ w^T) + residual (row-parallel weight, sequence-sharded output)."""

    @staticmethod
    def forward(ctx, comm, x, w, residual):
        x2 = x.reshape(-1, x.shape[-1])
        res2 = residual.reshape(-1, residual.shape[-1]) if residual is not None else None
        out = _gemm_rs(comm, x2 if x2.is_contiguous() else x2.contiguous(), w, residual=res2)
        ctx.comm = comm
        ctx.save_for_backward(x2, w)
        ctx.has_res = residual is not None
        return out.view(1, out.shape[0], out.shape[1])

    @staticmethod
    def backward(ctx, dy):
        comm = ctx.comm
        x2, w = ctx.saved_tensors
        d2 = dy.reshape(-1, dy.shape[-1])
        d2 = d2 if d2.is_contiguous() else d2.contiguous()
        dx, dyg = _ag_gemm(comm, d2, w, b_mn=True)          # all-gather(dY) -> GEMM with W [N, K_local] as [K=N, N=K_local]
        dw = N.gemm(dyg, x2, a_mn=True, b_mn=True, out_dtype=w.dtype)
        return None, dx.view(1, dx.shape[0], dx.shape[1]), dw, (dy if ctx.has_res else None)


class _LinearAR(Function):
    """all_reduce(x @ w^T) for a row-parallel weight with a REPLICATED output (plain Megatron TP without sequence
    parallelism: RowwiseParallel(), scripts/03_tensor_parallel_tp/tensor_parallel_vit.py:357-359).
    Two-shot all-reduce whose first half is fused into the GEMM: the epilogue scatters row blocks to their owners'
    slabs (GEMM -> reduce-scatter), the owners sum them, and a pull all-gather replicates the reduced blocks - the
    partial [M, N] product is never written locally."""

    @staticmethod
    def forward(ctx, comm, x, w, bias):
        x2 = x.reshape(-1, x.shape[-1])
        x2 = x2 if x2.is_contiguous() else x2.contiguous()
        chunk = _gemm_rs(comm, x2, w)
        if bias is not None:
            chunk = chunk + bias
        out = torch.empty((x2.shape[0], chunk.shape[1]), dtype=chunk.dtype, device=chunk.device)
        comm.all_gather(out, chunk)
        ctx.save_for_backward(x2, w)
        ctx.has_bias = bias is not None
        return out.view(*x.shape[:-1], out.shape[-1])

    @staticmethod
    def backward(ctx, dy):
        x2, w = ctx.saved_tensors
        d2 = dy.reshape(-1, dy.shape[-1])
        d2 = d2 if d2.is_contiguous() else d2.contiguous()
        dx = N.gemm(d2, w, b_mn=True)                      # dY is replicated: the input gradient is purely local
        dw = N.gemm(d2, x2, a_mn=True, b_mn=True, out_dtype=w.dtype)
        db = d2.float().sum(0).to(dy.dtype) if ctx.has_bias else None
        return None, dx.view(*dy.shape[:-1], dx.shape[-1]), dw, db


class LlamaBlockTPFn(Function):
    """One tensor-parallel + sequence-parallel Llama block as a single autograd node on the fused kernels.

    Input / output: this rank's sequence shard [1, S/T, D].  Per block, forward and backward together, the node issues

    * 4 all-gather -> GEMM kernels: ``[q|k|v]`` (ONE GEMM over the row-stacked wq|wk|wv shard), ``[w1|w3]`` (one GEMM over
      the stacked w1|w3 shard), and in backward the two row-parallel input gradients (gather dY, multiply by W);
    * 4 GEMM -> reduce-scatter kernels: wo and w2 (+ residual) forward, and the input gradients of the two stacked
      column-parallel projections (K = 3 D / T resp. 2 F / T: the partial products of q, k, v accumulate in TMEM);
    * RoPE in place inside the gathered projection output, attention on the local heads, SwiGLU on the two halves of
      the stacked output, RMSNorm on the sequence shard with the residual gradient folded into its backward.

    Reference plan: /root/reference/scripts/06_hybrid_parallelism/01_fsdp_tp_hybrid.py:126-152 (SequenceParallel norms,
    ColwiseParallel wq/wk/wv/w1/w3, RowwiseParallel wo/w2 with Shard(1) outputs); model llama2_model.py:207,267.
    """

    @staticmethod
    def forward(ctx, comm, x, freqs_real, pos_offset, eps, n_heads_local, n_kv_local, w_an, wq, wk, wv, wo, w_fn, w1, w2, w3):
        from ..models.fused_block import adjacent_rows

        S_loc, D = x.shape[1], x.shape[2]
        T = comm.world
        S = S_loc * T
        hd = wq.shape[0] // n_heads_local
        Nq, Nkv, Ffn = n_heads_local * hd, n_kv_local * hd, w1.shape[0]
        x2 = x.reshape(S_loc, D)
        wqkv, w13 = adjacent_rows((wq, wk, wv)), adjacent_rows((w1, w3))
        assert wqkv is not None and w13 is not None, "fused TP block: pack the weight shards first (enable_fused_tp does)"
        n1, rstd1 = N.rmsnorm_fwd(x2, w_an if w_an.dtype == x.dtype else w_an.to(x.dtype), eps)
        qkv, xg = _ag_gemm(comm, n1, wqkv)                                   # [S, Nq + 2 Nkv], gathered n1 [S, D]
        xg1 = xg.clone()                                                     # the workspace is reused by the next gather
        N.rope_strided(qkv, freqs_real, S, n_heads_local + n_kv_local, hd, pos_offset)
        q4 = qkv[:, :Nq].view(1, S, n_heads_local, hd)
        k4 = qkv[:, Nq:Nq + Nkv].view(1, S, n_kv_local, hd)
        v4 = qkv[:, Nq + Nkv:].view(1, S, n_kv_local, hd)
        scale = float(hd) ** -0.5
        o4, lse = N.flash_attn_fwd(q4, k4, v4, True, scale, 0, 0)
        h = _gemm_rs(comm, o4.view(S, Nq), wo, residual=x2)                  # [S/T, D] = x + attn(x), my rows
        n2, rstd2 = N.rmsnorm_fwd(h, w_fn if w_fn.dtype == x.dtype else w_fn.to(x.dtype), eps)
        ab, xg = _ag_gemm(comm, n2, w13)                                     # [S, 2 F/T]
        xg2 = xg.clone()
        s = N.swiglu_fwd(ab[:, :Ffn], ab[:, Ffn:])
        y = _gemm_rs(comm, s, w2, residual=h)
        ctx.comm = comm
        ctx.save_for_backward(x2, rstd1, xg1, qkv, o4, lse, h, rstd2, xg2, ab, s, freqs_real, w_an, wqkv, wo, w_fn, w13, w2)
        ctx.meta = (S_loc, S, D, hd, n_heads_local, n_kv_local, Ffn, pos_offset, scale, eps)
        ctx.params = (w_an, wq, wk, wv, wo, w_fn, w1, w2, w3)
        return y.view(1, S_loc, D)

    @staticmethod
    def backward(ctx, dy):
        from ..models.fused_block import _wgrad

        comm = ctx.comm
        (x2, rstd1, xg1, qkv, o4, lse, h, rstd2, xg2, ab, s, freqs_real, w_an, wqkv, wo, w_fn, w13, w2) = ctx.saved_tensors
        S_loc, S, D, hd, Hq, Hkv, Ffn, pos_offset, scale, eps = ctx.meta
        p_an, p_wq, p_wk, p_wv, p_wo, p_fn, p_w1, p_w2, p_w3 = ctx.params
        Nq, Nkv = Hq * hd, Hkv * hd
        dy2 = dy.reshape(S_loc, D)
        dy2 = dy2 if (dy2.is_contiguous() and dy2.dtype == torch.bfloat16) else dy2.to(torch.bfloat16).contiguous()
        wan = w_an if w_an.dtype == x2.dtype else w_an.to(x2.dtype)
        wfn = w_fn if w_fn.dtype == x2.dtype else w_fn.to(x2.dtype)
        # ---- feed-forward half ----
        ds, dyg = _ag_gemm(comm, dy2, w2, b_mn=True)                         # gather dY rows, ds = dY @ W2_shard  [S, F/T]
        (dw2,) = _wgrad(dyg, s, (p_w2,))                                     # uses the gathered dY before the next gather
        dab = torch.empty_like(ab)
        N.swiglu_bwd(ds, ab[:, :Ffn], ab[:, Ffn:], da=dab[:, :Ffn], db=dab[:, Ffn:])
        dn2 = _gemm_rs(comm, dab, w13, b_mn=True)                            # [S/T, D]: sum over ranks of dab @ W13_shard
        dw1, dw3 = _wgrad(dab, xg2, (p_w1, p_w3))
        dh, dw_fn = N.rmsnorm_bwd_res(dn2, h, wfn, rstd2, dy2)
        # ---- attention half ----
        do, dhg = _ag_gemm(comm, dh, wo, b_mn=True)                          # [S, Nq]
        (dwo,) = _wgrad(dhg, o4.view(S, Nq), (p_wo,))
        q4 = qkv[:, :Nq].view(1, S, Hq, hd)
        k4 = qkv[:, Nq:Nq + Nkv].view(1, S, Hkv, hd)
        v4 = qkv[:, Nq + Nkv:].view(1, S, Hkv, hd)
        dqkv = torch.empty_like(qkv)
        N.flash_attn_bwd(q4, k4, v4, o4, do.view(1, S, Hq, hd), lse, True, scale, 0, 0,
                         dq=dqkv[:, :Nq].view(1, S, Hq, hd), dk=dqkv[:, Nq:Nq + Nkv].view(1, S, Hkv, hd),
                         dv=dqkv[:, Nq + Nkv:].view(1, S, Hkv, hd))
        N.rope_strided(dqkv, freqs_real, S, Hq + Hkv, hd, pos_offset, backward=True)
        dn1 = _gemm_rs(comm, dqkv, wqkv, b_mn=True)
        dwq, dwk, dwv = _wgrad(dqkv, xg1, (p_wq, p_wk, p_wv))
        dx, dw_an = N.rmsnorm_bwd_res(dn1, x2, wan, rstd1, dh)
        return (None, dx.view(1, S_loc, D), None, None, None, None, None,
                dw_an.to(p_an.dtype), dwq, dwk, dwv, dwo, dw_fn.to(p_fn.dtype), dw1, dw2, dw3)


def block_tp_supported(blk, x, comm) -> bool:
    att, ffn = blk.attention, blk.feed_forward
    hd = att.head_dim
    return (fusable(x, comm) and hd in (64, 128) and all(l.bias is None for l in (att.wq, att.wk, att.wv, att.wo, ffn.w1, ffn.w2, ffn.w3))
            and att.wq.weight.shape[0] % hd == 0 and att.wk.weight.shape[0] % hd == 0)


def run_block_tp(blk, x, freqs_cis, pos_offset, comm):
    """x: this rank's sequence shard [1, S/T, D] of a block that went through ``llama_tp_plan(sequence_parallel=True)``."""
    att, ffn = blk.attention, blk.feed_forward
    hd = att.head_dim
    fr = torch.view_as_real(freqs_cis) if freqs_cis.is_complex() else freqs_cis
    if not fr.is_contiguous():
        fr = fr.contiguous()
    return LlamaBlockTPFn.apply(comm, x, fr, pos_offset, blk.attention_norm.eps, att.wq.weight.shape[0] // hd,
                                att.wk.weight.shape[0] // hd, blk.attention_norm.weight, att.wq.weight, att.wk.weight,
                                att.wv.weight, att.wo.weight, blk.ffn_norm.weight, ffn.w1.weight, ffn.w2.weight, ffn.w3.weight)


def linear_ar(comm, x, w, bias: Optional[torch.Tensor] = None):
    """Row-parallel linear with an all-reduced (replicated) output; rows of ``x`` must be divisible by the TP degree."""
    return _LinearAR.apply(comm, x, w, bias)


def ag_linear(comm, x, weights: List[torch.Tensor]):
    return _AGLinear.apply(comm, x, *weights)


def linear_rs(comm, x, w, residual: Optional[torch.Tensor] = None):
    return _LinearRS.apply(comm, x, w, residual)
