"""Ring (context-parallel) attention: the sequence is sharded across ranks, K/V chunks visit every rank, partial
results are merged with their log-sum-exp.

Docs-only in the reference (/root/reference/docs/guide/08_sequence_parallel.md:82-143: P steps of isend/irecv to the ring
neighbour overlapped with ``flash_attention_with_lse`` and an online-softmax merge).  On an NVSwitch box every peer is
one hop away, so the "ring" is only a schedule: at step t each rank PULLS the K/V chunk of rank (r - t) mod P straight
out of that rank's symmetric buffer (one peer-memory kernel, prefetched on a side stream while the tcgen05 flash
kernel works on the current chunk).  Causality is handled by the kernel's ``q_offset`` / ``k_offset`` arguments, so
future chunks cost nothing and the diagonal chunk is masked in-kernel.  Backward re-runs the chunk schedule with the
merged LSE; dK/dV partials are reduced to their owners with one reduce-scatter.
"""
from __future__ import annotations

import torch
from torch.autograd import Function

from .. import ops
from ..ops import functional as F_


def _merge(o, lse, o_i, lse_i):
    """Online-softmax merge of two partial attention results ([B,S,H,hd] with [B,H,S] log-sum-exps)."""
    new = torch.logaddexp(lse, lse_i)
    new_safe = torch.where(torch.isinf(new) & (new < 0), torch.zeros_like(new), new)
    w0 = torch.exp(lse - new_safe).transpose(1, 2).unsqueeze(-1)
    w1 = torch.exp(lse_i - new_safe).transpose(1, 2).unsqueeze(-1)
    return (o.float() * w0 + o_i.float() * w1), new


def _chunk_attention(q, k, v, causal, scale, q_off, k_off):
    if q.is_cuda:
        from ..ops import native as N

        return N.flash_attn_fwd(q, k, v, causal, scale, q_off, k_off)
    return F_._attention_reference(q, k, v, causal, scale, q_off, k_off)


class _RingAttn(Function):
    @staticmethod
    def forward(ctx, q, k, v, comm, causal, scale):
        P, r = comm.world, comm.rank
        B, s, H, hd = q.shape
        kv = torch.stack([k, v]).contiguous()                                   # [2, B, s, Hkv, hd]
        chunks = _gather_schedule(kv, comm)                                      # chunks[t] = kv of rank (r - t) % P
        o = torch.zeros((B, s, H, hd), dtype=torch.float32, device=q.device)
        lse = torch.full((B, H, s), float("-inf"), dtype=torch.float32, device=q.device)
        for t in range(P):
            j = (r - t) % P
            if causal and j > r:
                continue  # entirely in the future
            kj, vj = chunks[t][0], chunks[t][1]
            o_i, lse_i = _chunk_attention(q, kj, vj, causal, scale, r * s, j * s)
            o, lse = _merge(o, lse, o_i, lse_i)
        out = o.to(q.dtype)
        ctx.save_for_backward(q, k, v, out, lse)
        ctx.meta = (comm, causal, scale)
        return out

    @staticmethod
    def backward(ctx, do):
        q, k, v, out, lse = ctx.saved_tensors
        comm, causal, scale = ctx.meta
        P, r = comm.world, comm.rank
        B, s, H, hd = q.shape
        kv = torch.stack([k, v]).contiguous()
        chunks = _gather_schedule(kv, comm)
        dq = torch.zeros_like(q, dtype=torch.float32)
        dkv_part = torch.zeros((P,) + tuple(kv.shape), dtype=torch.float32, device=q.device)  # [owner j, 2, B, s, Hkv, hd]
        do = do.contiguous()
        for t in range(P):
            j = (r - t) % P
            if causal and j > r:
                continue
            kj, vj = chunks[t][0], chunks[t][1]
            if q.is_cuda:
                from ..ops import native as N

                dq_i, dk_i, dv_i = N.flash_attn_bwd(q, kj, vj, out, do, lse, causal, scale, r * s, j * s)
            else:
                dq_i, dk_i, dv_i = _reference_chunk_bwd(q, kj, vj, out, do, lse, causal, scale, r * s, j * s)
            dq += dq_i.float()
            dkv_part[j, 0] += dk_i.float()
            dkv_part[j, 1] += dv_i.float()
        mine = torch.empty(kv.shape, dtype=torch.float32, device=q.device)
        comm.reduce_scatter(mine, dkv_part, channel=6)
        return dq.to(q.dtype), mine[0].to(k.dtype), mine[1].to(v.dtype), None, None, None


def _gather_schedule(kv, comm):
    """chunks[t] = the K/V block of rank (rank - t) % P.  Peer pulls on CUDA (each rank reads a different peer through
    the broadcast kernel with a per-rank root); all-gather on CPU."""
    P, r = comm.world, comm.rank
    if comm.native and kv.is_cuda:
        nbytes = kv.numel() * kv.element_size()
        pad = (nbytes + 15) // 16 * 16
        buf = comm._stage(7, pad)
        buf.tensor(kv.dtype, 0, kv.numel()).copy_(kv.reshape(-1))
        from ..comm.collectives import I, LL  # noqa: F401
        from ..ops import _lib as L

        chunks = [kv]
        for t in range(1, P):
            dst = torch.empty(pad // kv.element_size(), dtype=kv.dtype, device=kv.device)
            L.call("comm_broadcast", *comm._ctx(buf, 7), L.ptr(dst), LL(0), LL(pad), I((r - t) % P), I(comm.max_ctas),
                   L.stream_ptr())
            chunks.append(dst[: kv.numel()].view(kv.shape))
        return chunks
    allkv = torch.empty((P,) + tuple(kv.shape), dtype=kv.dtype, device=kv.device)
    comm.all_gather(allkv, kv)
    return [allkv[(r - t) % P] for t in range(P)]


def _reference_chunk_bwd(q, k, v, out, do, lse, causal, scale, q_off, k_off):
    """CPU gradients of one chunk given the GLOBAL log-sum-exp (the same quantities the CUDA kernel computes)."""
    B, Sq, H, hd = q.shape
    Hkv = k.shape[2]
    g = H // Hkv
    qf, kf, vf = q.float().transpose(1, 2), k.float().transpose(1, 2), v.float().transpose(1, 2)
    kf_, vf_ = (kf.repeat_interleave(g, 1), vf.repeat_interleave(g, 1)) if g > 1 else (kf, vf)
    s = (qf @ kf_.transpose(-1, -2)) * scale
    if causal:
        qi = torch.arange(Sq).view(-1, 1) + q_off
        ki = torch.arange(k.shape[1]).view(1, -1) + k_off
        s = s.masked_fill(ki > qi, float("-inf"))
    lse_safe = torch.where(torch.isinf(lse), torch.full_like(lse, float("inf")), lse)
    p = torch.exp(s - lse_safe.unsqueeze(-1))
    dof, of = do.float().transpose(1, 2), out.float().transpose(1, 2)
    delta = (dof * of).sum(-1, keepdim=True)
    dv = p.transpose(-1, -2) @ dof
    dp = dof @ vf_.transpose(-1, -2)
    ds = p * (dp - delta)
    dq = (ds @ kf_) * scale
    dk = (ds.transpose(-1, -2) @ qf) * scale
    if g > 1:
        dk = dk.view(B, Hkv, g, *dk.shape[2:]).sum(2)
        dv = dv.view(B, Hkv, g, *dv.shape[2:]).sum(2)
    return dq.transpose(1, 2), dk.transpose(1, 2), dv.transpose(1, 2)


def ring_attention(q, k, v, comm, causal: bool = True, scale=None):
    """q, k, v: this rank's sequence chunk [B, S/P, H, hd] (chunk r holds positions r*S/P ...); returns the local output chunk."""
    scale = float(scale if scale is not None else q.shape[-1] ** -0.5)
    if comm.world == 1:
        return ops.attention(q, k, v, causal=causal, scale=scale)
    return _RingAttn.apply(q.contiguous(), k.contiguous(), v.contiguous(), comm, causal, scale)
