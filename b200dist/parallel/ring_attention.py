"""Ring (context-parallel) attention: the sequence is sharded across ranks, K/V chunks visit every rank, partial
results are merged with their log-sum-exp.

Docs-only in the reference (/root/reference/docs/guide/08_sequence_parallel.md:82-143: P steps of isend/irecv to the ring
neighbour overlapped with ``flash_attention_with_lse`` and an online-softmax merge).  On an NVSwitch box every peer is
one hop away, so the "ring" is only a schedule.  FORWARD on CUDA is ONE kernel (``_ring_fused_forward``): communication
CTAs inside the flash-attention grid stage the peers' K/V chunks into a local gathered buffer block by block while the
attention CTAs consume the blocks that have landed; the running (O, m, l) never leave the chip.  The chunk-loop below
(each rank PULLS the K/V chunk of rank (r - t) mod P out of that rank's symmetric buffer, one flash kernel per chunk, LSE
merge) remains for backward, for CPU tensors and for shapes the fused kernel does not take.  Causality is handled by the kernel's ``q_offset`` / ``k_offset`` arguments, so
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


_RING_FUSED = __import__("os").environ.get("B200DIST_RING_FUSED", "1") not in ("", "0")
_RING_CHANNEL = 10
_ring_ws = {}


def _fused_ok(q, k, comm) -> bool:
    return (_RING_FUSED and comm.native and comm.world > 1 and q.is_cuda and q.dtype == torch.bfloat16
            and q.shape[-1] in (64, 128) and q.shape[1] % 128 == 0 and k.shape[1] == q.shape[1]
            and q.shape[2] % k.shape[2] == 0)


def _ring_fused_forward(q, k, v, comm, causal, scale):
    """Ring attention forward as ONE kernel (csrc/flash_attn_fwd.cu, RingParams): this rank's K/V chunk is placed in a
    symmetric gathered buffer, communication CTAs inside the attention grid stage the peers' chunks into the local copy
    block by block (ascending key order, ready flag per 128-key block) and the attention CTAs consume the blocks as they
    land.  Returns (out, global log-sum-exp): no per-chunk partials, no merge."""
    import ctypes as C

    from ..ops import _lib as L

    P, r = comm.world, comm.rank
    B, s, H, hd = q.shape
    Hkv = k.shape[2]
    S_total = P * s
    n_el = 2 * B * S_total * Hkv * hd
    key = id(comm)
    ws = _ring_ws.get(key)
    if ws is None or ws["buf"].nbytes < n_el * 2 or ws["flags"].numel() < 2 * B * (S_total // 128):
        ws = {"buf": comm.symmetric_buffer(n_el * 2),                         # collective, same order on every rank
              "flags": torch.zeros(2 * B * (S_total // 128) + 8, dtype=torch.int32, device=q.device), "epoch": 0}
        _ring_ws[key] = ws
    kv = ws["buf"].tensor(torch.bfloat16, 0, n_el).view(2, B, S_total, Hkv, hd)
    kv[0, :, r * s:(r + 1) * s].copy_(k)
    kv[1, :, r * s:(r + 1) * s].copy_(v)
    qc = q if (q.stride(-1) == 1 and not any(st % 8 for st in q.stride()[:-1]) and not (q.data_ptr() & 15)) else q.contiguous()
    out = torch.empty((B, s, H, hd), dtype=q.dtype, device=q.device)
    lse = torch.empty((B, H, s), dtype=torch.float32, device=q.device)
    ws["epoch"] += 1
    comm._seq[_RING_CHANNEL] += 1
    I, LL, U, F = C.c_int, C.c_longlong, C.c_uint, C.c_float
    L.call("flash_attn_fwd_ring", L.ptr(qc), ws["buf"].ptr_array, L.ptr(out), L.ptr(lse), I(B), I(H), I(Hkv), I(s), I(s), I(hd),
           LL(qc.stride(0)), LL(qc.stride(1)), LL(qc.stride(2)), LL(out.stride(0)), LL(out.stride(1)), LL(out.stride(2)),
           F(scale), I(int(causal)), L.ptr(ws["flags"]), U(ws["epoch"]), I(8), comm.flags.ptr_array, L.ptr(comm.counter),
           I(_RING_CHANNEL), U(comm._seq[_RING_CHANNEL] & 0xFFFFFFFF), I(P), I(r), L.stream_ptr())
    return out, lse


class _RingAttn(Function):
    @staticmethod
    def forward(ctx, q, k, v, comm, causal, scale):
        P, r = comm.world, comm.rank
        B, s, H, hd = q.shape
        if _fused_ok(q, k, comm):
            out, lse = _ring_fused_forward(q, k, v, comm, causal, scale)
            ctx.save_for_backward(q, k, v, out, lse)
            ctx.meta = (comm, causal, scale)
            return out
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


# --------------------------------------------------------------------------- zig-zag (load-balanced causal) variant
def zigzag_positions(seq_len: int, world: int, rank: int):
    """Global token ranges held by ``rank`` under the zig-zag layout: the sequence is cut into 2 * world blocks and rank r
    owns blocks r and 2 * world - 1 - r, so every rank has the same amount of causal work (with contiguous chunks rank
    P-1 attends to P chunks and rank 0 to one).  Returns [(start, stop), (start, stop)]."""
    assert seq_len % (2 * world) == 0, "zig-zag ring attention needs seq_len % (2 * world) == 0"
    h = seq_len // (2 * world)
    return [(rank * h, (rank + 1) * h), ((2 * world - 1 - rank) * h, (2 * world - rank) * h)]


def zigzag_shard(x: torch.Tensor, world: int, rank: int, dim: int = 1) -> torch.Tensor:
    """Slice a full-sequence tensor into this rank's zig-zag shard (the two blocks concatenated along ``dim``)."""
    (a0, a1), (b0, b1) = zigzag_positions(x.shape[dim], world, rank)
    return torch.cat([x.narrow(dim, a0, a1 - a0), x.narrow(dim, b0, b1 - b0)], dim=dim).contiguous()


def zigzag_unshard(parts, dim: int = 1) -> torch.Tensor:
    """Inverse of ``zigzag_shard`` given the shards of all ranks in rank order."""
    world = len(parts)
    h = parts[0].shape[dim] // 2
    blocks = [None] * (2 * world)
    for r, p in enumerate(parts):
        blocks[r] = p.narrow(dim, 0, h)
        blocks[2 * world - 1 - r] = p.narrow(dim, h, h)
    return torch.cat(blocks, dim=dim)


class _ZigzagRingAttn(Function):
    """Causal ring attention over zig-zag shards: every (local q block, visiting k/v block) pair is one call of the chunk
    kernel with the blocks' global offsets; pairs that lie entirely in the future are skipped.  Each rank runs exactly
    2 * world + 1 block pairs (of which two are diagonal), independent of its rank."""

    @staticmethod
    def _pairs(P, r, t):
        """Block-pair list for visiting rank j = (r - t) % P: [(q_seg, kv_seg, q_block, kv_block)], future pairs dropped."""
        j = (r - t) % P
        qb, kb = (r, 2 * P - 1 - r), (j, 2 * P - 1 - j)
        return j, [(a, b, qb[a], kb[b]) for a in range(2) for b in range(2) if kb[b] <= qb[a]]

    @staticmethod
    def forward(ctx, q, k, v, comm, scale):
        P, r = comm.world, comm.rank
        B, s, H, hd = q.shape
        h = s // 2
        kv = torch.stack([k, v]).contiguous()                                   # [2, B, s, Hkv, hd]
        chunks = _gather_schedule(kv, comm)
        qs = [q[:, :h].contiguous(), q[:, h:].contiguous()]
        o = [torch.zeros((B, h, H, hd), dtype=torch.float32, device=q.device) for _ in range(2)]
        lse = [torch.full((B, H, h), float("-inf"), dtype=torch.float32, device=q.device) for _ in range(2)]
        for t in range(P):
            _j, pairs = _ZigzagRingAttn._pairs(P, r, t)
            for a, b, qblk, kblk in pairs:
                kj = chunks[t][0][:, b * h:(b + 1) * h].contiguous()
                vj = chunks[t][1][:, b * h:(b + 1) * h].contiguous()
                o_i, lse_i = _chunk_attention(qs[a], kj, vj, True, scale, qblk * h, kblk * h)
                o[a], lse[a] = _merge(o[a], lse[a], o_i, lse_i)
        out = torch.cat(o, dim=1).to(q.dtype)
        ctx.save_for_backward(q, k, v, out, lse[0], lse[1])
        ctx.meta = (comm, scale)
        return out

    @staticmethod
    def backward(ctx, do):
        q, k, v, out, lse0, lse1 = ctx.saved_tensors
        comm, scale = ctx.meta
        P, r = comm.world, comm.rank
        B, s, H, hd = q.shape
        h = s // 2
        kv = torch.stack([k, v]).contiguous()
        chunks = _gather_schedule(kv, comm)
        qs = [q[:, :h].contiguous(), q[:, h:].contiguous()]
        outs = [out[:, :h].contiguous(), out[:, h:].contiguous()]
        dos = [do[:, :h].contiguous(), do[:, h:].contiguous()]
        lses = [lse0, lse1]
        dq = [torch.zeros((B, h, H, hd), dtype=torch.float32, device=q.device) for _ in range(2)]
        dkv_part = torch.zeros((P,) + tuple(kv.shape), dtype=torch.float32, device=q.device)   # [owner, 2, B, s, Hkv, hd]
        for t in range(P):
            j, pairs = _ZigzagRingAttn._pairs(P, r, t)
            for a, b, qblk, kblk in pairs:
                kj = chunks[t][0][:, b * h:(b + 1) * h].contiguous()
                vj = chunks[t][1][:, b * h:(b + 1) * h].contiguous()
                if q.is_cuda:
                    from ..ops import native as N

                    dq_i, dk_i, dv_i = N.flash_attn_bwd(qs[a], kj, vj, outs[a], dos[a], lses[a], True, scale, qblk * h, kblk * h)
                else:
                    dq_i, dk_i, dv_i = _reference_chunk_bwd(qs[a], kj, vj, outs[a], dos[a], lses[a], True, scale,
                                                            qblk * h, kblk * h)
                dq[a] += dq_i.float()
                dkv_part[j, 0, :, b * h:(b + 1) * h] += dk_i.float()
                dkv_part[j, 1, :, b * h:(b + 1) * h] += dv_i.float()
        mine = torch.empty(kv.shape, dtype=torch.float32, device=q.device)
        comm.reduce_scatter(mine, dkv_part, channel=6)
        return torch.cat(dq, dim=1).to(q.dtype), mine[0].to(k.dtype), mine[1].to(v.dtype), None, None


def zigzag_ring_attention(q, k, v, comm, scale=None):
    """Causal ring attention on ZIG-ZAG shards (``zigzag_shard``): q, k, v are [B, S/P, H, hd] holding blocks r and
    2P-1-r of the sequence.  Same result as ``ring_attention`` on contiguous shards, but every rank does the same amount
    of work.  Built from the chunk kernels ``ring_attention`` uses; forward and backward checked
    against unsharded flash attention on CPU (gloo) and on 2 B200s (``profiles/zigzag_linear_ar_check_2gpu_r2.log``)."""
    scale = float(scale if scale is not None else q.shape[-1] ** -0.5)
    if comm.world == 1:
        return ops.attention(q, k, v, causal=True, scale=scale)
    assert q.shape[1] % 2 == 0
    return _ZigzagRingAttn.apply(q.contiguous(), k.contiguous(), v.contiguous(), comm, scale)
