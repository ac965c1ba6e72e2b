"""Fully-sharded data parallelism, per-block, with prefetch and a fused reduce-scatter + AdamW.

What the reference does (all through torch): FSDP1 ``FullyShardedDataParallel(FULL_SHARD,
size_based_auto_wrap_policy, MixedPrecision(bf16))`` in scripts/02_fully_sharded_fsdp/*.py:194-212 and
FSDP2 ``fully_shard(model, mesh=dp_mesh)`` in scripts/06_hybrid_parallelism/01_fsdp_tp_hybrid.py:155 —
all-gather parameters before forward and before backward, reduce-scatter gradients, step a
sharded optimizer.

What this module does on B200:

* **Units** — every block of a ``ModuleList`` (Llama ``layers``) is one unit, the remaining
  parameters (embedding, final norm, LM head) form the root unit.  A unit's parameters live in ONE
  flat bf16 buffer; each rank owns 1/world of it.
* **Shards in peer memory** — the bf16 shard sits in a symmetric buffer mapped on every rank, so
  the all-gather is a zero-copy pull over NVLink (``comm_allgather``) straight into the unit's flat
  buffer; parameters are views into that buffer, whose storage is released after forward
  (``reshard_after_forward``) and re-materialised before backward.
* **Prefetch** — while unit i computes, unit i+1 (forward) / i-1 (backward) is gathered on a side
  stream.  Buffers are (re)allocated on the compute stream and handed to the side stream through
  events, so the caching allocator never reuses memory that a peer pull still writes.
* **Fused gradient path** — when a unit's last gradient lands, gradients are packed into a
  symmetric staging buffer and ONE kernel per unit does reduce-scatter + 1/world scale + bf16->fp32
  cast + AdamW on the local shard + write-back of the bf16 shard (``comm_rs_adamw``): the reduced
  gradient never exists in HBM and the optimizer needs no separate pass.  ``optimizer.step()`` then
  only joins the streams.  (``fused_optimizer=False`` keeps reduce-scatter and the update separate
  for gradient accumulation.)
* 180 GB HBM3e: optimizer state is fp32 master + two moments per shard; for 7B on 8 GPUs that
  is 13.5 GB per rank, so ``reshard_after_forward=False`` (ZeRO-2 style, one gather per step) is
  also practical and selectable.

CPU tensors (gloo unit tests) run the same control flow on torch.distributed collectives.
"""
from __future__ import annotations

import contextlib
import os

import math
from typing import Dict, List, Optional

import torch
import torch.distributed as dist
from torch import nn

from ..comm import get_communicator
from ..ops import native as N


class _Unit:
    def __init__(self, name: str, module: nn.Module, params: List[nn.Parameter], names: List[str]):
        self.name, self.module, self.params, self.names = name, module, params, names
        self.offsets: List[int] = []
        self.total = 0
        self.shard_numel = 0
        self.full: Optional[torch.Tensor] = None
        self.shard_buf = None          # PeerBuffer (native) holding the compute-dtype shard
        self.shard: Optional[torch.Tensor] = None  # local view of the shard
        self.master = self.m = self.v = None
        self.grad_shard = None         # fp32 [shard_numel] (unfused path)
        self.sharded = True
        self.ag_event = None
        self.pending_grads = 0
        self.grad_ready: List[bool] = []
        self.rs_event = None
        self.index = 0
        self.stage_k = None            # staging buffer claimed for the current backward
        self.sink_written: List[bool] = []


def _default_units(module: nn.Module):
    """Each element of a top-level ModuleList becomes a unit (Llama ``layers``); the rest is root."""
    units = []
    for name, child in module.named_children():
        if isinstance(child, nn.ModuleList) and len(child) > 0:
            for i, blk in enumerate(child):
                units.append((f"{name}.{i}", blk))
    return units


def _ordered_named_parameters(module: nn.Module):
    """``named_parameters()`` with the module's ``adjacent_param_groups`` (tuples of parameter names) pulled together:
    the members of a group follow each other in the unit's flat buffer, so a consumer can address them as ONE row-stacked
    matrix (q|k|v and w1|w3 of a Llama block -> one GEMM each, models/fused_block.py)."""
    items = list(module.named_parameters())
    groups = getattr(module, "adjacent_param_groups", ())
    if not groups:
        return items
    by_name = dict(items)
    member = {n: g for g in groups for n in g if all(m in by_name for m in g)}
    out, done = [], set()
    for n, p in items:
        if n in done:
            continue
        g = member.get(n)
        if g is None:
            out.append((n, p)); done.add(n)
        else:
            for m in g:
                out.append((m, by_name[m])); done.add(m)
    return out


class FSDP(nn.Module):
    def __init__(self, module: nn.Module, process_group=None, units: Optional[List] = None,
                 reshard_after_forward: bool = True, fused_optimizer: bool = True,
                 comm_max_ctas: int = 64, sync_module_states: bool = True, copy_engine_gather: bool = True,
                 copy_engine_reduce: bool = True, replicate_group=None, nvls_reduce: bool = True):
        super().__init__()
        self.module = module
        self.comm = get_communicator(process_group)
        # HYBRID_SHARD (the reference only tabulates it, docs/guide/05_fully_sharded_fsdp.md:114-132): shard inside
        # ``process_group``, replicate across ``replicate_group``.  Each shard's reduced gradient is averaged over the
        # replicas before the optimizer step, which therefore runs unfused (reduce-scatter, all-reduce, then AdamW).
        self.replicate_comm = get_communicator(replicate_group) if replicate_group is not None else None
        if self.replicate_comm is not None and self.replicate_comm.world > 1:
            fused_optimizer = False
        else:
            self.replicate_comm = None
        self.world, self.rank = self.comm.world, self.comm.rank
        self.native = self.comm.native and next(module.parameters()).is_cuda
        self.reshard_after_forward = reshard_after_forward and self.world > 1
        self.fused_optimizer = fused_optimizer
        self.copy_engine_gather = copy_engine_gather   # parameter all-gathers as DMA peer copies (no SMs)
        self.copy_engine_reduce = copy_engine_reduce   # gradient slices pulled by DMA, then a local reduce+AdamW kernel
        self.device = next(module.parameters()).device
        self.dtype = next(module.parameters()).dtype
        if self.native:
            self.comm.max_ctas = comm_max_ctas
            self.ag_stream = torch.cuda.Stream(self.device)
            self.rs_stream = torch.cuda.Stream(self.device)
        self._opt = None
        self.nvls_grads = False
        self._step_t = 0
        self._in_backward = False
        self._bwd_seen = 0
        self.require_backward_grad_sync = True

        unit_mods = units if units is not None else _default_units(module)
        claimed = set()
        self.units: List[_Unit] = []
        for uname, um in unit_mods:
            ps, ns = [], []
            for n, p in _ordered_named_parameters(um):
                if id(p) not in claimed:
                    claimed.add(id(p)); ps.append(p); ns.append(f"{uname}.{n}")
            if ps:
                self.units.append(_Unit(uname, um, ps, ns))
        ps, ns = [], []
        for n, p in module.named_parameters():
            if id(p) not in claimed:
                claimed.add(id(p)); ps.append(p); ns.append(n)
        self.root_unit = _Unit("<root>", module, ps, ns) if ps else None
        # execution order: root first (embedding is used first), then blocks
        self.exec_units: List[_Unit] = ([self.root_unit] if self.root_unit else []) + self.units
        for i, u in enumerate(self.exec_units):
            u.index = i
        if sync_module_states and self.replicate_comm is not None:
            for u in self.exec_units:   # all replicas start from replica 0's weights
                for p in u.params:
                    self.replicate_comm.broadcast(p.data, src=0) if p.data.is_contiguous() else None
        if sync_module_states and self.world > 1:
            for u in self.exec_units:
                for p in u.params:
                    self.comm.broadcast(p.data, src=0) if p.data.is_contiguous() else None
        for u in self.exec_units:
            self._init_unit(u)
        max_total = max(u.total for u in self.exec_units)
        if self.native and self.world > 1:
            esize = torch.empty((), dtype=self.dtype).element_size()
            # When the platform offers NVSwitch multicast, the two staging buffers live behind a multicast object: the reduce-
            # scatter is then ONE multimem.ld_reduce pass over this rank's slice, fused with AdamW (comm_rs_adamw_nvls) -
            # no DMA staging of the peers' slabs, no scratch traffic in HBM, one arrive / depart instead of two barriers.
            self.grad_stage, self.nvls_grads = None, False
            if nvls_reduce and self.dtype == torch.bfloat16 and self.world >= int(os.environ.get("B200DIST_FSDP_NVLS_MIN_WORLD", "2")):
                bufs = [self.comm.nvls_buffer(max_total * esize) for _ in range(2)]
                if all(b is not None for b in bufs):
                    self.grad_stage, self.nvls_grads = bufs, True
                else:
                    for b in bufs:
                        if b is not None:
                            b.close()
            if self.grad_stage is None:
                self.grad_stage = [self.comm.symmetric_buffer(max_total * esize) for _ in range(2)]
            self._stage_events = [None, None]
        self._stage_idx = 0
        self._register_hooks()

    # ------------------------------------------------------------------ layout
    def _init_unit(self, u: _Unit) -> None:
        n = 0
        for p in u.params:
            u.offsets.append(n)
            n += (p.numel() + 7) // 8 * 8
        align = 8 * self.world
        u.total = (n + align - 1) // align * align
        u.shard_numel = u.total // self.world
        full = torch.zeros(u.total, dtype=self.dtype, device=self.device)
        for p, o in zip(u.params, u.offsets):
            full[o:o + p.numel()].copy_(p.data.reshape(-1))
        lo = self.rank * u.shard_numel
        if self.world == 1:
            u.shard = full  # nothing to shard: the "shard" is the flat buffer itself
        elif self.native:
            u.shard_buf = self.comm.symmetric_buffer(u.shard_numel * full.element_size())
            u.shard = u.shard_buf.tensor(self.dtype, 0, u.shard_numel)
            u.shard.copy_(full[lo:lo + u.shard_numel])
        else:
            u.shard = full[lo:lo + u.shard_numel].clone()
        u.master = full[lo:lo + u.shard_numel].to(torch.float32, copy=True)
        u.full = full
        for p, o in zip(u.params, u.offsets):
            p.data = full[o:o + p.numel()].view(p.shape)
        u.sharded = False
        u.grad_ready = [False] * len(u.params)
        u.pending_grads = len(u.params)
        if self.world > 1:
            self._free(u)

    def _free(self, u: _Unit) -> None:
        if self.world == 1 or u.sharded:
            return
        u.full.untyped_storage().resize_(0)
        u.sharded = True
        u.ag_event = None

    # ------------------------------------------------------------------ unshard / prefetch
    def _issue_gather(self, u: _Unit) -> None:
        """Allocate the flat buffer on the compute stream and start the all-gather (side stream)."""
        if not u.sharded or self.world == 1:
            return
        nbytes = u.total * u.full.element_size()
        u.full.untyped_storage().resize_(nbytes)
        u.sharded = False
        if self.native:
            ev = torch.cuda.Event()
            ev.record(torch.cuda.current_stream(self.device))
            with torch.cuda.stream(self.ag_stream):
                self.ag_stream.wait_event(ev)
                self.comm.all_gather_from(u.full, u.shard_buf, 0, u.shard_numel * u.full.element_size(), channel=1,
                                          copy_engine=self.copy_engine_gather)
                u.ag_event = torch.cuda.Event()
                u.ag_event.record(self.ag_stream)
        else:
            dist.all_gather_into_tensor(u.full, u.shard, group=self.comm.group)
            u.ag_event = None

    def _wait_gather(self, u: _Unit) -> None:
        if u.ag_event is not None:
            torch.cuda.current_stream(self.device).wait_event(u.ag_event)
            u.ag_event = None

    def _unshard(self, u: _Unit, prefetch: Optional[_Unit]) -> None:
        self._issue_gather(u)
        if prefetch is not None:
            self._issue_gather(prefetch)
        self._wait_gather(u)

    # ------------------------------------------------------------------ hooks
    def _register_hooks(self) -> None:
        for u in self.units:
            u.module.register_forward_pre_hook(self._make_pre_forward(u))
            u.module.register_forward_hook(self._make_post_forward(u))
            u.module.register_full_backward_pre_hook(self._make_pre_backward(u))
        for u in self.exec_units:
            u.sink_written = [False] * len(u.params)
            for i, p in enumerate(u.params):
                p.register_post_accumulate_grad_hook(self._make_grad_hook(u, i))
                # (not for the root unit: its first gradient (LM head) and last one (embedding) bracket the whole
                # backward pass, so it would pin one of the two staging buffers while the blocks rotate through them)
                if self.native and self.world > 1 and p.dim() == 2 and p.dtype == torch.bfloat16 and u is not self.root_unit:
                    p._b200_grad_sink = self._make_sink(u, i)
                    p._b200_group_sink = self._make_group_sink(u, i)

    def _next_unit(self, u: _Unit, step: int) -> Optional[_Unit]:
        j = u.index + step
        if 0 <= j < len(self.exec_units):
            nxt = self.exec_units[j]
            return nxt if nxt is not self.root_unit else None
        return None

    def _make_pre_forward(self, u: _Unit):
        def hook(mod, args):
            # during backward this is a recomputation (activation checkpointing): the next unit needed is the PREVIOUS one
            self._unshard(u, self._next_unit(u, -1 if self._in_backward else +1))
        return hook

    def _make_post_forward(self, u: _Unit):
        def hook(mod, args, out):
            # the last block is needed again immediately by backward: keep it (standard FSDP trick)
            last = torch.is_grad_enabled() and self.units and u is self.units[-1]
            if self._in_backward:
                return   # recomputed forward: the unit's backward runs next, it is freed after its gradients are reduced
            if self.reshard_after_forward and not last:
                self._free(u)
        return hook

    def _make_pre_backward(self, u: _Unit):
        def hook(mod, grad_out):
            self._begin_backward()
            self._unshard(u, self._next_unit(u, -1))
        return hook

    def _begin_backward(self) -> None:
        if not self._in_backward:
            self._in_backward = True
            torch.autograd.Variable._execution_engine.queue_callback(self.finish_backward)

    def _make_grad_hook(self, u: _Unit, slot: int):
        def hook(p):
            self._begin_backward()
            if not u.grad_ready[slot]:
                u.grad_ready[slot] = True
                u.pending_grads -= 1
            if u.pending_grads == 0:
                self._reduce_unit(u)
        return hook

    # ------------------------------------------------------------------ gradient path
    def _claim_stage(self, u: _Unit):
        """Pick one of the two symmetric staging buffers for this unit's gradients (first use in a backward)."""
        if u.stage_k is None:
            k = self._stage_idx
            self._stage_idx ^= 1
            if self._stage_events[k] is not None:  # the reduce-scatter that last read this buffer must be done
                torch.cuda.current_stream(self.device).wait_event(self._stage_events[k])
            u.stage_k = k
        return self.grad_stage[u.stage_k]

    def _make_sink(self, u: _Unit, slot: int):
        """Weight-gradient sink: the wgrad GEMM writes dW directly into the unit's symmetric staging slot."""
        def sink(dy2d, x2d) -> bool:
            p = u.params[slot]
            if not self.require_backward_grad_sync or p.grad is not None:
                return False  # gradient accumulation (no_sync micro-steps, or their final step): autograd accumulates in p.grad
            self._begin_backward()
            o = u.offsets[slot]
            view = self._claim_stage(u).tensor(self.dtype, 0, u.total)[o:o + p.numel()].view(p.shape)
            N.gemm(dy2d, x2d, a_mn=True, b_mn=True, out=view, accumulate=u.sink_written[slot])
            u.sink_written[slot] = True
            if not u.grad_ready[slot]:
                u.grad_ready[slot] = True
                u.pending_grads -= 1
            if u.pending_grads == 0:
                self._reduce_unit(u)
            return True
        return sink

    def _make_group_sink(self, u: _Unit, slot0: int):
        """Sink for several row-stacked parameters that are adjacent in the unit layout (q|k|v, w1|w3): ONE wgrad GEMM
        writes [sum rows, K] straight into the staging buffer.  Declines (False) unless ``params`` are exactly the unit's
        parameters slot0, slot0 + 1, ... laid out back to back."""
        def sink(dy2d, x2d, params) -> bool:
            n = len(params)
            if slot0 + n > len(u.params) or any(u.params[slot0 + j] is not params[j] for j in range(n)):
                return False
            if not self.require_backward_grad_sync or any(p.grad is not None for p in params):
                return False
            if any(u.offsets[slot0 + j] + params[j].numel() != u.offsets[slot0 + j + 1] for j in range(n - 1)):
                return False
            if len({u.sink_written[slot0 + j] for j in range(n)}) != 1:
                return False
            self._begin_backward()
            o, K = u.offsets[slot0], params[0].shape[1]
            rows = sum(p.shape[0] for p in params)
            view = self._claim_stage(u).tensor(self.dtype, 0, u.total)[o:o + rows * K].view(rows, K)
            N.gemm(dy2d, x2d, a_mn=True, b_mn=True, out=view, accumulate=u.sink_written[slot0])
            for j in range(n):
                u.sink_written[slot0 + j] = True
                if not u.grad_ready[slot0 + j]:
                    u.grad_ready[slot0 + j] = True
                    u.pending_grads -= 1
            if u.pending_grads == 0:
                self._reduce_unit(u)
            return True
        return sink

    def _reduce_unit(self, u: _Unit) -> None:
        u.pending_grads = -1  # launched
        if not self.require_backward_grad_sync:
            if u is not self.root_unit and self.reshard_after_forward:
                self._free(u)   # accumulation micro-step: gradients stay in p.grad, the gathered weights can go
            return
        hp = self._opt.hp if self._opt is not None else None
        t = self._step_t + 1
        if self.world == 1:
            self._local_update(u, hp, t)
            return
        if self.native:
            stage = self._claim_stage(u)
            k = u.stage_k
            cur = torch.cuda.current_stream(self.device)
            st = stage.tensor(self.dtype, 0, u.total)
            for i, (p, o) in enumerate(zip(u.params, u.offsets)):
                if u.sink_written[i]:
                    continue  # the wgrad GEMM already wrote this slot
                seg = st[o:o + p.numel()]
                if p.grad is None:
                    seg.zero_()
                else:
                    seg.copy_(p.grad.reshape(-1))
                p.grad = None
            ev = torch.cuda.Event()
            ev.record(cur)
            with torch.cuda.stream(self.rs_stream):
                self.rs_stream.wait_event(ev)
                # the fused kernel writes a bf16 compute copy next to the fp32 master: bf16 models only.  split=True: this
                # stream runs beside compute that may itself wait for other GPUs (fused TP GEMMs, ring attention), so
                # the waits for the peers are one-warp kernels and the bulk kernel never holds SMs while waiting
                if (self.nvls_grads and self.fused_optimizer and hp is not None and hp["kind"] == "adamw"):
                    self._ensure_state(u)
                    self.comm.rs_adamw_nvls_from(stage, 0, self.dtype, u.shard_numel, 1.0 / self.world, u.master, u.shard,
                                                 u.m, u.v, lr=hp["lr"], beta1=hp["betas"][0], beta2=hp["betas"][1],
                                                 eps=hp["eps"], wd=hp["weight_decay"], step=t, channel=2, split=True)
                elif self.nvls_grads:
                    if u.grad_shard is None:
                        u.grad_shard = torch.zeros(u.shard_numel, dtype=torch.float32, device=self.device)
                        self.comm.reduce_scatter_nvls_from(u.grad_shard, stage, 0, self.dtype, 1.0 / self.world, channel=2, split=True)
                    else:
                        tmp = torch.empty_like(u.grad_shard)
                        self.comm.reduce_scatter_nvls_from(tmp, stage, 0, self.dtype, 1.0 / self.world, channel=2, split=True)
                        u.grad_shard.add_(tmp)
                elif self.fused_optimizer and hp is not None and hp["kind"] == "adamw" and self.dtype == torch.bfloat16:
                    self._ensure_state(u)
                    self.comm.rs_adamw_from(stage, 0, self.dtype, u.shard_numel, 1.0 / self.world, u.master, u.shard,
                                            u.m, u.v, lr=hp["lr"], beta1=hp["betas"][0], beta2=hp["betas"][1],
                                            eps=hp["eps"], wd=hp["weight_decay"], step=t, channel=2,
                                            copy_engine=self.copy_engine_reduce)
                else:
                    if u.grad_shard is None:
                        u.grad_shard = torch.zeros(u.shard_numel, dtype=torch.float32, device=self.device)
                        self.comm.reduce_scatter_from(u.grad_shard, stage, 0, self.dtype, 1.0 / self.world, channel=2)
                    else:
                        tmp = torch.empty_like(u.grad_shard)
                        self.comm.reduce_scatter_from(tmp, stage, 0, self.dtype, 1.0 / self.world, channel=2)
                        u.grad_shard.add_(tmp)
                done = torch.cuda.Event()
                done.record(self.rs_stream)
            self._stage_events[k] = done
            u.rs_event = done
        else:
            flat = torch.zeros(u.total, dtype=torch.float32, device=self.device)
            for p, o in zip(u.params, u.offsets):
                if p.grad is not None:
                    flat[o:o + p.numel()].copy_(p.grad.reshape(-1))
                p.grad = None
            g = torch.empty(u.shard_numel, dtype=torch.float32, device=self.device)
            dist.reduce_scatter_tensor(g, flat, group=self.comm.group)
            g.mul_(1.0 / self.world)
            u.grad_shard = g if u.grad_shard is None else u.grad_shard.add_(g)
        if u is not self.root_unit:
            self._free(u)

    def _ensure_state(self, u: _Unit) -> None:
        if u.m is None:
            u.m = torch.zeros(u.shard_numel, dtype=torch.float32, device=self.device)
            u.v = torch.zeros(u.shard_numel, dtype=torch.float32, device=self.device)

    def _local_update(self, u: _Unit, hp, t) -> None:
        """world == 1: pack grads and keep them for the optimizer step."""
        flat = torch.zeros(u.total, dtype=torch.float32, device=self.device)
        for p, o in zip(u.params, u.offsets):
            if p.grad is not None:
                flat[o:o + p.numel()].copy_(p.grad.reshape(-1))
            p.grad = None
        u.grad_shard = flat if u.grad_shard is None else u.grad_shard.add_(flat)

    def finish_backward(self) -> None:
        if not self._in_backward:
            return
        for u in self.exec_units:
            if u.pending_grads > 0 and any(u.grad_ready):  # some params of the unit got no gradient
                self._reduce_unit(u)
        if self.native and self.world > 1:
            cur = torch.cuda.current_stream(self.device)
            for u in self.exec_units:
                if u.rs_event is not None:
                    cur.wait_event(u.rs_event)
                    u.rs_event = None
        for u in self.exec_units:
            u.grad_ready = [False] * len(u.params)
            u.pending_grads = len(u.params)
            u.sink_written = [False] * len(u.params)
            u.stage_k = None
        if self.root_unit is not None and self.world > 1:
            self._free(self.root_unit)
        self._in_backward = False

    def clip_grad_norm_(self, max_norm: float, norm_type: float = 2.0) -> torch.Tensor:
        """Scale all gradients so that their GLOBAL norm (over every shard of every rank) is at most ``max_norm``; returns
        the norm before clipping.  Call between ``backward()`` and ``optimizer.step()``.

        Needs gradients that still exist at that point: with the fused reduce-scatter + AdamW path (bf16 parameters on
        several GPUs, ``fused_optimizer=True``) every unit has already been updated when backward ends, so there is
        nothing left to clip - construct ``FSDP(model, fused_optimizer=False)`` when you clip."""
        self.finish_backward()
        if self.native and self.world > 1 and self.fused_optimizer:
            raise RuntimeError("clip_grad_norm_: the fused reduce-scatter + AdamW path applies each unit's update during "
                               "backward; build the wrapper with FSDP(model, fused_optimizer=False) to clip gradients")
        shards = [u.grad_shard for u in self.exec_units if u.grad_shard is not None]
        if not shards:
            return torch.zeros((), dtype=torch.float32, device=self.device)
        if norm_type == float("inf"):
            total = torch.stack([g.detach().abs().max().float() for g in shards]).max()
            if self.world > 1:
                dist.all_reduce(total, op=dist.ReduceOp.MAX, group=self.comm.group)
        else:
            total = torch.stack([g.detach().float().abs().pow(norm_type).sum() for g in shards]).sum()
            if self.world > 1:          # shards are disjoint slices of the flat gradients (padding is zero)
                dist.all_reduce(total, group=self.comm.group)
            total = total.pow(1.0 / norm_type)
        coef = (float(max_norm) / (total + 1e-6)).clamp(max=1.0)
        for g in shards:
            g.mul_(coef.to(g.dtype))
        return total

    @contextlib.contextmanager
    def no_sync(self):
        """Gradient accumulation: backward passes inside the block keep their (unsharded) gradients in ``p.grad`` and skip
        the reduce-scatter / optimizer fusion; the first backward outside reduces the accumulated sum (the reference only
        describes accumulation, docs/guide/troubleshooting.md:70-83; torch FSDP offers the same context manager)."""
        prev = self.require_backward_grad_sync
        self.require_backward_grad_sync = False
        try:
            yield
        finally:
            self.require_backward_grad_sync = prev

    # ------------------------------------------------------------------ forward
    def forward(self, *args, **kwargs):
        self.finish_backward()
        if self.root_unit is not None:
            first = self.units[0] if self.units else None
            self._unshard(self.root_unit, first)
        out = self.module(*args, **kwargs)
        if not torch.is_grad_enabled() and self.root_unit is not None:
            self._free(self.root_unit)
        return out

    # ------------------------------------------------------------------ optimizer
    def build_optimizer(self, kind: str = "adamw", lr: float = 1e-3, betas=(0.9, 0.999), eps: float = 1e-8,
                        weight_decay: float = 0.01, momentum: float = 0.9) -> "ShardedOptimizer":
        self._opt = ShardedOptimizer(self, dict(kind=kind, lr=lr, betas=tuple(betas), eps=eps,
                                                weight_decay=weight_decay, momentum=momentum))
        return self._opt

    # ------------------------------------------------------------------ state dicts
    def _gather_full(self, u: _Unit) -> torch.Tensor:
        if self.world == 1:
            return u.full
        out = torch.empty(u.total, dtype=self.dtype, device=self.device)
        if self.native:
            self.comm.all_gather_from(out, u.shard_buf, 0, u.shard_numel * out.element_size(), channel=1)
        else:
            dist.all_gather_into_tensor(out, u.shard, group=self.comm.group)
        return out

    def full_state_dict(self, rank0_only: bool = True, offload_to_cpu: bool = True) -> Dict[str, torch.Tensor]:
        """Collective.  Unsharded parameters (+ buffers) under the ORIGINAL module's names, gathered
        unit by unit (peak extra memory = one unit), like FSDP FULL_STATE_DICT with
        offload_to_cpu / rank0_only (reference multinode_fsdp_unet.py:285-298)."""
        self.finish_backward()
        sd: Dict[str, torch.Tensor] = {}
        keep = (self.rank == 0) or not rank0_only
        for u in self.exec_units:
            flat = self._gather_full(u)
            if keep:
                for p, o, n in zip(u.params, u.offsets, u.names):
                    t = flat[o:o + p.numel()].view(p.shape).clone()
                    sd[n] = t.cpu() if offload_to_cpu else t
            del flat
        if keep:
            pnames = {n for u in self.exec_units for n in u.names}
            ordered = {}
            for k, v in self.module.state_dict().items():  # persistent buffers + canonical key order
                if k in pnames:
                    ordered[k] = sd[k]
                else:
                    ordered[k] = v.detach().cpu() if offload_to_cpu else v.detach().clone()
            sd = ordered
        return sd

    def load_full_state_dict(self, sd: Dict[str, torch.Tensor]) -> None:
        """Every rank passes the same full state dict; each keeps its own shard."""
        self.finish_backward()
        for u in self.exec_units:
            flat = torch.zeros(u.total, dtype=self.dtype, device=self.device)
            for p, o, n in zip(u.params, u.offsets, u.names):
                flat[o:o + p.numel()].copy_(sd[n].reshape(-1).to(self.device, self.dtype))
            lo = self.rank * u.shard_numel
            u.shard.copy_(flat[lo:lo + u.shard_numel])
            u.master = flat[lo:lo + u.shard_numel].to(torch.float32, copy=True)
            if self.world == 1:
                u.full.copy_(flat)
        for n, b in self.module.named_buffers():
            if n in sd and b.shape == sd[n].shape:
                b.copy_(sd[n].to(b.device))
        if self.native and self.world > 1:
            torch.cuda.synchronize(self.device)
            self.comm.barrier()

    def sharded_state_dict(self):
        out = {}
        for u in self.exec_units:
            lo = self.rank * u.shard_numel
            for p, o, n in zip(u.params, u.offsets, u.names):
                a, b = max(o, lo), min(o + p.numel(), lo + u.shard_numel)
                if a < b:
                    out[n] = {"tensor": u.master[a - lo:b - lo].to(self.dtype).clone(), "global_shape": list(p.shape),
                              "dim": None, "index": self.rank, "parts": self.world, "flat_range": [a - o, b - o]}
        if self.rank == 0:
            pnames = {n for u in self.exec_units for n in u.names}
            for n, b in self.module.state_dict().items():
                if n not in pnames:
                    out[n] = {"tensor": b.detach().clone(), "global_shape": list(b.shape), "dim": None, "index": 0,
                              "parts": 1, "flat_range": None}
        return out


class ShardedOptimizer:
    """Optimizer facade for FSDP: state is fp32 master + moments per local shard.

    With the fused path the update already happened inside the reduce-scatter kernels during
    backward; ``step()`` joins the streams and advances the step count.  Otherwise ``step()`` runs
    the fused AdamW / SGD kernel on each shard's fp32 gradient.
    """

    def __init__(self, fsdp: FSDP, hp: dict):
        self.fsdp, self.hp = fsdp, hp
        self.param_groups = [hp]

    def zero_grad(self, set_to_none: bool = True) -> None:
        for u in self.fsdp.exec_units:
            for p in u.params:
                p.grad = None

    def step(self) -> None:
        f = self.fsdp
        f.finish_backward()
        f._step_t += 1
        t = f._step_t
        hp = self.hp
        for u in f.exec_units:
            if u.grad_shard is None:
                continue  # fused path already applied the update
            f._ensure_state(u)
            g = u.grad_shard
            if f.replicate_comm is not None:   # HYBRID_SHARD: average this shard's gradient over the replicas
                f.replicate_comm.all_reduce(g, scale=1.0 / f.replicate_comm.world, channel=5)
            if f.world == 1:
                master, target = u.master, u.full
            else:
                master, target = u.master, u.shard
            if g.is_cuda:
                if hp["kind"] == "adamw":
                    N.adamw_step(master, g, u.m, u.v, lr=hp["lr"], beta1=hp["betas"][0], beta2=hp["betas"][1],
                                 eps=hp["eps"], wd=hp["weight_decay"], step=t,
                                 p_lowp=target if target.dtype == torch.bfloat16 else None)
                    if target.dtype != torch.bfloat16:
                        target.copy_(master)
                else:
                    N.sgd_step(master, g, u.m, lr=hp["lr"], momentum=hp["momentum"], wd=hp["weight_decay"],
                               first_step=(t == 1), p_lowp=target if target.dtype == torch.bfloat16 else None)
                    if target.dtype != torch.bfloat16:
                        target.copy_(master)
            else:
                if hp["kind"] == "adamw":
                    b1, b2 = hp["betas"]
                    u.m.mul_(b1).add_(g, alpha=1 - b1)
                    u.v.mul_(b2).addcmul_(g, g, value=1 - b2)
                    master.mul_(1 - hp["lr"] * hp["weight_decay"])
                    denom = (u.v.sqrt() / math.sqrt(1 - b2 ** t)).add_(hp["eps"])
                    master.addcdiv_(u.m, denom, value=-hp["lr"] / (1 - b1 ** t))
                else:
                    d = g + hp["weight_decay"] * master
                    if t == 1:
                        u.m.copy_(d)
                    else:
                        u.m.mul_(hp["momentum"]).add_(d)
                    master.add_(u.m, alpha=-hp["lr"])
                target.copy_(master.to(target.dtype))
            u.grad_shard = None

    def state_dict(self):
        """THIS rank's optimizer shards (the per-rank payload of a sharded checkpoint).  Tagged with rank / world so that
        it can only be loaded back into the same position; a single-file checkpoint uses ``full_state_dict`` instead."""
        f = self.fsdp
        return {"format": "shards", "rank": f.rank, "world": f.world, "step": f._step_t,
                "hp": {k: v for k, v in self.hp.items()},
                "shards": {u.name: {"master": u.master, "m": u.m, "v": u.v} for u in f.exec_units}}

    def full_state_dict(self, rank0_only: bool = True):
        """Collective.  fp32 master weights and AdamW moments gathered unit by unit into UNSHARDED flat tensors (CPU),
        keyed by unit name, independent of the world size that wrote them: what a consolidated single-file checkpoint
        stores, so that any rank of any world size can cut its own slice back out (``load_state_dict``)."""
        f = self.fsdp
        f.finish_backward()
        keep = (f.rank == 0) or not rank0_only
        units = {}
        for u in f.exec_units:
            f._ensure_state(u)
            entry = {"numel": u.offsets[-1] + u.params[-1].numel() if u.params else 0}
            for key, t in (("master", u.master), ("m", u.m), ("v", u.v)):
                if f.world == 1:
                    full = t
                else:
                    full = torch.empty(u.total, dtype=torch.float32, device=t.device)
                    dist.all_gather_into_tensor(full, t.contiguous(), group=f.comm.group)
                if keep:
                    entry[key] = full[:entry["numel"]].detach().cpu().clone()
            if keep:
                units[u.name] = entry
        if not keep:
            return {}
        return {"format": "full", "step": f._step_t, "hp": {k: v for k, v in self.hp.items()}, "units": units}

    def load_state_dict(self, sd) -> None:
        f = self.fsdp
        fmt = sd.get("format", "shards")
        if fmt == "full":
            for u in f.exec_units:
                e = sd["units"].get(u.name)
                if e is None:
                    continue
                numel = u.offsets[-1] + u.params[-1].numel() if u.params else 0
                if int(e["numel"]) != numel:
                    raise ValueError(f"optimizer state of unit {u.name}: {e['numel']} elements, the model has {numel}")
                lo, hi = f.rank * u.shard_numel, min((f.rank + 1) * u.shard_numel, numel)
                f._ensure_state(u)
                for key, dst in (("master", u.master), ("m", u.m), ("v", u.v)):
                    dst.zero_()
                    if hi > lo:
                        dst[:hi - lo].copy_(e[key][lo:hi].to(dst.device))
            f._step_t = int(sd.get("step", 0))
            return
        if int(sd.get("world", f.world)) != f.world or int(sd.get("rank", f.rank)) != f.rank:
            raise ValueError(f"per-rank optimizer shards written by rank {sd.get('rank')} of {sd.get('world')} cannot be "
                             f"loaded on rank {f.rank} of {f.world}; use a consolidated (full_state_dict) checkpoint")
        f._step_t = int(sd.get("step", 0))
        for u in f.exec_units:
            s = sd.get("shards", {}).get(u.name)
            if s is None:
                continue
            if s["master"].numel() != u.master.numel():
                raise ValueError(f"optimizer shard of unit {u.name} has {s['master'].numel()} elements, expected {u.master.numel()}")
            u.master.copy_(s["master"].to(u.master.device))
            if s.get("m") is not None:
                f._ensure_state(u)
                u.m.copy_(s["m"].to(u.m.device))
                if s.get("v") is not None and u.v is not None:
                    u.v.copy_(s["v"].to(u.v.device))
