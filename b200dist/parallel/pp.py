"""Pipeline parallelism: stage wrapper + GPipe / 1F1B microbatch schedules over p2p send/recv.

Capability twin of what the reference does with ``torch.distributed.pipelining``
(/root/reference/scripts/04_pipeline_parallel_pp/02_pipeline_schedules.py:69-91,
03_pipeline_training.py:186-252): ``ScheduleGPipe`` / ``Schedule1F1B(stage, n_microbatches,
loss_fn)``; rank 0 calls ``step(inputs)``, the last rank ``step(target=..., losses=[])``, middle
ranks ``step()``.  Differences by design:

* no export tracing: a stage is any ``nn.Module``; helpers split ``nn.Sequential`` models and the
  Llama model by layer ranges (each rank only ever builds ITS stage, the reference builds the whole
  model on every rank, 03_pipeline_training.py:166-171);
* the schedule object is built once and reused every step (the reference rebuilds it per step, :236-240);
* activations/gradients travel as asynchronous NCCL p2p (``isend``/``irecv``) — by the north-star
  this path stays on NCCL — with the next receive posted before the current microbatch computes,
  so transfers overlap compute; gloo on CPU;
* PP x DP: pass ``dp_group`` and the stage's gradients are bucket-all-reduced across replicas
  (``b200dist.parallel.DDP``), synchronised only on each step's last backward.

``schedule_actions`` exposes the pure action list so the order can be unit-tested without GPUs.
"""
from __future__ import annotations

from typing import Callable, List, Optional, Sequence, Tuple

import torch
import torch.distributed as dist
from torch import nn

_DTYPES = [torch.float32, torch.bfloat16, torch.float16, torch.int64, torch.int32]


def schedule_actions(kind: str, stage: int, n_stages: int, n_microbatches: int) -> List[Tuple[str, int]]:
    """[("F", mb) | ("B", mb), ...] executed by ``stage`` for one training step."""
    M = n_microbatches
    if kind == "gpipe":
        return [("F", i) for i in range(M)] + [("B", i) for i in range(M)]
    if kind != "1f1b":
        raise ValueError(f"unknown schedule {kind!r}")
    warm = min(M, n_stages - 1 - stage)
    acts: List[Tuple[str, int]] = [("F", i) for i in range(warm)]
    f, b = warm, 0
    while f < M:  # steady state: one forward, one backward
        acts.append(("F", f)); f += 1
        acts.append(("B", b)); b += 1
    while b < M:  # cool-down
        acts.append(("B", b)); b += 1
    return acts


def bubble_fraction(n_stages: int, n_microbatches: int) -> float:
    """Idle fraction of a stage: (S-1)/(M+S-1)  (the reference's guide quotes (S-1)/M, bubble/ideal)."""
    return (n_stages - 1) / (n_microbatches + n_stages - 1)


class PipelineStage:
    def __init__(self, module: nn.Module, stage_index: int, num_stages: int, device=None, group=None,
                 dp_group=None):
        self.stage_index, self.num_stages = stage_index, num_stages
        self.group = group
        self.device = device if device is not None else next(module.parameters()).device
        self.is_first, self.is_last = stage_index == 0, stage_index == num_stages - 1
        ranks = dist.get_process_group_ranks(group) if (group is not None and dist.is_initialized()) else \
            list(range(dist.get_world_size() if dist.is_initialized() else 1))
        me = ranks.index(dist.get_rank()) if dist.is_initialized() else 0
        assert me == stage_index or num_stages == 1, "stage_index must equal this rank's position in the pp group"
        self.prev_rank = ranks[me - 1] if not self.is_first else None
        self.next_rank = ranks[me + 1] if not self.is_last else None
        if dp_group is not None and dist.get_world_size(dp_group) > 1:
            from .ddp import DDP

            module = DDP(module, process_group=dp_group)
        self.submod = module
        self._recv_meta = None   # (shape, dtype) of incoming activations
        self._out_meta = None

    def parameters(self):
        return self.submod.parameters()

    # shapes are exchanged once, on the first step
    def _send_meta(self, t: torch.Tensor) -> None:
        meta = torch.zeros(10, dtype=torch.int64)
        meta[0] = t.dim()
        meta[1:1 + t.dim()] = torch.tensor(t.shape)
        meta[9] = _DTYPES.index(t.dtype)
        dist.send(meta.to(self.device), self.next_rank, group=self.group)

    def _recv_meta_once(self):
        if self._recv_meta is None:
            meta = torch.zeros(10, dtype=torch.int64, device=self.device)
            dist.recv(meta, self.prev_rank, group=self.group)
            meta = meta.cpu()
            nd = int(meta[0])
            self._recv_meta = (tuple(int(x) for x in meta[1:1 + nd]), _DTYPES[int(meta[9])])
        return self._recv_meta


class _Schedule:
    kind = "gpipe"

    def __init__(self, stage: PipelineStage, n_microbatches: int, loss_fn: Optional[Callable] = None):
        self.stage, self.M, self.loss_fn = stage, n_microbatches, loss_fn
        self.actions = schedule_actions(self.kind, stage.stage_index, stage.num_stages, n_microbatches)

    def step(self, *args, target: Optional[torch.Tensor] = None, losses: Optional[list] = None):
        st, M = self.stage, self.M
        training = self.loss_fn is not None and torch.is_grad_enabled()
        mb_in = [None] * M
        if st.is_first:
            assert args, "the first stage needs the step inputs"
            chunks = [a.chunk(M) for a in args]
            mb_in = [tuple(c[i] for c in chunks) for i in range(M)]
        tgt = target.chunk(M) if (st.is_last and target is not None) else None
        saved_in: List[Optional[torch.Tensor]] = [None] * M
        saved_out: List[Optional[torch.Tensor]] = [None] * M
        outputs: List[Optional[torch.Tensor]] = [None] * M
        recv_act: dict = {}
        recv_grad: dict = {}
        pending_sends: list = []

        def post_recv_act(i):
            if st.is_first or i >= M or i in recv_act:
                return
            shape, dt = st._recv_meta_once()
            buf = torch.empty(shape, dtype=dt, device=st.device)
            recv_act[i] = (buf, dist.irecv(buf, st.prev_rank, group=st.group))

        def post_recv_grad(i):
            if st.is_last or i >= M or i in recv_grad or not training:
                return
            o = saved_out[i]
            buf = torch.empty_like(o)
            recv_grad[i] = (buf, dist.irecv(buf, st.next_rank, group=st.group))

        ddp = st.submod if hasattr(st.submod, "no_sync") else None
        n_back = sum(1 for a in self.actions if a[0] == "B")
        b_done = 0
        acts = self.actions if training else [a for a in self.actions if a[0] == "F"]
        for kind, i in acts:
            if kind == "F":
                if st.is_first:
                    x = mb_in[i]
                else:
                    post_recv_act(i)
                    buf, work = recv_act.pop(i)
                    work.wait()
                    post_recv_act(i + 1)  # prefetch the next activation while this microbatch computes
                    if training and buf.is_floating_point():
                        buf.requires_grad_(True)
                    saved_in[i] = buf
                    x = (buf,)
                if ddp is not None and training:
                    with ddp.no_sync():  # forward never syncs; buffers are broadcast once per step below
                        out = st.submod(*x)
                else:
                    out = st.submod(*x)
                if st.is_last:
                    outputs[i] = out
                    if training:
                        loss = self.loss_fn(out, tgt[i])
                        saved_out[i] = loss
                        if losses is not None:
                            losses.append(loss.detach())
                else:
                    if st._out_meta is None:
                        st._send_meta(out)
                        st._out_meta = True
                    saved_out[i] = out
                    pending_sends.append(dist.isend(out.detach(), st.next_rank, group=st.group))
            else:  # backward
                last_b = b_done == n_back - 1
                ctx = ddp.no_sync() if (ddp is not None and not last_b) else _null()
                with ctx:
                    if st.is_last:
                        (saved_out[i] / M).backward()
                    else:
                        post_recv_grad(i)
                        g, work = recv_grad.pop(i)
                        work.wait()
                        torch.autograd.backward(saved_out[i], g)
                b_done += 1
                saved_out[i] = None
                if not st.is_first:
                    gin = saved_in[i].grad
                    pending_sends.append(dist.isend(gin, st.prev_rank, group=st.group))
                    saved_in[i] = None
        for w in pending_sends:
            w.wait()
        if ddp is not None and training:
            ddp.finish_backward()
        if st.is_last:
            return torch.cat([o.detach() for o in outputs]) if all(o is not None for o in outputs) else None
        return None


class _null:
    def __enter__(self):
        return self

    def __exit__(self, *a):
        return False


class ScheduleGPipe(_Schedule):
    kind = "gpipe"


class Schedule1F1B(_Schedule):
    kind = "1f1b"


def split_sequential(model: nn.Sequential, num_stages: int, stage_index: int) -> nn.Sequential:
    """Contiguous, near-equal slice of an ``nn.Sequential`` for ``stage_index``."""
    n = len(model)
    per, rem = divmod(n, num_stages)
    start = stage_index * per + min(stage_index, rem)
    stop = start + per + (1 if stage_index < rem else 0)
    return nn.Sequential(*list(model.children())[start:stop])
