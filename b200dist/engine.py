"""Training-step engine: the public call a user makes to run one optimizer step.

``Trainer.train_step(inputs, targets)`` does zero_grad → forward → loss → backward → (gradient
sync by the parallel wrapper) → optimizer step on device tensors; ``train_step_host`` is the
end-to-end variant that also stages the batch from pinned host memory on a copy stream and
returns the loss as a Python float (one D2H read), which is what ``bench.py`` reports as ``e2e``.

The hot loop of the reference recipes (e.g. /root/reference/scripts/main.py:334-339,
scripts/06_hybrid_parallelism/01_fsdp_tp_hybrid.py:171-180) has exactly this shape.
"""
from __future__ import annotations

import os
from typing import Callable, Optional

import torch

from . import ops


class InjectedFault(RuntimeError):
    """Raised by the fault-injection hook (``B200DIST_INJECT_FAULT=rank:step``)."""


def _fault_spec():
    """``B200DIST_INJECT_FAULT="rank:step"`` makes that rank fail right before that (0-based) optimizer step - a test
    hook for restart-from-checkpoint drills (the reference has no fault injection; its only recovery
    path is resuming from snapshot.pt, scripts/01_data_parallel_ddp/multinode_ddp_basic.py:144-155)."""
    spec = os.environ.get("B200DIST_INJECT_FAULT", "")
    if not spec:
        return None
    r, s = spec.split(":")
    return int(r), int(s)


def lm_loss(logits: torch.Tensor, targets: torch.Tensor) -> torch.Tensor:
    return ops.cross_entropy(logits, targets)


class Trainer:
    def __init__(self, model: torch.nn.Module, optimizer: torch.optim.Optimizer,
                 loss_fn: Callable[[torch.Tensor, torch.Tensor], torch.Tensor] = lm_loss,
                 device: Optional[torch.device] = None, cuda_graph: bool = False, graph_warmup: int = 3):
        """``cuda_graph=True`` (single-GPU models, static shapes): after ``graph_warmup`` eager steps the whole step -
        zero_grad, forward, loss, backward, optimizer - is captured into ONE CUDA graph and replayed from then on, which
        removes the per-kernel launch cost of launch-bound models (several hundred launches per step for a ResNet).
        Needs an optimizer whose step-dependent scalars live on the device (``FusedAdamW(capturable=True)``; FusedSGD bakes
        its learning rate into the graph).  Wrappers that drive side streams from autograd hooks (DDP / FSDP) are not
        captured: the flag is ignored for them.  Checked by ``tools/cuda_graph_check.py`` on a B200: loss sequence identical to the
        eager step, 4.71 -> 1.13 ms per step on a 4-layer, dim-512 Llama (``profiles/cuda_graph_check_r2.log``)."""
        self.model, self.optimizer, self.loss_fn = model, optimizer, loss_fn
        self._graph_wanted = bool(cuda_graph) and not callable(getattr(model, "finish_backward", None))
        self._graph_warmup = max(1, int(graph_warmup))
        self._graph = None
        self._static = None
        self.device = device or next(model.parameters()).device
        self._copy_stream = torch.cuda.Stream(self.device) if self.device.type == "cuda" else None
        self._staged = None
        self.steps_done = 0
        self._fault = _fault_spec()
        self._rank = int(os.environ.get("RANK", "0"))

    # ---------------------------------------------------------------- device-resident step
    def train_step(self, inputs: torch.Tensor, targets: torch.Tensor) -> torch.Tensor:
        if self._fault is not None and self._fault == (self._rank, self.steps_done):
            raise InjectedFault(f"injected fault on rank {self._rank} before step {self.steps_done}")
        if self._graph_wanted and inputs.is_cuda:
            return self._graphed_step(inputs, targets)
        return self._eager_step(inputs, targets)

    def _graphed_step(self, inputs: torch.Tensor, targets: torch.Tensor) -> torch.Tensor:
        opt = self.optimizer
        if self._graph is None:
            if self.steps_done < self._graph_warmup:
                return self._eager_step(inputs, targets)      # allocator, lazy state and func attributes settle here
            self._static = (torch.empty_like(inputs), torch.empty_like(targets))
            self._static[0].copy_(inputs); self._static[1].copy_(targets)
            prepare = getattr(opt, "prepare_step", None)
            if callable(prepare):
                opt._external_prepare = True
                prepare()
            torch.cuda.synchronize(self.device)
            self._graph = torch.cuda.CUDAGraph()
            with torch.cuda.graph(self._graph):
                self._static_loss = self._eager_step(self._static[0], self._static[1], count=False)
            # capture records the work without running it: replay once so this call is a real step
            self._graph.replay()
            self.steps_done += 1
            return self._static_loss
        if inputs.shape != self._static[0].shape or targets.shape != self._static[1].shape:
            raise ValueError("cuda_graph=True needs static shapes: got "
                             f"{tuple(inputs.shape)} after capturing {tuple(self._static[0].shape)}")
        self._static[0].copy_(inputs, non_blocking=True)
        self._static[1].copy_(targets, non_blocking=True)
        prepare = getattr(opt, "prepare_step", None)
        if callable(prepare):
            prepare()
        self._graph.replay()
        self.steps_done += 1
        return self._static_loss

    def _eager_step(self, inputs: torch.Tensor, targets: torch.Tensor, count: bool = True) -> torch.Tensor:
        self.optimizer.zero_grad(set_to_none=True)
        out = self.model(inputs)
        loss = self.loss_fn(out, targets)
        loss.backward()
        finish = getattr(self.model, "finish_backward", None)
        if callable(finish):
            finish()  # parallel wrappers: wait for gradient collectives issued from hooks
        self.optimizer.step()
        if count:
            self.steps_done += 1
        return loss.detach()

    # ---------------------------------------------------------------- host -> device -> host step
    def stage(self, inputs_host: torch.Tensor, targets_host: torch.Tensor):
        """Start the H2D copy of the next batch on the copy stream (inputs must be pinned)."""
        if self._copy_stream is None:
            self._staged = (inputs_host, targets_host, None)
            return
        with torch.cuda.stream(self._copy_stream):
            x = inputs_host.to(self.device, non_blocking=True)
            y = targets_host.to(self.device, non_blocking=True)
            ev = torch.cuda.Event()
            ev.record(self._copy_stream)
        self._staged = (x, y, ev)

    def train_step_host(self, inputs_host: torch.Tensor, targets_host: torch.Tensor) -> float:
        if self._staged is None:
            self.stage(inputs_host, targets_host)
        x, y, ev = self._staged
        self._staged = None
        if ev is not None:
            torch.cuda.current_stream(self.device).wait_event(ev)
            x.record_stream(torch.cuda.current_stream(self.device))
            y.record_stream(torch.cuda.current_stream(self.device))
        loss = self.train_step(x, y)
        return float(loss.item())  # D2H read of the step result
