"""Data parallelism: bucketed gradient all-reduce overlapped with backward.

Behavioural twin of ``torch.nn.parallel.DistributedDataParallel`` as the reference uses it
(/root/reference/scripts/main.py:258-260; scripts/01_data_parallel_ddp/multinode_ddp_basic.py:148;
multinode_ddp_unet.py:309; distributed_dataloader.py:230-234 — all defaults):

* ctor: parameters and buffers are broadcast from rank 0 (one coalesced flat buffer each);
* parameters are assigned to buckets in REVERSE registration order, first bucket 1 MiB, the rest
  ``bucket_cap_mb`` (25 MiB);
* a post-accumulate-grad hook per parameter copies the gradient into its bucket slot; when the
  last slot of a bucket lands, the bucket's all-reduce is issued on a side stream, so it overlaps
  the rest of backward.  The 1/world scale (and an optional bf16->fp32 cast) is FUSED into the
  all-reduce kernel (``b200dist.comm`` two-shot / one-shot over NVLink peer memory) instead of
  being a separate divide pass;
* ``broadcast_buffers=True`` re-broadcasts floating-point buffers (BatchNorm running stats) from
  rank 0 before every forward, coalesced into one message;
* ``no_sync()`` for gradient accumulation, ``finish_backward()`` to join the side stream (the
  ``Trainer`` calls it; it is also called lazily from the next forward / optimizer hook).

On CPU tensors (gloo) the same code runs with torch.distributed collectives.
"""
from __future__ import annotations

import contextlib
from typing import List, Optional

import torch
import torch.distributed as dist
from torch import nn

from ..comm import get_communicator


class _Bucket:
    def __init__(self, params: List[nn.Parameter], dtype: torch.dtype, device: torch.device, index: int):
        self.params = params
        self.index = index
        self.offsets = []
        n = 0
        for p in params:
            self.offsets.append(n)
            n += (p.numel() + 7) // 8 * 8  # keep every slot 16/32-byte aligned
        self.numel = n = (n + 63) // 64 * 64   # whole 16-byte vectors per rank for any world <= 8 (NVLS slices)
        self.dtype = dtype
        self.mc_offset = None                  # byte offset inside the DDP multicast buffer when NVLS is in use
        self.flat = torch.zeros(n, dtype=dtype, device=device)
        self.views = [self.flat[o:o + p.numel()].view(p.shape) for o, p in zip(self.offsets, params)]
        self.pending = len(params)
        self.ready = [False] * len(params)
        self.event = None
        self.launched = False

    def rebase(self, flat: torch.Tensor, mc_offset: int) -> None:
        """Move the bucket into ``flat`` (a view of the multicast buffer's local copy)."""
        self.flat, self.mc_offset = flat, mc_offset
        self.views = [flat[o:o + p.numel()].view(p.shape) for o, p in zip(self.offsets, self.params)]

    def reset(self):
        self.pending = len(self.params)
        self.ready = [False] * len(self.params)
        self.event = None
        self.launched = False


class DistributedDataParallel(nn.Module):
    def __init__(self, module: nn.Module, device_ids=None, output_device=None, process_group=None,
                 bucket_cap_mb: float = 25.0, first_bucket_mb: float = 1.0, broadcast_buffers: bool = True,
                 comm_channel: int = 0):
        super().__init__()
        self.module = module
        self.process_group = process_group
        self.broadcast_buffers = broadcast_buffers
        self.comm = get_communicator(process_group)
        self.world = self.comm.world
        self.require_backward_grad_sync = True
        self._channel = comm_channel
        params = [p for p in module.parameters() if p.requires_grad]
        self._params = params
        self.device = params[0].device if params else torch.device("cpu")
        self._stream = torch.cuda.Stream(self.device) if self.device.type == "cuda" else None
        self._sync_module_states()
        self._build_buckets(params, bucket_cap_mb, first_bucket_mb)
        self._mc = None
        self._place_buckets_in_multicast_memory()
        self._hooks = []
        for b in self.buckets:
            for i, p in enumerate(b.params):
                self._hooks.append(p.register_post_accumulate_grad_hook(self._make_hook(b, i)))
        self._in_backward = False

    # ------------------------------------------------------------------ setup
    def _flat_broadcast(self, tensors: List[torch.Tensor]) -> None:
        if self.world == 1 or not tensors:
            return
        by_dtype = {}
        for t in tensors:
            by_dtype.setdefault(t.dtype, []).append(t)
        for dt, ts in by_dtype.items():
            flat = torch.cat([t.detach().reshape(-1) for t in ts])
            self.comm.broadcast(flat, src=0)
            off = 0
            for t in ts:
                t.detach().copy_(flat[off:off + t.numel()].view(t.shape))
                off += t.numel()

    def _sync_module_states(self) -> None:
        self._flat_broadcast([p for p in self.module.parameters()])
        self._flat_broadcast([b for b in self.module.buffers() if b.is_floating_point()])

    def _build_buckets(self, params, cap_mb, first_mb) -> None:
        self.buckets: List[_Bucket] = []
        cur, cur_bytes, cur_dtype = [], 0, None
        cap = int(first_mb * (1 << 20))
        for p in reversed(params):
            nbytes = p.numel() * p.element_size()
            if cur and (cur_bytes + nbytes > cap or p.dtype != cur_dtype):
                self.buckets.append(_Bucket(cur, cur_dtype, self.device, len(self.buckets)))
                cur, cur_bytes = [], 0
                cap = int(cap_mb * (1 << 20))
            cur.append(p)
            cur_bytes += nbytes
            cur_dtype = p.dtype
        if cur:
            self.buckets.append(_Bucket(cur, cur_dtype, self.device, len(self.buckets)))

    def _place_buckets_in_multicast_memory(self) -> None:
        """From ``nvls_min_world`` ranks up the buckets live inside ONE NVSwitch multicast buffer, so a bucket all-reduce
        is a single zero-copy NVLS kernel: ``multimem.ld_reduce`` of the rank's slice (summed in the switch), x 1/world,
        ``multimem.st`` to every replica.  Otherwise (fewer ranks, no multicast support, CPU)
        buckets are ordinary tensors reduced through the staging buffer by the two-shot pull kernel."""
        comm = self.comm
        if not (comm.native and self.device.type == "cuda" and comm.world >= comm.nvls_min_world and self.buckets):
            return
        if not comm._nvls_ready():
            return
        offs, total = [], 0
        for b in self.buckets:
            offs.append(total)
            total += (b.numel * b.flat.element_size() + 255) // 256 * 256
        self._mc = comm.nvls_buffer(total)
        if self._mc is None:
            return
        for b, o in zip(self.buckets, offs):
            b.rebase(self._mc.tensor(b.dtype, o, b.numel), o)

    def _reduce(self, bucket: _Bucket, scale: float) -> None:
        if bucket.mc_offset is not None:
            # split: the reducer stream runs beside backward kernels that may wait for other GPUs themselves (halo exchange)
            self.comm.all_reduce_nvls_inplace(self._mc, bucket.mc_offset, bucket.numel, bucket.dtype, scale, self._channel,
                                              split=True)
        else:
            self.comm.all_reduce(bucket.flat, scale=scale, channel=self._channel)

    # ------------------------------------------------------------------ backward hooks
    def _make_hook(self, bucket: _Bucket, slot: int):
        def hook(p: torch.Tensor):
            if not self.require_backward_grad_sync or self.world == 1:
                return
            g = p.grad
            view = bucket.views[slot]
            if g.data_ptr() != view.data_ptr():
                view.copy_(g)
                p.grad = view  # gradient_as_bucket_view: the optimizer reads the reduced bucket directly
            if not bucket.ready[slot]:
                bucket.ready[slot] = True
                bucket.pending -= 1
            if not self._in_backward:
                self._in_backward = True
                # join the side stream when this backward pass ends, even without a Trainer
                torch.autograd.Variable._execution_engine.queue_callback(self.finish_backward)
            if bucket.pending == 0 and not bucket.launched:
                self._launch(bucket)
        return hook

    def _launch(self, bucket: _Bucket) -> None:
        bucket.launched = True
        scale = 1.0 / self.world
        if self._stream is None:
            self._reduce(bucket, scale)
            return
        ev = torch.cuda.Event()
        ev.record(torch.cuda.current_stream(self.device))
        with torch.cuda.stream(self._stream):
            self._stream.wait_event(ev)
            self._reduce(bucket, scale)
            bucket.event = torch.cuda.Event()
            bucket.event.record(self._stream)

    def finish_backward(self) -> None:
        """Join the communication stream; flush buckets whose parameters produced no gradient."""
        if not self._in_backward:
            return
        for b in self.buckets:
            if not b.launched and self.require_backward_grad_sync and self.world > 1:
                for i, p in enumerate(b.params):
                    if not b.ready[i]:
                        b.views[i].zero_()
                        p.grad = b.views[i]
                self._launch(b)
        if self._stream is not None:
            cur = torch.cuda.current_stream(self.device)
            for b in self.buckets:
                if b.event is not None:
                    cur.wait_event(b.event)
        for b in self.buckets:
            b.reset()
        self._in_backward = False

    @contextlib.contextmanager
    def no_sync(self):
        prev = self.require_backward_grad_sync
        self.require_backward_grad_sync = False
        try:
            yield
        finally:
            self.require_backward_grad_sync = prev

    # ------------------------------------------------------------------ forward
    def forward(self, *args, **kwargs):
        self.finish_backward()
        if self.broadcast_buffers and self.world > 1 and self.training:
            bufs = [b for b in self.module.buffers() if b.is_floating_point()]
            if bufs:
                self._flat_broadcast(bufs)
        return self.module(*args, **kwargs)

    def state_dict(self, *args, **kwargs):
        # reference checkpoints strip the ``module.`` prefix (utils/checkpointing.py:45-49); keep
        # DDP-style prefixed keys here so ``main.py --resume`` style files also round-trip.
        return super().state_dict(*args, **kwargs)


DDP = DistributedDataParallel
