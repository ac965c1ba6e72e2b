"""Debug aids for the peer-memory data plane (the reference has no counterpart - its only
debugging advice is NCCL_DEBUG, docs/guide/troubleshooting.md).

* ``CollectiveOrderChecker`` - collective-ORDER hash.  The device collectives match each other by (channel, sequence
  number) only; if one rank skips or reorders a call the symptom is a watchdog trap much later.  With
  ``B200DIST_DEBUG_COLLECTIVES=1`` every ``Communicator`` call is folded into a per-channel digest (op, bytes) and
  ``Communicator.barrier()`` compares the digests of all ranks on the host, reporting the first diverging call.
* ``sync_debug()`` / ``B200DIST_DEBUG_SYNC=1`` - serialising mode: every native call is followed by a stream
  synchronize and an error check, so an asynchronous fault is attributed to the op that caused it, and a bug that
  disappears in this mode is a missing stream dependency (a race).
* ``StreamHazardTracker`` - host-side race detector for buffers shared between streams (staging buffers, FSDP flat
  buffers): declare reads and writes with the stream they were issued on; an access that is not ordered after the previous
  conflicting access by an event the tracker has seen is reported.
* ``assert_replicas_equal`` - bitwise comparison of a tensor across a group (DDP replicas, TP-replicated parameters).
"""
from __future__ import annotations

import contextlib
import hashlib
import os
from typing import Dict, List, Optional, Tuple

import torch
import torch.distributed as dist


def collectives_debug_enabled() -> bool:
    return os.environ.get("B200DIST_DEBUG_COLLECTIVES", "0") not in ("", "0")


class CollectiveOrderError(RuntimeError):
    pass


class CollectiveOrderChecker:
    """Rolling digest of the collective call sequence per channel, plus a short tail of readable records."""

    def __init__(self, keep: int = 32):
        self.keep = keep
        self.digest: Dict[int, bytes] = {}
        self.count: Dict[int, int] = {}
        self.tail: Dict[int, List[str]] = {}

    def record(self, channel: int, op: str, nbytes: int) -> None:
        rec = f"{op}:{nbytes}"
        h = hashlib.blake2b(self.digest.get(channel, b""), digest_size=8)
        h.update(rec.encode())
        self.digest[channel] = h.digest()
        self.count[channel] = self.count.get(channel, 0) + 1
        t = self.tail.setdefault(channel, [])
        t.append(f"#{self.count[channel]} {rec}")
        if len(t) > self.keep:
            del t[0]

    def state(self) -> Dict[int, Tuple[int, str]]:
        return {c: (self.count[c], self.digest[c].hex()) for c in self.digest}

    def verify(self, group=None) -> None:
        """Host-side comparison across ``group`` (collective).  Raises CollectiveOrderError on divergence."""
        if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size(group) == 1:
            return
        mine = (self.state(), {c: list(t) for c, t in self.tail.items()})
        everyone: list = [None] * dist.get_world_size(group)
        dist.all_gather_object(everyone, mine, group=group)
        ref_state, ref_tail = everyone[0]
        for r, (st, tail) in enumerate(everyone):
            if st == ref_state:
                continue
            for c in sorted(set(st) | set(ref_state)):
                if st.get(c) != ref_state.get(c):
                    a, b = ref_tail.get(c, []), tail.get(c, [])
                    first = next((i for i, (x, y) in enumerate(zip(a, b)) if x != y), min(len(a), len(b)))
                    raise CollectiveOrderError(
                        f"collective order diverged on channel {c}: rank 0 issued {ref_state.get(c, (0,))[0]} calls, rank {r} "
                        f"issued {st.get(c, (0,))[0]}; first differing record near the tail: rank0={a[first:first + 3]} "
                        f"rank{r}={b[first:first + 3]}")


_SYNC = os.environ.get("B200DIST_DEBUG_SYNC", "0") not in ("", "0")


def sync_debug_enabled() -> bool:
    return _SYNC


@contextlib.contextmanager
def sync_debug(enabled: bool = True):
    """Serialise every native call inside the block (stream synchronize + error check after each)."""
    global _SYNC
    prev, _SYNC = _SYNC, enabled
    try:
        yield
    finally:
        _SYNC = prev


def after_native_call(name: str) -> None:
    """Called by ops._lib.call when sync debugging is on."""
    if torch.cuda.is_available():
        try:
            torch.cuda.current_stream().synchronize()
        except RuntimeError as e:  # attribute the asynchronous fault to this op
            raise RuntimeError(f"asynchronous CUDA error surfaced right after b200_{name}: {e}") from e


class StreamHazardError(RuntimeError):
    pass


class StreamHazardTracker:
    """Happens-before bookkeeping for buffers touched from several streams.

    Usage: ``t.write(buf, stream)`` / ``t.read(buf, stream)`` when work touching ``buf`` is enqueued,
    ``t.record(event_id, stream)`` / ``t.wait(event_id, stream)`` next to the real event calls.  Streams are any
    hashable ids.  A conflicting access (write-after-read, read-after-write, write-after-write) on a different stream
    that is not ordered by a recorded->waited event raises StreamHazardError.  Pure host logic: usable in CPU tests.
    """

    def __init__(self):
        self._clock: Dict[object, Dict[object, int]] = {}    # stream -> vector clock
        self._events: Dict[object, Dict[object, int]] = {}   # event -> snapshot of the recording stream's clock
        self._last_write: Dict[object, Tuple[object, int]] = {}
        self._last_reads: Dict[object, List[Tuple[object, int]]] = {}

    def _tick(self, stream) -> int:
        vc = self._clock.setdefault(stream, {})
        vc[stream] = vc.get(stream, 0) + 1
        return vc[stream]

    def _ordered(self, access: Tuple[object, int], stream) -> bool:
        s, t = access
        return s == stream or self._clock.get(stream, {}).get(s, 0) >= t

    def record(self, event, stream) -> None:
        self._tick(stream)
        self._events[event] = dict(self._clock[stream])

    def wait(self, event, stream) -> None:
        snap = self._events.get(event)
        if snap is None:
            raise StreamHazardError(f"stream {stream!r} waits on event {event!r} that was never recorded")
        vc = self._clock.setdefault(stream, {})
        for s, t in snap.items():
            vc[s] = max(vc.get(s, 0), t)

    def read(self, buf, stream, what: str = "") -> None:
        w = self._last_write.get(buf)
        if w is not None and not self._ordered(w, stream):
            raise StreamHazardError(f"read-after-write race on {buf!r} {what}: written on stream {w[0]!r}, read on "
                                    f"{stream!r} without an event dependency")
        self._last_reads.setdefault(buf, []).append((stream, self._tick(stream)))

    def write(self, buf, stream, what: str = "") -> None:
        w = self._last_write.get(buf)
        if w is not None and not self._ordered(w, stream):
            raise StreamHazardError(f"write-after-write race on {buf!r} {what}: streams {w[0]!r} and {stream!r}")
        for r in self._last_reads.get(buf, []):
            if not self._ordered(r, stream):
                raise StreamHazardError(f"write-after-read race on {buf!r} {what}: still being read on stream {r[0]!r} "
                                        f"when stream {stream!r} overwrites it")
        self._last_write[buf] = (stream, self._tick(stream))
        self._last_reads[buf] = []


def assert_replicas_equal(t: torch.Tensor, group=None, name: str = "tensor") -> None:
    """Raise if ``t`` is not bitwise identical on every rank of ``group`` (collective)."""
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size(group) == 1:
        return
    flat = t.detach().reshape(-1).contiguous()
    probe = flat.view(torch.uint8) if flat.dtype != torch.bool else flat.to(torch.uint8)
    # order-sensitive 64-bit checksum, cheap enough for every parameter
    idx = torch.arange(1, probe.numel() + 1, device=probe.device, dtype=torch.int64)
    digest = torch.stack([(probe.to(torch.int64) * (idx % 65521 + 1)).sum(), probe.to(torch.int64).sum()])
    lo, hi = digest.clone(), digest.clone()
    dist.all_reduce(lo, op=dist.ReduceOp.MIN, group=group)
    dist.all_reduce(hi, op=dist.ReduceOp.MAX, group=group)
    if not torch.equal(lo, hi):
        raise AssertionError(f"{name} differs between the ranks of the group (checksum min {lo.tolist()} max {hi.tolist()})")
