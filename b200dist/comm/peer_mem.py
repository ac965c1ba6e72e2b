"""Symmetric peer memory: one device buffer per rank, mapped into every rank of the group.

Control plane = ``torch.distributed`` (handle exchange with ``all_gather_object``); data plane =
CUDA IPC mappings of plain ``cudaMalloc`` memory, reachable over NVLink 5 / NVSwitch from inside
our kernels with ordinary loads and stores.  This plays the role NCCL's internal buffer
registration plays for the reference.
"""
from __future__ import annotations

import ctypes as C
from typing import List, Optional

import torch
import torch.distributed as dist

from ..ops import _lib as L


class _RawCuda:
    def __init__(self, ptr: int, nbytes: int):
        self.__cuda_array_interface__ = {"shape": (nbytes,), "typestr": "|u1", "data": (ptr, False),
                                         "version": 3, "strides": None}


def _wrap(ptr: int, nbytes: int, device: torch.device) -> torch.Tensor:
    return torch.as_tensor(_RawCuda(ptr, nbytes), device=device)


class PeerBuffer:
    """``nbytes`` of zero-initialised device memory on every rank of ``group``; ``ptrs[r]`` is the
    address at which rank r's buffer is mapped in THIS process."""

    def __init__(self, nbytes: int, group=None, device: Optional[torch.device] = None):
        self.group = group
        self.rank = dist.get_rank(group) if dist.is_initialized() else 0
        self.world = dist.get_world_size(group) if dist.is_initialized() else 1
        self.device = device or torch.device("cuda", torch.cuda.current_device())
        self.nbytes = (int(nbytes) + 255) // 256 * 256
        lib = L.load()
        p = C.c_void_p()
        with torch.cuda.device(self.device):
            L.check(lib.b200_peer_alloc(C.c_size_t(self.nbytes), C.byref(p)), "peer_alloc")
        self.local_ptr = p.value
        self.ptrs: List[int] = [0] * self.world
        self.ptrs[self.rank] = self.local_ptr
        self._imported: List[int] = []
        if self.world > 1:
            h = (C.c_ubyte * 64)()
            L.check(lib.b200_peer_export(C.c_void_p(self.local_ptr), h), "peer_export")
            handles = [None] * self.world
            dist.all_gather_object(handles, bytes(h), group=group)
            with torch.cuda.device(self.device):
                for r, hb in enumerate(handles):
                    if r == self.rank:
                        continue
                    q = C.c_void_p()
                    buf = (C.c_ubyte * 64).from_buffer_copy(hb)
                    L.check(lib.b200_peer_import(buf, C.byref(q)), "peer_import")
                    self.ptrs[r] = q.value
                    self._imported.append(q.value)
        self.ptr_array = (C.c_void_p * self.world)(*self.ptrs)
        self._local = _wrap(self.local_ptr, self.nbytes, self.device)

    def tensor(self, dtype: torch.dtype = torch.uint8, offset_bytes: int = 0, numel: Optional[int] = None):
        """Typed view of the LOCAL buffer."""
        item = torch.empty((), dtype=dtype).element_size()
        if numel is None:
            numel = (self.nbytes - offset_bytes) // item
        return self._local[offset_bytes:offset_bytes + numel * item].view(dtype)

    def close(self):
        lib = L.load()
        for q in self._imported:
            lib.b200_peer_close(C.c_void_p(q))
        self._imported = []
        if self.local_ptr:
            self._local = None
            lib.b200_peer_free(C.c_void_p(self.local_ptr))
            self.local_ptr = 0
