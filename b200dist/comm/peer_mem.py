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
        if offset_bytes < 0 or offset_bytes + numel * item > self.nbytes:
            raise ValueError(f"view of {numel} x {dtype} at byte {offset_bytes} exceeds the {self.nbytes}-byte buffer")
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


def multicast_supported() -> bool:
    """True when this GPU can join NVSwitch multicast teams (NVLS) and the library exposes the entry points."""
    if not (torch.cuda.is_available() and L.available()):
        return False
    v = C.c_int(0)
    try:
        return L.load().b200_mc_supported(C.byref(v)) == 0 and v.value == 1
    except AttributeError:
        return False


class MulticastBuffer:
    """Symmetric buffer behind an NVSwitch multicast object: ``tensor()`` views the local copy, ``mc_ptr`` is the
    multicast address that ``multimem.ld_reduce`` / ``multimem.st`` kernels use (csrc/multicast.cu).

    Collective over ``group``.  Rank 0 creates the object and exports it as a POSIX file descriptor; the descriptor
    travels to the other processes over an AF_UNIX socket (SCM_RIGHTS), the socket path over torch.distributed.  After
    every phase the ranks agree on success, so a refusal on ANY rank (not permitted, system not ready, ...) raises
    RuntimeError on ALL ranks instead of leaving the others blocked; callers then fall back to the pull kernels."""

    def __init__(self, nbytes: int, group=None, device: Optional[torch.device] = None):
        self.group = group
        self.rank = dist.get_rank(group)
        self.world = dist.get_world_size(group)
        self.device = device or torch.device("cuda", torch.cuda.current_device())
        self._state = None
        self._lib = L.load()
        with torch.cuda.device(self.device):
            handle = C.c_ulonglong(0)
            self._agree(self._phase_create(int(nbytes), handle), "create / share the multicast object")
            self._agree(self._try(lambda: L.check(self._lib.b200_mc_add_device(handle), "mc_add_device")), "add the device")
            lp, mp, st = C.c_void_p(), C.c_void_p(), C.c_void_p()
            # binding blocks until every device has been added: the agreement above doubles as that barrier
            self._agree(self._try(lambda: L.check(self._lib.b200_mc_bind_map(
                handle, C.c_size_t(self.nbytes), C.byref(lp), C.byref(mp), C.byref(st)), "mc_bind_map")), "bind and map memory")
        self.local_ptr, self.mc_ptr, self._state = lp.value, mp.value, st
        self._local = _wrap(self.local_ptr, self.nbytes, self.device)
        torch.cuda.synchronize(self.device)
        self._agree(None, "finish")

    @staticmethod
    def _try(fn) -> Optional[str]:
        try:
            fn()
            return None
        except Exception as e:  # noqa: BLE001
            return str(e)

    def _agree(self, err: Optional[str], what: str) -> None:
        flag = torch.tensor([0 if err is None else 1], device=self.device, dtype=torch.int32)
        dist.all_reduce(flag, op=dist.ReduceOp.MAX, group=self.group)
        if flag.item():
            raise RuntimeError(f"NVSwitch multicast unavailable (could not {what}): {err or 'refused on another rank'}")

    def _phase_create(self, nbytes: int, handle) -> Optional[str]:
        import os
        import socket
        import tempfile
        import time

        lib = self._lib
        path, err, fd = [None], None, C.c_int(-1)
        rounded = C.c_size_t(0)
        err = self._try(lambda: L.check(lib.b200_mc_granularity(C.c_size_t(nbytes), C.c_int(self.world), C.byref(rounded)),
                                        "mc_granularity"))
        self.nbytes = int(rounded.value) if err is None else nbytes
        srv = None
        if self.rank == 0 and err is None:
            err = self._try(lambda: L.check(lib.b200_mc_create(C.c_size_t(self.nbytes), C.c_int(self.world), C.byref(handle),
                                                               C.byref(fd)), "mc_create"))
            if err is None:
                path[0] = os.path.join(tempfile.gettempdir(),
                                       f"b200dist_mc_{os.getpid()}_{int(time.time() * 1e6) & 0xffffff}.sock")
                srv = socket.socket(socket.AF_UNIX, socket.SOCK_STREAM)
                srv.bind(path[0])
                srv.listen(self.world)
        src = dist.get_global_rank(self.group, 0) if self.group is not None else 0
        dist.broadcast_object_list(path, src=src, group=self.group)   # None = rank 0 could not create the object
        if path[0] is None:
            return err or "rank 0 could not create the multicast object"
        if self.rank == 0:
            for _ in range(self.world - 1):
                conn, _a = srv.accept()
                socket.send_fds(conn, [b"mc"], [fd.value])
                conn.close()
            srv.close()
            os.unlink(path[0])
            os.close(fd.value)
            return None
        if err is not None:   # still drain the connection so rank 0 does not block in accept()
            pass
        cli = socket.socket(socket.AF_UNIX, socket.SOCK_STREAM)
        cli.connect(path[0])
        _msg, fds, _flags, _addr = socket.recv_fds(cli, 16, 1)
        cli.close()
        if err is None:
            err = self._try(lambda: L.check(lib.b200_mc_import(C.c_int(fds[0]), C.byref(handle)), "mc_import"))
        os.close(fds[0])
        return err

    def tensor(self, dtype: torch.dtype = torch.uint8, offset_bytes: int = 0, numel: Optional[int] = None):
        item = torch.empty((), dtype=dtype).element_size()
        if numel is None:
            numel = (self.nbytes - offset_bytes) // item
        if offset_bytes < 0 or offset_bytes + numel * item > self.nbytes:
            raise ValueError(f"view of {numel} x {dtype} at byte {offset_bytes} exceeds the {self.nbytes}-byte buffer")
        return self._local[offset_bytes:offset_bytes + numel * item].view(dtype)

    def close(self):
        if self._state is not None:
            self._local = None
            L.load().b200_mc_release(self._state)
            self._state = None
