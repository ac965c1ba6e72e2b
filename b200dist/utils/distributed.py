"""Launcher detection and process-group bring-up (device-agnostic).

Same public tuple API as the reference helper so recipes port 1:1:
``init_distributed(backend, verbose) -> (world_rank, world_size, local_rank)`` and
``get_rank_info() -> (local_rank, world_size, world_rank, launcher)``
(reference: /root/reference/utils/distributed.py:26-177).

Differences, all deliberate:
* ONE implementation (the reference carries five private copies of this logic, e.g.
  scripts/main.py:28-60) and it never calls ``torch.cuda.set_device`` on a box without
  CUDA, so every entrypoint also runs on CPU/gloo.
* rendezvous defaults to 127.0.0.1 when nothing provides MASTER_ADDR (single NVSwitch
  box; container hostnames may not resolve).
* ``torch.distributed`` is only the CONTROL plane (rank discovery, handle exchange,
  gloo for CPU tests, NCCL p2p for pipeline stages).  The data plane for collectives
  is ``b200dist.comm`` (peer memory over NVLink).
"""
from __future__ import annotations

import datetime
import os
import socket
from typing import Optional, Tuple

import torch
import torch.distributed as dist

_LAUNCHER: Optional[str] = None


def _env_int(*names: str) -> Optional[int]:
    for n in names:
        v = os.environ.get(n)
        if v is not None and v != "":
            try:
                return int(v)
            except ValueError:
                continue
    return None


def _device_count() -> int:
    return torch.cuda.device_count() if torch.cuda.is_available() else 0


def get_rank_info() -> Tuple[int, int, int, str]:
    """Resolve ``(local_rank, world_size, world_rank, launcher)``.

    Priority (reference utils/distributed.py:41-100): torchrun env → Open MPI env →
    Cray PMI/PALS env → mpi4py → single process.
    """
    env = os.environ
    # 1. torchrun / torch.distributed.run
    if "LOCAL_RANK" in env and "RANK" in env:
        return int(env["LOCAL_RANK"]), int(env.get("WORLD_SIZE", "1")), int(env["RANK"]), "torchrun"
    # 2. Open MPI (mpirun / mpiexec)
    if "OMPI_COMM_WORLD_LOCAL_RANK" in env:
        local = int(env["OMPI_COMM_WORLD_LOCAL_RANK"])
        size = int(env["OMPI_COMM_WORLD_SIZE"])
        rank = int(env["OMPI_COMM_WORLD_RANK"])
        _set_master_addr_from_mpi(rank)
        return local, size, rank, "openmpi"
    # 3. Cray MPICH / PALS
    if "PMI_RANK" in env:
        rank = int(env["PMI_RANK"])
        size = int(env.get("PMI_SIZE", "1"))
        local = _env_int("PMI_LOCAL_RANK", "PALS_LOCAL_RANKID")
        if local is None:
            ngpu = _device_count()
            local = rank % ngpu if ngpu > 0 else 0
        _set_master_addr_from_mpi(rank)
        return local, size, rank, "cray-mpich"
    # 4. mpi4py (optional)
    if env.get("B200DIST_DISABLE_MPI4PY") != "1":
        try:
            from mpi4py import MPI  # type: ignore

            comm = MPI.COMM_WORLD
            size = comm.Get_size()
            if size > 1:
                rank = comm.Get_rank()
                local = comm.Split_type(MPI.COMM_TYPE_SHARED).Get_rank()
                _set_master_addr_from_mpi(rank)
                return local, size, rank, "mpi4py"
        except ImportError:
            pass
    # 5. single process
    return 0, 1, 0, "single"


def _own_address() -> str:
    try:
        return socket.gethostbyname(socket.gethostname())
    except OSError:
        return "127.0.0.1"


def _set_master_addr_from_mpi(world_rank: int) -> None:
    """Rank 0 publishes its address (MPI bcast when mpi4py exists), port defaults to 29500."""
    if "MASTER_ADDR" not in os.environ:
        addr = None
        try:
            from mpi4py import MPI  # type: ignore

            addr = _own_address() if world_rank == 0 else None
            addr = MPI.COMM_WORLD.bcast(addr, root=0)
        except ImportError:
            addr = _own_address()
        os.environ["MASTER_ADDR"] = addr or "127.0.0.1"
    os.environ.setdefault("MASTER_PORT", "29500")


def init_distributed(backend: str = "nccl", verbose: bool = True,
                     timeout_s: Optional[float] = None) -> Tuple[int, int, int]:
    """Bind the device, create the default process group, return (rank, world, local_rank).

    ``backend="nccl"`` silently becomes ``gloo`` when CUDA is absent, so the same recipe
    runs on a CPU box.  Initialisation is skipped for ``world_size == 1``
    (reference utils/distributed.py:144).
    """
    global _LAUNCHER
    local_rank, world_size, world_rank, launcher = get_rank_info()
    _LAUNCHER = launcher
    use_cuda = torch.cuda.is_available()
    if use_cuda:
        torch.cuda.set_device(local_rank % max(1, torch.cuda.device_count()))
    backend = os.environ.get("B200DIST_BACKEND", backend)   # lets test drivers repeat a script per backend
    if backend == "nccl" and not use_cuda:
        backend = "gloo"
    if world_size > 1 and not dist.is_initialized():
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        os.environ.setdefault("MASTER_PORT", "29500")
        kwargs = {}
        if timeout_s is not None:
            kwargs["timeout"] = datetime.timedelta(seconds=timeout_s)
        if backend == "nccl":
            # eager NCCL init on the bound device avoids the multi-second first-collective stall
            kwargs["device_id"] = torch.device("cuda", torch.cuda.current_device())
        try:
            dist.init_process_group(backend=backend, init_method="env://", world_size=world_size,
                                    rank=world_rank, **kwargs)
        except TypeError:
            kwargs.pop("device_id", None)
            dist.init_process_group(backend=backend, init_method="env://", world_size=world_size,
                                    rank=world_rank, **kwargs)
    if verbose and world_rank == 0:
        dev = torch.cuda.get_device_name() if use_cuda else "cpu"
        print(f"[b200dist] launcher={launcher} world_size={world_size} backend={backend} device={dev}",
              flush=True)
    return world_rank, world_size, local_rank


def cleanup_distributed() -> None:
    if dist.is_available() and dist.is_initialized():
        dist.destroy_process_group()


def get_rank() -> int:
    return dist.get_rank() if dist.is_available() and dist.is_initialized() else 0


def get_world_size() -> int:
    return dist.get_world_size() if dist.is_available() and dist.is_initialized() else 1


def is_main_rank() -> bool:
    return get_rank() == 0


def print_rank0(*args, **kwargs) -> None:
    if is_main_rank():
        kwargs.setdefault("flush", True)
        print(*args, **kwargs)


def launcher_name() -> str:
    return _LAUNCHER or get_rank_info()[3]


def default_device(local_rank: Optional[int] = None) -> torch.device:
    if torch.cuda.is_available():
        if local_rank is None:
            return torch.device("cuda", torch.cuda.current_device())
        return torch.device("cuda", local_rank % torch.cuda.device_count())
    return torch.device("cpu")


def barrier() -> None:
    if dist.is_available() and dist.is_initialized():
        if dist.get_backend() == "nccl":
            dist.barrier(device_ids=[torch.cuda.current_device()])
        else:
            dist.barrier()
