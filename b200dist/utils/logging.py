"""Rank-aware logging helpers (reference: /root/reference/utils/logging.py:19-65;
duplicates fsdp_tp/log_utils.py and scripts/06_hybrid_parallelism/log_utils.py are folded in)."""
from __future__ import annotations

import logging
import sys

import torch

_FMT = "%(asctime)s %(levelname)s %(name)s: %(message)s"
_configured = False


def _configure() -> None:
    global _configured
    if not _configured:
        logging.basicConfig(format=_FMT, level=logging.INFO, stream=sys.stdout)
        _configured = True


def get_logger(name: str | None = None) -> logging.Logger:
    _configure()
    return logging.getLogger(name or "b200dist")


def rank_log(rank: int, logger: logging.Logger, msg: str) -> None:
    """Log ``msg`` on rank 0 only."""
    if rank == 0:
        logger.info(" %s", msg)


def verify_min_gpu_count(min_gpus: int = 2) -> bool:
    """True when at least ``min_gpus`` CUDA devices are visible."""
    return torch.cuda.is_available() and torch.cuda.device_count() >= min_gpus


def log_memory(rank: int, logger: logging.Logger, tag: str = "") -> dict:
    """Device-memory line on rank 0 (allocated / reserved / peak, GiB) - the snippet the reference's hybrid README shows
    (scripts/06_hybrid_parallelism/README.md:285-290) as a helper.  Symmetric (peer / multicast) buffers are allocated by
    the native library and are not part of the caching allocator's numbers.  Returns the figures on every rank."""
    import torch

    if not torch.cuda.is_available():
        return {}
    gib = 1 << 30
    m = {"allocated_gib": torch.cuda.memory_allocated() / gib, "reserved_gib": torch.cuda.memory_reserved() / gib,
         "peak_gib": torch.cuda.max_memory_allocated() / gib}
    rank_log(rank, logger, f"memory{(' ' + tag) if tag else ''}: allocated {m['allocated_gib']:.2f} GiB, reserved "
                           f"{m['reserved_gib']:.2f} GiB, peak {m['peak_gib']:.2f} GiB")
    return m
