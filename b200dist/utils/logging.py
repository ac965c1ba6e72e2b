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
