"""Shared runtime utilities (same public names as the reference ``utils`` package,
/root/reference/utils/__init__.py:9-16)."""
from . import debug
from .checkpointing import (consolidate_sharded_checkpoint, load_checkpoint, load_sharded_checkpoint, load_snapshot,
                            save_checkpoint, save_sharded_checkpoint, save_snapshot)
from .config import TrainingConfig
from .distributed import (cleanup_distributed, get_rank_info, init_distributed, is_main_rank, print_rank0)
from .logging import get_logger, log_memory, rank_log, verify_min_gpu_count
from .profiling import print_profiler_summary, training_profiler
from .redirect import redirect
from .sampler import DistributedSampler

__all__ = ["get_rank_info", "init_distributed", "cleanup_distributed", "is_main_rank", "print_rank0", "get_logger",
           "rank_log", "log_memory", "verify_min_gpu_count", "save_checkpoint", "load_checkpoint", "save_sharded_checkpoint",
           "load_sharded_checkpoint", "consolidate_sharded_checkpoint", "save_snapshot", "load_snapshot",
           "debug", "TrainingConfig", "training_profiler", "print_profiler_summary", "redirect", "DistributedSampler"]
