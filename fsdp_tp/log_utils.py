"""rank_log / get_logger / verify_min_gpu_count under the module name the reference examples import
(fsdp_tp/log_utils.py, duplicated as scripts/06_hybrid_parallelism/log_utils.py)."""
import sys
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from b200dist.utils.logging import get_logger, rank_log, verify_min_gpu_count  # noqa: E402,F401
