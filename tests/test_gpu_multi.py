"""Multi-GPU tests (need >= 2 B200s): peer-memory collectives vs NCCL, FSDP / DDP / TP parity."""
import os
import subprocess
import sys
from pathlib import Path

import pytest

pytestmark = [pytest.mark.gpu, pytest.mark.multigpu]
ROOT = Path(__file__).resolve().parent.parent


def _torchrun(script, nproc=2, args=(), timeout=600):
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={nproc}",
           "--master-addr", "127.0.0.1", "--master-port", "29561", str(ROOT / script), *args]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=timeout, cwd=str(ROOT))
    assert r.returncode == 0, r.stdout[-4000:] + "\n" + r.stderr[-4000:]
    return r.stdout


def test_collectives_match_nccl():
    out = _torchrun("tools/comm_check.py")
    assert "total failures 0" in out


def test_parallel_strategies_match_single_gpu():
    out = _torchrun("tools/parallel_check.py")
    assert "ALL OK" in out


def test_zigzag_ring_and_fused_rowwise_linear():
    out = _torchrun("tools/microbench/zigzag_linear_ar_check.py")
    assert "failures 0" in out


def test_fused_gemms_with_ragged_widths():
    out = _torchrun("tools/microbench/ag_gemm_ragged_check.py")
    assert "failures 0" in out
