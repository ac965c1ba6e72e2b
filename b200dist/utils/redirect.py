"""Per-rank stdout/stderr capture, including output written by native code
(reference: /root/reference/utils/redirect.py:5-38)."""
from __future__ import annotations

import contextlib
import os
import sys


@contextlib.contextmanager
def redirect(*, log_prefix: str = "std", rank_env: str = "RANK", path: str = "."):
    """Send this process's fds 1/2 to ``{path}/{log_prefix}.{rank}.out|err``.

    ``os.dup2`` rebinds the OS-level descriptors, so text printed by NCCL, the CUDA
    driver or our own kernels' ``printf`` lands in the same files as Python prints.
    """
    rank = os.environ.get(rank_env, "0")
    os.makedirs(path, exist_ok=True)
    out_path = os.path.join(path, f"{log_prefix}.{rank}.out")
    err_path = os.path.join(path, f"{log_prefix}.{rank}.err")
    sys.stdout.flush()
    sys.stderr.flush()
    saved_out_fd, saved_err_fd = os.dup(1), os.dup(2)
    saved_out, saved_err = sys.stdout, sys.stderr
    fout = open(out_path, "w", buffering=1)
    ferr = open(err_path, "w", buffering=1)
    try:
        os.dup2(fout.fileno(), 1)
        os.dup2(ferr.fileno(), 2)
        sys.stdout, sys.stderr = fout, ferr
        yield out_path, err_path
    finally:
        try:
            sys.stdout.flush()
            sys.stderr.flush()
        finally:
            os.dup2(saved_out_fd, 1)
            os.dup2(saved_err_fd, 2)
            os.close(saved_out_fd)
            os.close(saved_err_fd)
            sys.stdout, sys.stderr = saved_out, saved_err
            fout.close()
            ferr.close()
