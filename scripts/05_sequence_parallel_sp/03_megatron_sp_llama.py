#!/usr/bin/env python
"""Megatron-style sequence parallelism inside the TP group (the only executable SP of the reference, inside
01_fsdp_tp_hybrid.py:118,128,137): the hybrid recipe with dp = 1."""
import runpy
import sys
from pathlib import Path

if __name__ == "__main__":
    sys.argv += ["--fsdp_size", "1"] if "--fsdp_size" not in sys.argv else []
    runpy.run_path(str(Path(__file__).resolve().parent.parent / "06_hybrid_parallelism" / "01_fsdp_tp_hybrid.py"), run_name="__main__")
