#!/usr/bin/env python
"""FSDP U-Net on ERA5-like data with a gathered full-state-dict checkpoint
(/root/reference/scripts/02_fully_sharded_fsdp/multinode_fsdp_unet.py:202-213,285-298): every rank enters the gather,
tensors are offloaded to CPU, rank 0 writes the consolidated file."""
import runpy
import sys
from pathlib import Path

if __name__ == "__main__":
    # same training loop as the DDP recipe with --strategy fsdp; then the FULL_STATE_DICT-style save
    target = Path(__file__).resolve().parent.parent / "01_data_parallel_ddp" / "multinode_ddp_unet.py"
    if "--strategy" not in sys.argv:
        sys.argv += ["--strategy", "fsdp"]
    if "--save_every" not in sys.argv:
        sys.argv += ["--save_every", "1"]
    runpy.run_path(str(target), run_name="__main__")
