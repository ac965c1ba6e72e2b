#!/usr/bin/env python
"""2-D parallel Llama example (TP + SP inside a group of ranks, FSDP across groups) - the older, mpi4py-launched twin of
scripts/06_hybrid_parallelism/01_fsdp_tp_hybrid.py in the reference (fsdp_tp/fsdp_tp_example.py).  Launcher detection here
is the shared one (torchrun / Open MPI / PMI / mpi4py), so this file simply runs that recipe:
    mpiexec -n 8 python fsdp_tp/fsdp_tp_example.py --tp 4        or       scripts/launch.sh 8 fsdp_tp/fsdp_tp_example.py --tp 4"""
import runpy
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT / "scripts"))
runpy.run_path(str(ROOT / "scripts" / "06_hybrid_parallelism" / "01_fsdp_tp_hybrid.py"), run_name="__main__")
