#!/usr/bin/env python
"""Ring / context-parallel attention demo (reference: docs/guide/08_sequence_parallel.md:82-143, pseudo-code only)."""
import importlib.util
import sys
from pathlib import Path

if __name__ == "__main__":
    spec = importlib.util.spec_from_file_location("ulysses_demo", Path(__file__).with_name("01_ulysses_attention.py"))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    mod.main("ring")
