#!/usr/bin/env python
"""Load-balanced causal ring attention: every rank holds blocks r and 2P-1-r of the sequence (zig-zag) and runs the same
2P+1 block pairs (plain ring attention gives rank P-1 P chunks of work and rank 0 one).  Not in the reference."""
import importlib.util
from pathlib import Path

if __name__ == "__main__":
    spec = importlib.util.spec_from_file_location("ulysses_demo", Path(__file__).with_name("01_ulysses_attention.py"))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    mod.main("zigzag")
