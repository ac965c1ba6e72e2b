"""Import-compatible location of the Llama-2 model (the reference keeps it at fsdp_tp/llama2_model.py with a byte-identical
copy under scripts/06_hybrid_parallelism/): ``from llama2_model import Transformer, ModelArgs`` keeps working.  The
implementation lives in b200dist.models.llama2 (same ModelArgs fields, parameter names and forward contract)."""
import sys
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from b200dist.models.llama2 import (Attention, FeedForward, ModelArgs, RMSNorm, Transformer,  # noqa: E402,F401
                                    TransformerBlock, TransformerStage)
from b200dist.ops import precompute_freqs_cis, rope  # noqa: E402,F401


def apply_rotary_emb(xq, xk, freqs_cis):
    """Reference signature (:74-100): rotate q and k ([B, S, H, hd], adjacent-pair convention) by ``freqs_cis``."""
    return rope(xq, freqs_cis), rope(xk, freqs_cis)
