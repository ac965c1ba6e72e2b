"""Hot ops: hand-written sm_100a kernels behind differentiable PyTorch functions."""
from .functional import (attention, attention_dropout, cross_entropy, embedding, gelu, layer_norm, linear, precompute_freqs_cis,
                         relu, rms_norm, rope, silu, swiglu)
from .optim import FusedAdamW, FusedSGD

__all__ = ["attention", "attention_dropout", "cross_entropy", "embedding", "gelu", "layer_norm", "linear", "precompute_freqs_cis",
           "relu", "rms_norm", "rope", "silu", "swiglu", "FusedAdamW", "FusedSGD"]
