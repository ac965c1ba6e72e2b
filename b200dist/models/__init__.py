"""Model zoo (state-dict compatible with the reference / torchvision where a counterpart exists)."""
from .llama2 import ModelArgs, Transformer, model_args_for, PRESETS as LLAMA_PRESETS

__all__ = ["ModelArgs", "Transformer", "model_args_for", "LLAMA_PRESETS"]
