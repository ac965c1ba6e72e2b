"""Model zoo (state-dict compatible with the reference / torchvision where a counterpart exists)."""
from .llama2 import ModelArgs, Transformer, TransformerStage, model_args_for, PRESETS as LLAMA_PRESETS
from .pp_transformer import PipelineTransformer
from .resnet import build_resnet, resnet18, resnet50, resnet101, resnet152
from .toy import FourBlockMLP, SimpleModel, StageModule, ToyMLP, ToyModel, regression_head
from .unet import SimpleUNet, latitude_weighted_mse
from .vit import SimpleViT

__all__ = ["ModelArgs", "Transformer", "TransformerStage", "model_args_for", "LLAMA_PRESETS", "PipelineTransformer", "build_resnet",
           "resnet18", "resnet50", "resnet101", "resnet152", "FourBlockMLP", "SimpleModel", "StageModule", "ToyMLP",
           "ToyModel", "regression_head", "SimpleUNet", "latitude_weighted_mse", "SimpleViT"]
