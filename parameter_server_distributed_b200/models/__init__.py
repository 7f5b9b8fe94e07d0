"""Model families of the BASELINE.json configs (the reference itself ships no model)."""
from .embedding import EmbeddingMLP, EmbeddingMLPConfig
from .gpt2 import GPT2, GPT2Config, lm_loss
from .mlp import MLP, MLPConfig, mse_loss
from .resnet import ResNet, ce_loss, resnet50, resnet_tiny

__all__ = ["MLP", "MLPConfig", "mse_loss", "GPT2", "GPT2Config", "lm_loss", "ResNet", "resnet50",
           "resnet_tiny", "ce_loss", "EmbeddingMLP", "EmbeddingMLPConfig"]
