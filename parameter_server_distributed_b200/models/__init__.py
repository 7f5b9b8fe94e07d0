"""Model families of the BASELINE.json configs (the reference itself ships no model)."""
from .mlp import MLP, MLPConfig, mse_loss

__all__ = ["MLP", "MLPConfig", "mse_loss"]
