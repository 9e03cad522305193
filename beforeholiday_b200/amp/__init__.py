from .scaler import LossScaler

__all__ = ["LossScaler"]
