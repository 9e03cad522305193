from .api import initialize, load_state_dict, master_params, scale_loss, state_dict
from .optimizer_hooks import process_optimizer
from .scaler import LossScaler

__all__ = ["LossScaler", "initialize", "scale_loss", "master_params", "state_dict", "load_state_dict", "process_optimizer"]
