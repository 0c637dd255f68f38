from .trainer import Trainer, save_model_state

__all__ = ["Trainer", "save_model_state"]
