"""Training surface of the reference (css/trainer/__init__.py) on the native step."""
from .engine import ConformerTrainEngine
from .iteration_types import train_one_epoch, validate
from .lr_scheduler import LRScheduler
from .optim import FusedAdam, GradientExchange

__all__ = ["ConformerTrainEngine", "FusedAdam", "GradientExchange", "LRScheduler", "train_one_epoch", "validate"]
