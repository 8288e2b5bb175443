"""``from src.training.trainer import Trainer, MetaTrainer, WalkForwardTrainer`` (src/__init__.py:19)."""
from vhmm_b200.training import Trainer  # noqa: F401
from vhmm_b200 import _refcompat as _rc


def __getattr__(name):
    if name in ("MetaTrainer", "WalkForwardTrainer"):
        return _rc.delegated(__name__, "src/training/trainer.py", name)
    raise AttributeError(f"module {__name__!r} has no attribute {name!r}")
