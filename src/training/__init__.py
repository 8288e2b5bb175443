"""``from src.training import Trainer`` (examples/train_example.py:7, src/training/__init__.py:3).
``Trainer`` is the fused-step mirror of src/training/trainer.py:9-43 (vhmm_b200.training.Trainer);
``MetaTrainer`` / ``WalkForwardTrainer`` (:46-110, out of scope) come from the reference checkout."""
from vhmm_b200.training import Trainer  # noqa: F401
from vhmm_b200 import _refcompat as _rc


def __getattr__(name):
    if name in ("MetaTrainer", "WalkForwardTrainer"):
        return _rc.delegated(__name__, "src/training/trainer.py", name)
    raise AttributeError(f"module {__name__!r} has no attribute {name!r}")
