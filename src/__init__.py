"""Packaged-layout shim of the reference (``from src.models import VAE_HMM``, ``from src.training import
Trainer``, ``from src.utils import RandomChunkDataset, collate_fn``; examples/train_example.py:5-8,
src/__init__.py:11-19).  ``src.models``, ``src.utils`` and ``src.training`` are this repo's; every other
sub-package of the reference's ``src`` (``src.portfolio``, ``src.backtesting``: out of scope, pure-torch
consumers) resolves to the reference checkout behind this repo on ``sys.path``, whose ``src`` directory is
appended to this package's search path."""
import os as _os

from vhmm_b200 import _refcompat as _rc

for _r in _rc.reference_roots():
    _p = _os.path.join(_r, "src")
    if _os.path.isdir(_p) and _p not in __path__:
        __path__.append(_p)

__version__ = "1.0.0"
_LAZY = {
    "VAE_HMM": "src.models", "Encoder": "src.models", "Decoder": "src.models", "Prior": "src.models",
    "Trainer": "src.training", "MetaTrainer": "src.training", "WalkForwardTrainer": "src.training",
    "RegimePortfolioOptimizer": "src.portfolio", "AttentionPortfolioOptimizer": "src.portfolio",
    "TransformerPortfolioOptimizer": "src.portfolio", "BayesianPortfolioOptimizer": "src.portfolio",
    "RegimeDeltaHedger": "src.portfolio", "DynamicDeltaHedger": "src.portfolio",
}


def __getattr__(name):
    mod = _LAZY.get(name)
    if mod is None:
        raise AttributeError(f"module 'src' has no attribute {name!r}")
    import importlib
    try:
        return getattr(importlib.import_module(mod), name)
    except ModuleNotFoundError as e:
        raise ImportError(f"cannot import name '{name}' from 'src': {mod} is served by the reference checkout, "
                          f"which is not on PYTHONPATH behind this repo") from e
