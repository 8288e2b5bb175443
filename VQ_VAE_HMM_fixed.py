"""Drop-in module name: every reference call site does ``from VQ_VAE_HMM_fixed import VAE_HMM,
train_model, collate_fn, RandomChunkDataset`` (train.py:4, data_loader.py:6,
training_pipeline/train.py:16, evaluation/evaluate.py:19, inference_api/app.py:10, app.py:5, api.py:4,
tests/smoke_test.py:11, examples/*.py).  Putting this repo on PYTHONPATH ahead of the reference switches
those imports to the B200 implementation without touching the callers.

The names of the reference module that are NOT on the hot path (VQ_VAE_HMM_fixed.py:182-260:
``RegimePortfolioOptimizer``, ``sharpe_loss``, ``RegimeLSTMOptimizer``, ``regime_aware_sharpe_loss``,
``train_portfolio_optimizer``; train.py:4 imports two of them) are pure-torch consumers of the model and
stay the reference's own code: they are resolved lazily from the reference checkout behind this repo on
``sys.path`` (vhmm_b200/_refcompat.py)."""
from vhmm_b200 import (VAE_HMM, Encoder, Prior, Decoder, RandomChunkDataset, collate_fn,  # noqa: F401
                       train_model, device)
from vhmm_b200 import _refcompat as _rc

_DELEGATED = ("RegimePortfolioOptimizer", "sharpe_loss", "RegimeLSTMOptimizer", "regime_aware_sharpe_loss",
              "train_portfolio_optimizer")


def __getattr__(name):
    if name in _DELEGATED:
        return _rc.delegated(__name__, "VQ_VAE_HMM_fixed.py", name)
    raise AttributeError(f"module {__name__!r} has no attribute {name!r}")
