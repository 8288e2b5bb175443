"""Drop-in module name: every reference call site does ``from VQ_VAE_HMM_fixed import VAE_HMM,
train_model, collate_fn, RandomChunkDataset`` (train.py:4, data_loader.py:6,
training_pipeline/train.py:16, evaluation/evaluate.py:19, inference_api/app.py:10,
tests/smoke_test.py:11).  Putting this repo on PYTHONPATH ahead of the reference switches those
imports to the B200 implementation without touching the callers."""
from vhmm_b200 import (VAE_HMM, Encoder, Prior, Decoder, RandomChunkDataset, collate_fn,  # noqa: F401
                       train_model, device)
