"""Packaged-layout shim of the reference (``from src.models import VAE_HMM``,
``from src.utils import RandomChunkDataset, collate_fn``; examples/train_example.py:5-8)."""
