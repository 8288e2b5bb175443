"""``from src.utils.data import ...`` (src/utils/__init__.py:3; src/utils/data.py:7-44)."""
from vhmm_b200 import RandomChunkDataset, collate_fn, save_checkpoint, load_checkpoint  # noqa: F401
