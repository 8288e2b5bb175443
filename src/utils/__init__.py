from vhmm_b200 import RandomChunkDataset, collate_fn, save_checkpoint, load_checkpoint  # noqa: F401


def _collate_cuda_default(batch, device='cuda'):  # src/utils/data.py:29 default device
    return collate_fn(batch, device)
