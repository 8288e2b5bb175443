"""vhmm_b200: B200-native (sm_100a) implementation of the VAE_HMM hot path of
yashnaray/VQ-VAE-HMM-model behind the reference's own Python API.

    from vhmm_b200 import VAE_HMM, collate_fn, RandomChunkDataset, train_model

Kernels live in csrc/ and are reached through the C ABI declared in include/vhmm.h
(libvhmm.so, bound with ctypes in _lib.py).  There is no CPU fallback.
"""
from ._lib import VhmmError, build, lib, LIB_PATH  # noqa: F401
from . import functional  # noqa: F401
from .model import VAE_HMM, Encoder, Prior, Decoder  # noqa: F401
from .data import (RandomChunkDataset, collate_fn, train_model, device,  # noqa: F401
                   save_checkpoint, load_checkpoint)
from . import parallel  # noqa: F401
from . import training  # noqa: F401
from . import serving  # noqa: F401
from .training import DeviceCorpus, DeviceChunkLoader, FusedAdam, Trainer, train_model_fused  # noqa: F401

__all__ = ["VAE_HMM", "Encoder", "Prior", "Decoder", "RandomChunkDataset", "collate_fn",
           "train_model", "device", "save_checkpoint", "load_checkpoint", "functional",
           "parallel", "training", "serving", "DeviceCorpus", "DeviceChunkLoader", "FusedAdam", "Trainer",
           "train_model_fused", "build", "lib", "VhmmError"]
