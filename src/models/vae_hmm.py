"""``from src.models.vae_hmm import VAE_HMM, Encoder, Decoder, Prior`` (src/__init__.py:11, src/models/__init__.py:2)."""
from vhmm_b200 import VAE_HMM, Encoder, Prior, Decoder  # noqa: F401
