from vhmm_b200 import VAE_HMM, Encoder, Prior, Decoder  # noqa: F401  (src/models/vae_hmm.py:8-126)
