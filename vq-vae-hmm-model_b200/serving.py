"""Serving path of the reference, host side: the body of the ``/infer`` endpoint of
inference_api/app.py:57-77 -- encode -> softmax -> decode on one (C,T) series, results as nested lists
under the keys ``mu``, ``logvar``, ``regime_probs`` -- and a batched form for request queues.

The reference runs every request as its own (1,C,T) tensor.  Batching requests of DIFFERENT lengths into
one zero-padded tensor would change the last steps of the shorter ones (the convolutions see the padded
region of the batch, SURVEY.md finding 4), so ``infer_many`` only fuses requests of EQUAL length into one
(B,C,T) call -- bit-identical to B single calls -- and keeps the reference's per-request semantics for
the rest.  No web framework is involved here (FastAPI / gunicorn stay the reference's own).
"""
from __future__ import annotations

import json
from collections import defaultdict
from typing import Dict, List, Sequence

import torch

from .model import VAE_HMM


def load_inference_model(cfg_path: str = 'inference_config.json', device: str = 'cuda') -> VAE_HMM:
    """inference_api/app.py:24-47: model dims and checkpoint path from the JSON config."""
    with open(cfg_path, 'r') as f:
        cfg = json.load(f)
    m = cfg.get('model', {})
    model = VAE_HMM(input_dim=m.get('input_dim', 5), hidden_dim=m.get('hidden_dim', 64), K=m.get('K', 3),
                    hidden_dim2=m.get('hidden_dim2', 32), u_dim=m.get('u_dim', None),
                    trans_hidden=m.get('trans_hidden', 128))
    state = torch.load(cfg.get('checkpoint_path', 'checkpoints/vae_hmm_trained.pth'), map_location='cpu')
    model.load_state_dict(state)
    return model.to(device).eval()


def _forward(model: VAE_HMM, x: torch.Tensor):
    with torch.no_grad():
        (mu, logvar), q = model(x)          # encode -> softmax -> decode (VQ_VAE_HMM_fixed.py:139-143)
    return mu, logvar, q


def infer(model: VAE_HMM, x_list: Sequence[Sequence[float]]) -> Dict[str, list]:
    """One request: ``x_list`` is the (C,T) series of InferRequest.x (app.py:20-21, :57-77)."""
    dev = next(model.parameters()).device
    x = torch.tensor(x_list, dtype=torch.float32).unsqueeze(0).to(dev)
    mu, logvar, q = _forward(model, x)
    return {'mu': mu.squeeze(0).cpu().tolist(), 'logvar': logvar.squeeze(0).cpu().tolist(),
            'regime_probs': q.squeeze(0).cpu().tolist()}


def infer_many(model: VAE_HMM, requests: Sequence[Sequence[Sequence[float]]]) -> List[Dict[str, list]]:
    """A queue of requests: those of equal length share one (B,C,T) launch; results in request order."""
    dev = next(model.parameters()).device
    groups = defaultdict(list)
    tensors = []
    for i, r in enumerate(requests):
        t = torch.tensor(r, dtype=torch.float32)
        tensors.append(t)
        groups[tuple(t.shape)].append(i)
    out: List[Dict[str, list]] = [None] * len(requests)
    for _, idx in groups.items():
        x = torch.stack([tensors[i] for i in idx]).to(dev)
        mu, logvar, q = _forward(model, x)
        mu, logvar, q = mu.cpu(), logvar.cpu(), q.cpu()
        for j, i in enumerate(idx):
            out[i] = {'mu': mu[j].tolist(), 'logvar': logvar[j].tolist(), 'regime_probs': q[j].tolist()}
    return out
