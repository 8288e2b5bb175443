"""Oracle for the *pinned* path: a functional CPU restatement (torch ops on CPU tensors)
of the reference's Encoder / Prior / Decoder / mean-field ELBO.

TEST INFRASTRUCTURE ONLY (see oracle/__init__.py).  Every function cites the reference
lines it restates (paths relative to /root/reference).  Parameters travel as a plain
``dict`` keyed exactly like the reference ``state_dict``, so a
reference checkpoint can be fed in unchanged.  Works in float32 or float64 (dtype follows
the parameters), which is how the fp32 noise floor is measured.
"""
from __future__ import annotations

import math
from typing import Dict, Tuple

import torch
import torch.nn.functional as F

PARAM_ORDER = (
    "encoder.conv1.weight", "encoder.conv1.bias",
    "encoder.conv2.weight", "encoder.conv2.bias",
    "encoder.to_logits.weight", "encoder.to_logits.bias",
    "prior.log_prior",
    "prior.transition_net.0.weight", "prior.transition_net.0.bias",
    "prior.transition_net.2.weight", "prior.transition_net.2.bias",
    "decoder.embeddings.weight",
    "decoder.conv1.weight", "decoder.conv1.bias",
    "decoder.conv2.weight", "decoder.conv2.bias",
    "decoder.to_params.weight", "decoder.to_params.bias",
)


def param_shapes(C: int, H: int, K: int, H2: int, U: int, TH: int = 128) -> Dict[str, tuple]:
    """Shapes of the 18 tensors; VQ_VAE_HMM_fixed.py:33-36, :47-57, :76-79, :96-98."""
    return {
        "encoder.conv1.weight": (H, C, 3), "encoder.conv1.bias": (H,),
        "encoder.conv2.weight": (H2, H, 3), "encoder.conv2.bias": (H2,),
        "encoder.to_logits.weight": (K, H2, 1), "encoder.to_logits.bias": (K,),
        "prior.log_prior": (K,),
        "prior.transition_net.0.weight": (TH, U), "prior.transition_net.0.bias": (TH,),
        "prior.transition_net.2.weight": (K * K, TH), "prior.transition_net.2.bias": (K * K,),
        "decoder.embeddings.weight": (K, H),
        "decoder.conv1.weight": (H, H, 3), "decoder.conv1.bias": (H,),
        "decoder.conv2.weight": (H, H, 3), "decoder.conv2.bias": (H,),
        "decoder.to_params.weight": (2 * C, H, 1), "decoder.to_params.bias": (2 * C,),
    }


def init_params(C, H, K, H2, U, TH=128, seed=0, dtype=torch.float32) -> Dict[str, torch.Tensor]:
    """Deterministic synthetic parameters (NOT torch's default init; the golden fixtures
    carry the reference's own default-init and checkpoint weights).  Uniform(+-1/sqrt(fan_in))
    like Conv1d/Linear defaults, N(0,1) embeddings, small random log_prior."""
    g = torch.Generator().manual_seed(seed)
    out = {}
    for name, shp in param_shapes(C, H, K, H2, U, TH).items():
        if name == "decoder.embeddings.weight":
            t = torch.randn(shp, generator=g)
        elif name == "prior.log_prior":
            t = 0.1 * torch.randn(shp, generator=g)
        else:
            fan_in = 1
            for s in (shp[1:] if len(shp) > 1 else ()):
                fan_in *= s
            if len(shp) == 1:  # bias: fan_in of the matching weight
                w = param_shapes(C, H, K, H2, U, TH)[name.replace("bias", "weight")]
                fan_in = 1
                for s in w[1:]:
                    fan_in *= s
            bound = 1.0 / math.sqrt(fan_in)
            t = (torch.rand(shp, generator=g) * 2 - 1) * bound
        out[name] = t.to(dtype)
    return out


def encode(p, x):
    """logits (B,K,T).  VQ_VAE_HMM_fixed.py:38-41."""
    h = F.relu(F.conv1d(x, p["encoder.conv1.weight"], p["encoder.conv1.bias"], padding=1))
    h = F.relu(F.conv1d(h, p["encoder.conv2.weight"], p["encoder.conv2.bias"], padding=1))
    return F.conv1d(h, p["encoder.to_logits.weight"], p["encoder.to_logits.bias"])


def prior(p, u, U: int):
    """(log_pi (K,), log_A (B,T,K,K)).  VQ_VAE_HMM_fixed.py:59-71 incl. the permute rule :64-65."""
    K = p["prior.log_prior"].shape[0]
    if u is None:
        raise ValueError("u required for non-stationary transitions")
    if u.dim() == 3 and u.shape[1] == U:
        u = u.permute(0, 2, 1)
    B, T, _ = u.shape
    a = F.relu(F.linear(u.reshape(B * T, -1), p["prior.transition_net.0.weight"],
                        p["prior.transition_net.0.bias"]))
    m = F.linear(a, p["prior.transition_net.2.weight"], p["prior.transition_net.2.bias"])
    log_A = F.log_softmax(m.view(B, T, K, K), dim=-1)
    return F.log_softmax(p["prior.log_prior"], dim=-1), log_A


def decode(p, q):
    """(mu, logvar), each (B,C,T).  VQ_VAE_HMM_fixed.py:81-90."""
    e = torch.matmul(q.permute(0, 2, 1), p["decoder.embeddings.weight"]).permute(0, 2, 1)
    g = F.relu(F.conv1d(e, p["decoder.conv1.weight"], p["decoder.conv1.bias"], padding=1))
    g = F.relu(F.conv1d(g, p["decoder.conv2.weight"], p["decoder.conv2.bias"], padding=1))
    prm = F.conv1d(g, p["decoder.to_params.weight"], p["decoder.to_params.bias"])
    C = prm.shape[1] // 2
    return prm[:, :C, :], prm[:, C:, :]


def forward(p, x):
    """((mu, logvar), q).  VQ_VAE_HMM_fixed.py:139-143."""
    q = F.softmax(encode(p, x), dim=1)
    mu, logvar = decode(p, q)
    return (mu, logvar), q


def loss_terms(p, x, u, lengths, U: int) -> Tuple[torch.Tensor, torch.Tensor, torch.Tensor]:
    """(recon_loss, prior_loss, entropy) exactly as normalised by the reference.
    VQ_VAE_HMM_fixed.py:111-135."""
    B, C, T = x.shape
    if lengths is None:
        raise ValueError("lengths required")
    valid = torch.arange(T, device=x.device)[None, :] < lengths.to(x.device)[:, None]   # :111
    log_pi, log_A = prior(p, u, U)                                           # :112
    logits = encode(p, x)                                                    # :113
    q = F.softmax(logits, dim=1)                                             # :114
    mu, logvar = decode(p, q)                                                # :115
    var = logvar.exp().clamp(min=1e-8)                                       # :118
    nll = 0.5 * (torch.log(2 * math.pi * var) + (mu - x) ** 2 / var)         # :119
    vf = valid.to(x.dtype)
    recon = (nll * vf.unsqueeze(1)).sum() / (valid.sum() * C).clamp(min=1.0)  # :120
    init = (q[:, :, 0] * log_pi.unsqueeze(0)).sum(dim=1)                     # :123
    qp = q[:, :, :-1].permute(0, 2, 1).unsqueeze(-1)                         # :125
    qn = q[:, :, 1:].permute(0, 2, 1).unsqueeze(-2)                          # :126
    trans = (qp * qn * log_A[:, 1:]).sum(dim=(2, 3))                         # :127
    tmask = (valid[:, 1:] & valid[:, :-1]).to(x.dtype)                       # :128
    trans = (trans * tmask).sum(dim=1)                                       # :129
    prior_loss = -(init + trans).mean()                                      # :131
    ent = -(q * F.log_softmax(logits, dim=1)).sum(dim=1)                     # :134
    ent = (ent * vf).sum() / B                                               # :135
    return recon, prior_loss, ent


def compute_loss(p, x, u, lengths, U: int, beta: float = 1.0) -> torch.Tensor:
    """recon + beta*(prior_loss - entropy).  VQ_VAE_HMM_fixed.py:137."""
    recon, prior_loss, ent = loss_terms(p, x, u, lengths, U)
    return recon + beta * (prior_loss - ent)


def loss_and_grads(p, x, u, lengths, U: int, beta: float = 1.0):
    """Loss, its three terms and d loss / d param for the 18 tensors (autograd on CPU),
    i.e. what ``loss.backward()`` leaves in ``.grad`` at VQ_VAE_HMM_fixed.py:155-156."""
    pp = {k: v.detach().clone().requires_grad_(True) for k, v in p.items()}
    recon, prior_loss, ent = loss_terms(pp, x, u, lengths, U)
    loss = recon + beta * (prior_loss - ent)
    grads = torch.autograd.grad(loss, [pp[k] for k in PARAM_ORDER], allow_unused=True)
    gd = {k: (g if g is not None else torch.zeros_like(pp[k])) for k, g in zip(PARAM_ORDER, grads)}
    return loss.detach(), (recon.detach(), prior_loss.detach(), ent.detach()), gd


def argmax_regimes(q):
    """Hard regimes, first-index tie-break.  backtesting.py:155, src/backtesting.py:107."""
    return q.argmax(dim=1)


def synth_batch(B, C, U, T, seed=1234, ragged=True, min_len=None, dtype=torch.float32):
    """Seeded synthetic batch with the collate_fn contract (zero tails, int64 CPU lengths).
    SURVEY.md section 8(d): x,u ~ N(0,1), generator seed 1234."""
    g = torch.Generator().manual_seed(seed)
    x = torch.randn(B, C, T, generator=g)
    u = torch.randn(B, U, T, generator=g)
    if ragged:
        lo = max(1, T // 2) if min_len is None else min_len
        lengths = torch.randint(lo, T + 1, (B,), generator=g)
        lengths[0] = T  # collate pads to the batch max
    else:
        lengths = torch.full((B,), T, dtype=torch.long)
    keep = (torch.arange(T)[None, :] < lengths[:, None]).unsqueeze(1)
    x = x * keep
    u = u * keep
    return x.to(dtype), u.to(dtype), lengths.long()
