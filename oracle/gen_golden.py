#!/usr/bin/env python3
"""Generate tests/golden/*.npz by running the UNMODIFIED reference in this container.

TEST INFRASTRUCTURE ONLY.  Run from the repo root, in the authoring container (needs
/root/reference, which does not exist on the GPU box):

    PYTHONDONTWRITEBYTECODE=1 python oracle/gen_golden.py

The reference module /root/reference/VQ_VAE_HMM_fixed.py is imported by file path under a
private module name (nothing is copied).  Each fixture holds seeded inputs, the reference's
18 parameter tensors and the reference's own outputs:
  logits, q, mu, logvar, log_pi, log_A, loss(beta), the three loss terms, the 18 gradients
  (fp32), plus loss and gradients from a float64 rerun (.double()) = the noise-floor truth.
Cases:
  base_fresh    VAE_HMM(5,64,3,32,u_dim=4), torch.manual_seed(0) default init, B=6 T=50 ragged
  base_ckpt     same architecture, weights = models/vae_hmm.pt (small variances), B=6 T=50 ragged
  base_collate  base_fresh weights, batch produced by the reference RandomChunkDataset+collate_fn
                (random.seed(0)), B=8, lengths 20..60 -> arbitrary T (not a multiple of 4)
  smoke_cfg     VAE_HMM(5,8,3,4,u_dim=2,trans_hidden=8) of tests/smoke_test.py:19-26, B=3 T=16
  scaled_small  VAE_HMM(5,256,32,128,u_dim=16) (cfg 5 dims, H2=128 assumed), B=2 T=24 ragged
"""
import importlib.util
import os
import random
import sys

import numpy as np
import torch
import torch.nn.functional as F

REF = "/root/reference"
OUT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "tests", "golden")


def load_reference():
    spec = importlib.util.spec_from_file_location("_ref_vqvaehmm", os.path.join(REF, "VQ_VAE_HMM_fixed.py"))
    mod = importlib.util.module_from_spec(spec)
    sys.dont_write_bytecode = True
    spec.loader.exec_module(mod)
    return mod


def synth(B, C, U, T, seed, min_len):
    g = torch.Generator().manual_seed(seed)
    x = torch.randn(B, C, T, generator=g)
    u = torch.randn(B, U, T, generator=g)
    lengths = torch.randint(min_len, T + 1, (B,), generator=g)
    lengths[0] = T
    keep = (torch.arange(T)[None, :] < lengths[:, None]).unsqueeze(1)
    return x * keep, u * keep, lengths.long()


def run_case(ref, name, ctor, x, u, lengths, beta, state_dict=None, seed=0):
    torch.manual_seed(seed)
    model = ref.VAE_HMM(*ctor["args"], **ctor["kwargs"]).to("cpu")
    if state_dict is not None:
        model.load_state_dict(state_dict)
    model.train()
    out = {"x": x.numpy(), "u": u.numpy(), "lengths": lengths.numpy(), "beta": np.float64(beta)}
    out["ctor"] = np.array([ctor["args"][0], ctor["args"][1], ctor["args"][2], ctor["args"][3],
                            ctor["kwargs"]["u_dim"], ctor["kwargs"].get("trans_hidden", 128)], np.int64)
    for k, v in model.state_dict().items():
        out["param:" + k] = v.detach().numpy().copy()
    with torch.no_grad():
        logits = model.encode(x)
        (mu, logvar), q = model(x)
        log_pi, log_A = model.prior(u)
    out.update(logits=logits.numpy(), q=q.numpy(), mu=mu.numpy(), logvar=logvar.numpy(),
               log_pi=log_pi.numpy(), log_A=log_A.numpy(), regimes=q.argmax(dim=1).numpy())
    # loss and gradients (fp32)
    model.zero_grad()
    loss = model.compute_loss(x, u, lengths, beta)
    loss.backward()
    out["loss"] = loss.detach().numpy()
    for k, p in model.named_parameters():
        out["grad:" + k] = (p.grad if p.grad is not None else torch.zeros_like(p)).numpy().copy()
    # the three terms, through beta=0 / differences of the reference's own scalar
    with torch.no_grad():
        l0 = model.compute_loss(x, u, lengths, 0.0)
        l1 = model.compute_loss(x, u, lengths, 1.0)
    out["recon"] = l0.numpy()
    out["prior_minus_entropy"] = (l1.double() - l0.double()).numpy()
    # float64 rerun
    m64 = ref.VAE_HMM(*ctor["args"], **ctor["kwargs"]).double()
    m64.load_state_dict({k: v.double() for k, v in model.state_dict().items()})
    loss64 = m64.compute_loss(x.double(), u.double(), lengths, beta)
    loss64.backward()
    out["loss64"] = loss64.detach().numpy()
    for k, p in m64.named_parameters():
        out["grad64:" + k] = (p.grad if p.grad is not None else torch.zeros_like(p)).numpy().copy()
    with torch.no_grad():
        out["recon64"] = m64.compute_loss(x.double(), u.double(), lengths, 0.0).numpy()
    path = os.path.join(OUT, name + ".npz")
    np.savez_compressed(path, **out)
    print(f"{name}: loss={float(loss):.6f} loss64={float(loss64):.9f} T={x.shape[2]} -> {path}")


def main():
    os.makedirs(OUT, exist_ok=True)
    ref = load_reference()
    ref.device = torch.device("cpu")
    base = {"args": (5, 64, 3, 32), "kwargs": {"u_dim": 4}}
    x, u, L = synth(6, 5, 4, 50, 1234, 20)
    run_case(ref, "base_fresh", base, x, u, L, beta=0.7)
    sd = torch.load(os.path.join(REF, "models", "vae_hmm.pt"), map_location="cpu")
    run_case(ref, "base_ckpt", base, x, u, L, beta=1.0, state_dict=sd)
    # the real dataset + collate path (cfg 1 style), arbitrary T
    random.seed(0)
    g = torch.Generator().manual_seed(1234)
    xs = torch.randn(10, 5, 120, generator=g)
    us = torch.randn(10, 4, 120, generator=g)
    ds = ref.RandomChunkDataset(xs, us, min_len=20, max_len=60)
    batch = [ds[i] for i in range(8)]
    xb, ub, Lb = ref.collate_fn(batch)
    run_case(ref, "base_collate", base, xb, ub, Lb, beta=0.25)
    smoke = {"args": (5, 8, 3, 4), "kwargs": {"u_dim": 2, "trans_hidden": 8}}
    x, u, L = synth(3, 5, 2, 16, 7, 5)
    run_case(ref, "smoke_cfg", smoke, x, u, L, beta=1.0)
    scaled = {"args": (5, 256, 32, 128), "kwargs": {"u_dim": 16}}
    x, u, L = synth(2, 5, 16, 24, 99, 10)
    run_case(ref, "scaled_small", scaled, x, u, L, beta=0.5)


if __name__ == "__main__":
    main()
