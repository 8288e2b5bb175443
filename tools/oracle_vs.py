#!/usr/bin/env python3
"""GPU-box diagnostic: fused and layerwise gradients vs the float64 oracle on a random base-config batch."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import torch
import vhmm_b200 as V
from vhmm_b200 import functional as Fn
from oracle import elbo as O
from conftest import rel_err
B, T = int(sys.argv[1]), int(sys.argv[2])
torch.manual_seed(0)
m = V.VAE_HMM(5, 64, 3, 32, u_dim=4)
p = {k: v.detach().clone() for k, v in m.state_dict().items()}
m = m.cuda()
gen = torch.Generator().manual_seed(7)
x = torch.randn(B, 5, T, generator=gen); u = torch.randn(B, 4, T, generator=gen)
L = torch.full((B,), T, dtype=torch.int64)
p64 = {k: v.double() for k, v in p.items()}
ref, _, g64 = O.loss_and_grads(p64, x.double(), u.double(), L, 4, 0.7)
for impl in ("layerwise", "fused"):
    Fn.elbo_select(impl)
    m.zero_grad(set_to_none=True)
    loss = m.compute_loss(x.cuda(), u.cuda(), L, 0.7); loss.backward()
    errs = {k: rel_err(q.grad.double().cpu(), g64[k]) for k, q in m.named_parameters()}
    w = max(errs.items(), key=lambda kv: kv[1])
    print(f"{impl:9s} B={B} T={T} loss rel {abs(float(loss)-float(ref))/abs(float(ref)):.2e} worst grad {w[1]:.2e} ({w[0]}); decoder.conv1.bias {errs['decoder.conv1.bias']:.2e} |g|={float(g64['decoder.conv1.bias'].norm()):.3e}")
