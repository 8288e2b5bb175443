#!/usr/bin/env python3
"""cfg-2-sized (B=65536, T=512) comparison of the flat ELBO gradient between two builds of the
library (default FLUSH vs tools/bin/libvhmm_flush16.so) -- run twice with VHMM_LIB set / unset:
    python tools/flush_check.py save /tmp/g_a.pt ; VHMM_LIB=... python tools/flush_check.py save /tmp/g_b.pt
    python tools/flush_check.py cmp /tmp/g_a.pt /tmp/g_b.pt"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch

if sys.argv[1] == "save":
    import vhmm_b200 as V
    from vhmm_b200 import functional as Fn
    torch.manual_seed(0)
    m = V.VAE_HMM(5, 64, 3, 32, u_dim=4).cuda()
    g = torch.Generator(device="cuda").manual_seed(1234)
    B, T = 65536, 512
    x = torch.randn(B, 5, T, device="cuda", generator=g); u = torch.randn(B, 4, T, device="cuda", generator=g)
    L = torch.full((B,), T, dtype=torch.int64)
    loss, flat = m.loss_and_flat_grads(x, u, L, 1.0)
    torch.save({"loss": loss.cpu(), "flat": flat.cpu(), "off": Fn.param_offsets(m._dims)}, sys.argv[2])
    print("saved", float(loss))
else:
    a, b = torch.load(sys.argv[2]), torch.load(sys.argv[3])
    off = a["off"]
    from vhmm_b200.functional import PARAM_ORDER
    print("loss", float(a["loss"]), float(b["loss"]))
    worst = 0.0
    for i, k in enumerate(PARAM_ORDER):
        ga, gb = a["flat"][off[i]:off[i + 1]].double(), b["flat"][off[i]:off[i + 1]].double()
        r = float((ga - gb).norm() / gb.norm().clamp_min(1e-30))
        worst = max(worst, r)
        print(f"{k:36s} rel diff {r:.2e}  |g| {float(gb.norm()):.3e}")
    print("worst", worst)
