#!/usr/bin/env python3
"""one step of the cfg-5 model at B=1024 (for ncu captures of the tcgen05 layer kernels)"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import vhmm_b200 as V
B, T = int(sys.argv[1]) if len(sys.argv) > 1 else 1024, 1024
torch.manual_seed(0)
m = V.VAE_HMM(5, 256, 32, 128, u_dim=16).cuda()
g = torch.Generator(device="cuda").manual_seed(1)
x = torch.randn(B, 5, T, device="cuda", generator=g); u = torch.randn(B, 16, T, device="cuda", generator=g)
L = torch.full((B,), T, dtype=torch.int64)
for _ in range(2):
    m.zero_grad(set_to_none=True); loss = m.compute_loss(x, u, L, 1.0); loss.backward()
torch.cuda.synchronize(); print(float(loss))
