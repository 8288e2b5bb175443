#!/usr/bin/env python3
"""One launch of each kernel that round 2 added or that carries a headline (for ncu --set full):
   ELBO step at cfg 2 (decoder / enc_prior / enc_bwd tile kernels), the fused forward-only tile kernel, the K = 32
   warp kernels of hmm_wide.cuh, the one-pass posterior at cfg 3 (B reduced).   python tools/prof_r2.py"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
import vhmm_b200 as V
from vhmm_b200 import functional as Fn

torch.manual_seed(0)
m = V.VAE_HMM(5, 64, 3, 32, u_dim=4).cuda()
B, T = 65536, 512
g = torch.Generator(device="cuda").manual_seed(1)
x = torch.randn(B, 5, T, device="cuda", generator=g); u = torch.randn(B, 4, T, device="cuda", generator=g)
L = torch.full((B,), T, dtype=torch.int64)
for _ in range(2):
    m.zero_grad(); loss = m.compute_loss(x, u, L, 1.0); loss.backward()
with torch.no_grad():
    for _ in range(2):
        m(x)
torch.cuda.synchronize()
del x, u
Fn.release_workspaces(); torch.cuda.empty_cache()
K, Bk, Tk = 32, 2048, 1024
lp = torch.log_softmax(torch.randn(K, device="cuda", generator=g), 0)
la = torch.log_softmax(torch.randn(Bk, Tk, K, K, device="cuda", generator=g), -1)
lb = -0.5 * torch.randn(Bk, Tk, K, device="cuda", generator=g) ** 2 - 1
Lk = torch.full((Bk,), Tk, dtype=torch.int64, device="cuda")
for _ in range(2):
    Fn.hmm_forward(lp, la, lb, Lk); Fn.viterbi(lp, la, lb, Lk); Fn.hmm_posterior(lp, la, lb, Lk, want_xi=False)
torch.cuda.synchronize()
print("ok")
