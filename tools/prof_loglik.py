#!/usr/bin/env python3
"""One call each of the fused front end and the fused log-likelihood kernel (for ncu captures)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import vhmm_b200 as V
B = int(sys.argv[1]) if len(sys.argv) > 1 else 1 << 18
T = 256
m = V.VAE_HMM(5, 64, 3, 32, u_dim=4).cuda().eval()
g = torch.Generator(device="cuda").manual_seed(1)
x = torch.randn(B, 5, T, device="cuda", generator=g); u = torch.randn(B, 4, T, device="cuda", generator=g)
with torch.no_grad():
    for _ in range(2):
        m.hmm_inputs(x, u)
        m.hmm_log_likelihood(x, u)
torch.cuda.synchronize()
