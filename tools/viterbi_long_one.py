#!/usr/bin/env python3
"""one vhmm_viterbi call on B=64, T=65536, K=3 (for an ncu capture of hv::vit_long_kernel)"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from vhmm_b200 import functional as Fn
B, T, K = 64, 65536, 3
g = torch.Generator(device="cuda").manual_seed(0)
la = torch.log_softmax(torch.randn(B, T, K, K, device="cuda", generator=g), -1)
lb = torch.randn(B, T, K, device="cuda", generator=g)
lp = torch.log_softmax(torch.randn(K, device="cuda", generator=g), -1)
L = torch.full((B,), T, dtype=torch.int64, device="cuda")
for _ in range(2): Fn.viterbi(lp, la, lb, L)
torch.cuda.synchronize()
print("ok")
