#!/usr/bin/env python3
"""cfg 4 (B=64, T=65536, K=3): vhmm_hmm_scan log-likelihood and all-prefix alpha, a few calls each (for ncu launch lists)"""
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
for want in (False, True):
    for _ in range(3): Fn.hmm_scan(lp, la, lb, L, want_alpha=want)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(50): Fn.hmm_scan(lp, la, lb, L, want_alpha=want)
    e1.record(); torch.cuda.synchronize()
    print("want_alpha", want, "%.4f ms" % (e0.elapsed_time(e1) / 50), flush=True)
