#!/usr/bin/env python3
"""Two launches of every exact-HMM stage at the bench shapes (cfg 3: B=2^20, T=256, K=3; cfg 4: B=64, T=65536), for an ncu
launch list (per-kernel times behind the stage lines of bench.py).   python tools/prof_hmm_stages.py"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from vhmm_b200 import functional as Fn
g = torch.Generator(device="cuda").manual_seed(0)
def inputs(B, T, K):
    la = torch.log_softmax(torch.randn(B, T, K, K, device="cuda", generator=g), -1)
    lb = -0.5 * torch.randn(B, T, K, device="cuda", generator=g) ** 2 - 1
    lp = torch.log_softmax(torch.randn(K, device="cuda", generator=g), -1)
    return lp, la, lb, torch.full((B,), T, dtype=torch.int64, device="cuda")
for (B, T, K) in ((1 << 20, 256, 3), (64, 65536, 3)):
    lp, la, lb, L = inputs(B, T, K)
    for _ in range(2):
        Fn.hmm_forward(lp, la, lb, L)
        Fn.viterbi(lp, la, lb, L)
        Fn.hmm_posterior(lp, la, lb, L, want_xi=False)
        Fn.hmm_posterior(lp, la, lb, L, want_xi=True)
        Fn.hmm_decode(lp, la, lb, L)
        Fn.hmm_scan(lp, la, lb, L)
        Fn.hmm_scan(lp, la, lb, L, want_alpha=True)
    torch.cuda.synchronize()
    del la, lb
    Fn.release_workspaces(); torch.cuda.empty_cache()
print("ok")
