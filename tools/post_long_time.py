#!/usr/bin/env python3
"""vhmm_hmm_posterior (gamma) on LONG sequences: B=64, T=65536 (cfg-4 shape) and B=4096, T=4096"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from vhmm_b200 import functional as Fn
K = 3
for B, T in ((64, 65536), (4096, 4096)):
    g = torch.Generator(device="cuda").manual_seed(0)
    la = torch.log_softmax(torch.randn(B, T, K, K, device="cuda", generator=g), -1)
    lb = torch.randn(B, T, K, device="cuda", generator=g)
    lp = torch.log_softmax(torch.randn(K, device="cuda", generator=g), -1)
    L = torch.full((B,), T, dtype=torch.int64, device="cuda")
    for _ in range(2): Fn.hmm_posterior(lp, la, lb, L, want_xi=False)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(5): out = Fn.hmm_posterior(lp, la, lb, L, want_xi=False)
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 5
    print("B", B, "T", T, "gamma %.3f ms" % ms, "= %.0f GB/s on 60 B/step" % (60 * B * T / ms / 1e6), flush=True)
