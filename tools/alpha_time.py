#!/usr/bin/env python3
"""vhmm_hmm_forward with all-prefix alpha at B=2^20, T=256, K=3 (VHMM_ALPHA=sweep selects the thread-per-sequence sweep)"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from vhmm_b200 import functional as Fn
B, T, K = 1 << 20, 256, 3
g = torch.Generator(device="cuda").manual_seed(0)
la = torch.log_softmax(torch.randn(B, T, K, K, device="cuda", generator=g), -1)
lb = torch.randn(B, T, K, device="cuda", generator=g)
lp = torch.log_softmax(torch.randn(K, device="cuda", generator=g), -1)
L = torch.full((B,), T, dtype=torch.int64, device="cuda")
for _ in range(3): Fn.hmm_forward(lp, la, lb, L, want_alpha=True)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(10): ll, al = Fn.hmm_forward(lp, la, lb, L, want_alpha=True)
e1.record(); torch.cuda.synchronize()
print(os.environ.get("VHMM_ALPHA", "onepass"), "%.3f ms" % (e0.elapsed_time(e1) / 10), "checksum %.6e" % float(al[::4097].double().sum()))
