#!/usr/bin/env python3
"""vhmm_hmm_scan (segment products + carry) against the warp-per-sequence forward for 4 < K <= 32 on few long sequences"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from vhmm_b200 import functional as Fn
g = torch.Generator(device="cuda").manual_seed(0)
def timed(fn, n=3):
    for _ in range(2): fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n): out = fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n, out
for (B, T, K) in ((4, 16384, 32), (64, 2048, 32), (256, 1024, 32), (1, 65536, 32), (4, 16384, 16), (4, 16384, 8), (64, 2048, 6)):
    la = torch.log_softmax(torch.randn(B, T, K, K, device="cuda", generator=g), -1)
    lb = torch.randn(B, T, K, device="cuda", generator=g) * 2 - 3
    lp = torch.log_softmax(torch.randn(K, device="cuda", generator=g), -1)
    L = torch.full((B,), T, dtype=torch.int64, device="cuda")
    ms1, a = timed(lambda: Fn.hmm_scan(lp, la, lb, L))
    ms2, b = timed(lambda: Fn.hmm_forward(lp, la, lb, L, time_parallel=False))
    print("B=%d T=%d K=%d scan %.3f ms  warp-per-sequence %.3f ms  max rel diff %.2e" %
          (B, T, K, ms1, ms2, float(((a - b) / b).abs().max())), flush=True)
    del la, lb
