#!/usr/bin/env python3
"""vhmm_viterbi for 4 < K <= 32 on few long sequences (VHMM_WIDE_SCAN=off: the warp-per-sequence kernel with one step in flight)"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from vhmm_b200 import functional as Fn
g = torch.Generator(device="cuda").manual_seed(0)
for (B, T, K) in ((4, 16384, 32), (64, 2048, 32), (4, 16384, 16), (4, 16384, 8)):
    la = torch.log_softmax(torch.randn(B, T, K, K, device="cuda", generator=g), -1)
    lb = torch.randn(B, T, K, device="cuda", generator=g) * 2 - 3
    lp = torch.log_softmax(torch.randn(K, device="cuda", generator=g), -1)
    L = torch.full((B,), T, dtype=torch.int64, device="cuda")
    for _ in range(2): out = Fn.viterbi(lp, la, lb, L)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(3): out = Fn.viterbi(lp, la, lb, L)
    e1.record(); torch.cuda.synchronize()
    print(os.environ.get("VHMM_WIDE_SCAN", "on"), "B=%d T=%d K=%d viterbi %.3f ms  checksum %d score %.4f" %
          (B, T, K, e0.elapsed_time(e1) / 3, int(out[0].to(torch.int64).sum()), float(out[1][0])), flush=True)
