#!/usr/bin/env python3
"""time vhmm_viterbi / vhmm_hmm_decode at BASELINE.json configs[2] (B=2^20, T=256, K=3); VHMM_LIB selects an experimental build"""
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
ref = None
for name, fn in (("viterbi", lambda: Fn.viterbi(lp, la, lb, L)), ("decode", lambda: Fn.hmm_decode(lp, la, lb, L))):
    for _ in range(3): fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(20): out = fn()
    e1.record(); torch.cuda.synchronize()
    print(os.environ.get("VHMM_LIB", "default").split("/")[-1], name, "%.3f ms" % (e0.elapsed_time(e1) / 20),
          "path checksum", int(out[0].to(torch.int64).sum()), flush=True)
