#!/usr/bin/env python3
"""time vhmm_viterbi / vhmm_hmm_decode on few long sequences (B=64, T=65536 and smaller); VHMM_VITERBI=sweep = thread per sequence"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from vhmm_b200 import functional as Fn
K = 3
g = torch.Generator(device="cuda").manual_seed(0)
for (B, T) in ((64, 65536), (8, 65536), (1, 65536), (512, 4096), (2048, 1024), (64, 256)):
    la = torch.log_softmax(torch.randn(B, T, K, K, device="cuda", generator=g), -1)
    lb = torch.randn(B, T, K, device="cuda", generator=g)
    lp = torch.log_softmax(torch.randn(K, device="cuda", generator=g), -1)
    L = torch.full((B,), T, dtype=torch.int64, device="cuda")
    for name, fn in (("viterbi", lambda: Fn.viterbi(lp, la, lb, L)), ("decode", lambda: Fn.hmm_decode(lp, la, lb, L))):
        for _ in range(2): fn()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(5): out = fn()
        e1.record(); torch.cuda.synchronize()
        print(os.environ.get("VHMM_VITERBI", "auto"), "B=%d T=%d" % (B, T), name, "%.3f ms" % (e0.elapsed_time(e1) / 5),
              "path checksum", int(out[0].to(torch.int64).sum()), flush=True)
