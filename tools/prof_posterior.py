#!/usr/bin/env python3
"""One call of each posterior build on a cfg-3-shaped batch (for ncu):  python tools/prof_posterior.py [B] [T] [xi]"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from vhmm_b200 import functional as Fn
from bench_hmm import make_inputs

B = int(sys.argv[1]) if len(sys.argv) > 1 else 1 << 18
T = int(sys.argv[2]) if len(sys.argv) > 2 else 256
xi = len(sys.argv) > 3 and sys.argv[3] == "xi"
log_pi, log_A, log_b, L = make_inputs(B, T, 3)
for _ in range(2):
    Fn.hmm_posterior(log_pi, log_A, log_b, L, want_xi=xi)
torch.cuda.synchronize()
