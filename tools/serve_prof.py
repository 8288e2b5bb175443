#!/usr/bin/env python3
"""serving path: one (C,T) request through serving.infer -- latency per request and the kernels behind it"""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import vhmm_b200 as V
from vhmm_b200 import serving
torch.manual_seed(0)
m = V.VAE_HMM(5, 64, 3, 32, u_dim=4).cuda().eval()
for T in (60, 512):
    x = torch.randn(5, T).tolist()
    for _ in range(20): serving.infer(m, x)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    n = 200
    for _ in range(n): serving.infer(m, x)
    torch.cuda.synchronize()
    print("T", T, "infer: %.1f us per request (wall, incl. list <-> tensor conversion)" % ((time.perf_counter() - t0) * 1e6 / n), flush=True)
    xt = torch.tensor(x).unsqueeze(0).cuda()
    with torch.no_grad():
        for _ in range(20): m(xt)
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        for _ in range(n): m(xt)
        torch.cuda.synchronize()
    print("T", T, "model(x) on a resident tensor: %.1f us per call" % ((time.perf_counter() - t0) * 1e6 / n), flush=True)
