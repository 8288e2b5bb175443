#!/usr/bin/env python3
"""e2e step time (pinned host batch -> loss on the host) against the number of streaming chunks."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import vhmm_b200 as V
from vhmm_b200 import functional as Fn
torch.manual_seed(0)
m = V.VAE_HMM(5, 64, 3, 32, u_dim=4).cuda()
B, T = 65536, 512
xh = torch.randn(B, 5, T).pin_memory(); uh = torch.randn(B, 4, T).pin_memory()
L = torch.full((B,), T, dtype=torch.int64); Ld = L.cuda()
norms = Fn.elbo_norms(Ld, B, 5, T)
for nch in (0, 8, 16):
    chunk = -(-B // nch) if nch else 0     # 0: the default geometric schedule
    def step():
        s, g = Fn.elbo_raw_streamed(m._dims, m._params(), xh, uh, Ld, norms, 1.0, True, chunk=chunk)
        return float(Fn.loss_from_sums(s, norms, 1.0))
    for _ in range(2): step()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(4): step()
    e1.record(); torch.cuda.synchronize()
    print(nch, "chunks:", e0.elapsed_time(e1) / 4, "ms")
