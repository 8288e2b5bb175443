#!/usr/bin/env python3
"""scaled model (cfg-5 dims) on few long sequences: hmm_inputs and hmm_nll fwd+bwd, for a launch list"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import vhmm_b200 as V
torch.manual_seed(0)
m = V.VAE_HMM(5, 256, 32, 128, u_dim=16, trans_hidden=128).cuda()
B = int(sys.argv[1]) if len(sys.argv) > 1 else 4
T = int(sys.argv[2]) if len(sys.argv) > 2 else 16384
g = torch.Generator(device="cuda").manual_seed(1)
x = torch.randn(B, 5, T, device="cuda", generator=g); u = torch.randn(B, 16, T, device="cuda", generator=g)
L = torch.full((B,), T, dtype=torch.int64)
def step():
    m.zero_grad(set_to_none=True)
    loss = m.hmm_nll(x, u, L); loss.backward(); return loss
for _ in range(2): step()
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(2): loss = step()
e1.record(); torch.cuda.synchronize()
print("B", B, "T", T, "hmm_nll fwd+bwd %.3f ms" % (e0.elapsed_time(e1) / 2), "loss", float(loss.detach()))
