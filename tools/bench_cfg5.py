#!/usr/bin/env python3
"""BASELINE.json configs[4] (scaled model K=32, hidden_dim=256, hidden_dim2=128 (assumed), u_dim=16,
B=32768, T=1024) on the layer-wise path (the fused tile kernels are instantiated for the base dims only)."""
import json, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import vhmm_b200 as V
B = int(sys.argv[1]) if len(sys.argv) > 1 else 32768
conv = sys.argv[2] if len(sys.argv) > 2 else "auto"
from vhmm_b200 import functional as Fn
Fn.conv_select(conv)
T = 1024
torch.manual_seed(0)
m = V.VAE_HMM(5, 256, 32, 128, u_dim=16).cuda()
g = torch.Generator(device="cuda").manual_seed(1)
x = torch.randn(B, 5, T, device="cuda", generator=g); u = torch.randn(B, 16, T, device="cuda", generator=g)
L = torch.full((B,), T, dtype=torch.int64)
def step():
    m.zero_grad(set_to_none=True)
    loss = m.compute_loss(x, u, L, 1.0); loss.backward(); return loss
step(); torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from bench import ClockSampler
smp = ClockSampler(0, interval=0.02); smp.start()
n_it = 3
e0.record()
for _ in range(n_it):
    loss = step()
e1.record(); torch.cuda.synchronize()
clocks = smp.stop()
ms = e0.elapsed_time(e1) / n_it
flop = 3872256
print(json.dumps({"config": f"cfg5 scaled: VAE_HMM(5,256,K=32,128,u_dim=16) B={B} T={T}, layer-wise path, layer kernels: {conv}", "ms_per_step": ms, "clocks": clocks,
                  "sequence_steps_per_s": B * T / (ms * 1e-3), "algorithmic_tflops": B * T * flop / (ms * 1e-3) / 1e12,
                  "loss": float(loss.detach()), "peak_mem_gb": torch.cuda.max_memory_allocated() / 1e9}))
