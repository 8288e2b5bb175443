#!/usr/bin/env python3
"""model.hmm_log_likelihood / get_regime_probs on FEW LONG sequences (B=64, T=65536, base dims)"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import vhmm_b200 as V
from vhmm_b200 import functional as Fn
torch.manual_seed(0)
m = V.VAE_HMM(5, 64, 3, 32, u_dim=4).cuda().eval()
B, T = 64, 65536
g = torch.Generator(device="cuda").manual_seed(1)
x = torch.randn(B, 5, T, device="cuda", generator=g); u = torch.randn(B, 4, T, device="cuda", generator=g)
L = torch.full((B,), T, dtype=torch.int64, device="cuda")
def timed(fn, n=5):
    for _ in range(2): fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n): out = fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n, out
with torch.no_grad():
    ms, ll = timed(lambda: m.hmm_log_likelihood(x, u, L))
    print("hmm_log_likelihood %.3f ms" % ms, float(ll[0]))
    mu_tab, lv_tab = Fn.emission_table(m._dims, m._params(), T, x.device)
    log_pi = torch.log_softmax(m.prior.log_prior.detach().float(), dim=0)
    ms2, ll2 = timed(lambda: Fn.hmm_loglik_fused(m._dims, m._params(), log_pi, x, u, L, mu_tab, lv_tab), 2)
    print("fused warp-per-sequence kernel %.3f ms" % ms2, float(ll2[0]), "max |diff| %.3e" % float((ll - ll2).abs().max()))
    ms3, g3 = timed(lambda: m.get_regime_probs(x, u, L))
    print("get_regime_probs %.3f ms" % ms3)
