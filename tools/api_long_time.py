#!/usr/bin/env python3
"""every model-level call on FEW LONG sequences (B=64, T=65536 and B=4, T=16384, base dims): ms per call, to spot a path that
was built for B >> #SM.   python tools/api_long_time.py"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import vhmm_b200 as V
from vhmm_b200 import functional as Fn
torch.manual_seed(0)
def timed(fn, n=3):
    for _ in range(2): fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n
DIMS = {"base": (5, 64, 3, 32, 4, 128), "cfg5": (5, 256, 32, 128, 16, 128), "k4": (6, 48, 4, 24, 3, 64)}
which = sys.argv[1] if len(sys.argv) > 1 else "base"
C, Hd, K, H2, U, TH = DIMS[which]
for (B, T) in (((64, 65536), (4, 16384)) if which == "base" else ((4, 16384), (64, 2048))):
    m = V.VAE_HMM(C, Hd, K, H2, u_dim=U, trans_hidden=TH).cuda()
    g = torch.Generator(device="cuda").manual_seed(1)
    x = torch.randn(B, C, T, device="cuda", generator=g); u = torch.randn(B, U, T, device="cuda", generator=g)
    L = torch.full((B,), T, dtype=torch.int64, device="cuda")
    def nll():
        m.zero_grad(set_to_none=True); m.hmm_nll(x, u, L).backward()
    def elbo():
        m.zero_grad(set_to_none=True); m.compute_loss(x, u, L, 0.7).backward()
    calls = [("hmm_inputs", lambda: m.hmm_inputs(x, u)), ("hmm_log_likelihood", lambda: m.hmm_log_likelihood(x, u, L)),
             ("get_regime_probs", lambda: m.get_regime_probs(x, u, L)), ("viterbi", lambda: m.viterbi(x, u, L)),
             ("decode_regimes", lambda: m.decode_regimes(x, u, L)), ("regime_probs_meanfield", lambda: m.regime_probs_meanfield(x)),
             ("transition_probs", lambda: m.transition_probs(u))]
    with torch.no_grad():
        for name, fn in calls:
            try:
                print(which, "B=%d T=%d %-24s %.3f ms" % (B, T, name, timed(fn)), flush=True)
            except Exception as e:
                print("B=%d T=%d %-24s FAILED %s" % (B, T, name, str(e)[:120]), flush=True)
    for name, fn in (("hmm_nll fwd+bwd", nll), ("compute_loss fwd+bwd", elbo)):
        try:
            print(which, "B=%d T=%d %-24s %.3f ms" % (B, T, name, timed(fn)), flush=True)
        except Exception as e:
            print("B=%d T=%d %-24s FAILED %s" % (B, T, name, str(e)[:120]), flush=True)
    del x, u, m
    Fn.release_workspaces(); torch.cuda.empty_cache()
