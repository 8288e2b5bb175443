#!/usr/bin/env python3
"""cfg-3 pipeline from raw (x, u): Prior (u -> log_A), emission table + log-probabilities (x -> log_b), then
decode (Viterbi path + gamma) -- the model-level calls get_regime_probs / viterbi / decode_regimes make."""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import torch
import vhmm_b200 as V
from vhmm_b200 import functional as Fn
from bench_hmm import timed
B = int(sys.argv[1]) if len(sys.argv) > 1 else 1 << 20
T = 256
torch.manual_seed(0)
m = V.VAE_HMM(5, 64, 3, 32, u_dim=4).cuda().eval()
g = torch.Generator(device="cuda").manual_seed(1)
x = torch.randn(B, 5, T, device="cuda", generator=g); u = torch.randn(B, 4, T, device="cuda", generator=g)
L = torch.full((B,), T, dtype=torch.int64, device="cuda")
with torch.no_grad():
    t_prior = timed(lambda: m.prior(u), iters=3, warmup=1)
    mu_tab, lv_tab = Fn.emission_table(m._dims, m._params(), T, x.device)
    t_emis = timed(lambda: Fn.emission_logprob(x, mu_tab, lv_tab), iters=3, warmup=1)
    t_inputs = timed(lambda: m.hmm_inputs(x, u), iters=3, warmup=1)
    log_pi, log_A, log_b = m.hmm_inputs(x, u)
    t_dec = timed(lambda: Fn.hmm_decode(log_pi, log_A, log_b, L), iters=3, warmup=1)
    t_all = timed(lambda: m.decode_regimes(x, u, L), iters=3, warmup=1)
    t_ll2 = timed(lambda: Fn.hmm_forward(*m.hmm_inputs(x, u), L), iters=3, warmup=1)       # front end + recursion
    t_ll1 = timed(lambda: m.hmm_log_likelihood(x, u, L), iters=3, warmup=1)                 # one fused kernel
    d = float((m.hmm_log_likelihood(x, u, L) - Fn.hmm_forward(*m.hmm_inputs(x, u), L)).abs().max())
steps = B * T
print(json.dumps({"config": f"cfg3 pipeline B={B} T={T} K=3 from raw x,u", "prior_ms": t_prior, "emission_logprob_ms": t_emis,
                  "hmm_inputs_ms": t_inputs, "decode_ms": t_dec, "decode_regimes_ms": t_all, "loglik_two_stage_ms": t_ll2,
                  "loglik_fused_ms": t_ll1, "loglik_fused_steps_per_s": B * T / (t_ll1 * 1e-3), "loglik_max_abs_diff": d, "steps_per_s_end_to_end": steps / (t_all * 1e-3)}))
