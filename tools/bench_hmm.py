#!/usr/bin/env python3
"""HBM-roofline measurements of the exact-HMM stage kernels (BASELINE.json configs
[2] and [3]) through the C ABI.  Prints one JSON line per kernel: algorithmic bytes per step,
CUDA-event time, achieved GB/s and the fraction of the measured copy peak.

    python tools/bench_hmm.py [--B 1048576 --T 256 --K 3] [--scanB 64 --scanT 65536]
Inputs are synthetic row-normalised log_A and Gaussian log_b (resident in HBM, far larger than L2).
"""
import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402

import vhmm_b200 as V  # noqa: E402
from vhmm_b200 import functional as Fn  # noqa: E402


def peak_hbm():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        return json.load(open(p))["hbm_gbs"], "measured"
    return 6650.0, "fallback"


def timed(fn, iters=5, warmup=3):
    for _ in range(warmup):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters


def make_inputs(B, T, K, seed=0, ragged=False):
    g = torch.Generator(device="cuda").manual_seed(seed)
    log_pi = torch.log_softmax(torch.randn(K, device="cuda", generator=g), 0)
    log_A = torch.empty(B, T, K, K, device="cuda")
    chunk = max(1, (1 << 26) // (T * K * K))
    for b0 in range(0, B, chunk):
        n = min(chunk, B - b0)
        log_A[b0:b0 + n] = torch.log_softmax(torch.randn(n, T, K, K, device="cuda", generator=g), -1)
    log_b = -0.5 * torch.randn(B, T, K, device="cuda", generator=g) ** 2 - 1.0
    if ragged:
        L = torch.randint(T // 2, T + 1, (B,), device="cuda", generator=g)
    else:
        L = torch.full((B,), T, dtype=torch.int64, device="cuda")
    return log_pi, log_A, log_b, L


def report(name, bytes_per_step, steps, ms, pk, src, extra=None):
    gbs = bytes_per_step * steps / (ms * 1e-3) / 1e9
    line = {"kernel": name, "bytes_per_step": bytes_per_step, "steps": steps, "ms": ms, "steps_per_s": steps / (ms * 1e-3),
            "roofline": {"bound": "hbm", "achieved": gbs, "peak": pk, "unit": "GB/s", "frac": gbs / pk, "peak_source": src}}
    if extra:
        line.update(extra)
    print(json.dumps(line), flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--B", type=int, default=1 << 20)
    ap.add_argument("--T", type=int, default=256)
    ap.add_argument("--K", type=int, default=3)
    ap.add_argument("--scanB", type=int, default=64)
    ap.add_argument("--scanT", type=int, default=65536)
    ap.add_argument("--ragged", action="store_true")
    a = ap.parse_args()
    torch.cuda.set_device(0)
    pk, src = peak_hbm()
    K = a.K
    log_pi, log_A, log_b, L = make_inputs(a.B, a.T, K, ragged=a.ragged)
    steps = int(L.sum())
    cfg = {"config": f"B={a.B} T={a.T} K={K} ragged={a.ragged}"}
    inb = 4 * (K * K + K)
    report("hmm_forward (loglik)", inb, steps, timed(lambda: Fn.hmm_forward(log_pi, log_A, log_b, L)), pk, src, cfg)
    report("viterbi (uint8 path)", inb + 1, steps, timed(lambda: Fn.viterbi(log_pi, log_A, log_b, L)), pk, src, cfg)
    report("hmm_decode (viterbi path + gamma)", inb + 4 * K + 1, steps,
           timed(lambda: Fn.hmm_decode(log_pi, log_A, log_b, L)), pk, src, cfg)
    report("hmm_posterior (gamma)", inb + 4 * K, steps,
           timed(lambda: Fn.hmm_posterior(log_pi, log_A, log_b, L, want_xi=False)), pk, src, cfg)
    if a.B * a.T * K * K * 4 < 40e9:
        report("hmm_posterior (gamma, xi = beta-recursion gradient)", 2 * inb, steps,
               timed(lambda: Fn.hmm_posterior(log_pi, log_A, log_b, L, want_xi=True), iters=3, warmup=2), pk, src, cfg)
    del log_A, log_b
    Fn.release_workspaces()
    torch.cuda.empty_cache()
    log_pi, log_A, log_b, L = make_inputs(a.scanB, a.scanT, K, seed=1)
    steps = a.scanB * a.scanT
    cfg = {"config": f"B={a.scanB} T={a.scanT} K={K} (long sequences, small batch)"}
    report("hmm_scan (log-semiring parallel scan, loglik)", inb, steps, timed(lambda: Fn.hmm_scan(log_pi, log_A, log_b, L)), pk, src, cfg)
    report("hmm_scan (all-prefix alpha)", inb + 4 * K, steps,
           timed(lambda: Fn.hmm_scan(log_pi, log_A, log_b, L, want_alpha=True)), pk, src, cfg)
    report("hmm_forward sequential (same inputs, for comparison)", inb, steps,
           timed(lambda: Fn.hmm_forward(log_pi, log_A, log_b, L, time_parallel=False)), pk, src, cfg)


if __name__ == "__main__":
    main()
