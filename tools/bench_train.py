#!/usr/bin/env python3
"""Measurements for SURVEY.md 8(f) rows f1 / f2 (one JSON line each):

  gather      vhmm_gather_chunks at cfg-2 size (B=65536, T=512, ragged lengths 256..512) from a
              corpus resident in HBM: achieved GB/s on the algorithmic bytes (4 (C+U) per valid step
              read + 4 (C+U) per padded step written) against the measured copy peak.
  train_cfg1  one epoch of BASELINE.json configs[0] (batch 64, chunks of 20..200 steps, 1000 items):
              host path  = RandomChunkDataset + collate_fn + compute_loss/backward + torch Adam +
                           loss.item() per batch (the reference's train_model structure on our kernels)
              fused path = DeviceChunkLoader + loss_and_flat_grads + vhmm_adam_step, one sync per epoch
  train_cfg2  the fused training step at cfg 2 (B=65536, T=512): gather + ELBO fwd+bwd + Adam.

    python tools/bench_train.py
"""
import json
import os
import random
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np  # noqa: E402
import torch  # noqa: E402

import vhmm_b200 as V  # noqa: E402
from bench_hmm import peak_hbm, timed  # noqa: E402


def main():
    torch.cuda.set_device(0)
    pk, src = peak_hbm()
    C, U = 5, 4
    # ---- f1: gather at cfg-2 size ----------------------------------------------------------
    g = torch.Generator(device="cuda").manual_seed(0)
    n_seq, S = 4096, 4096
    corpus = V.DeviceCorpus.__new__(V.DeviceCorpus)
    corpus.C, corpus.U, corpus.device = C, U, torch.device("cuda", 0)
    corpus.seq_lens = [S] * n_seq
    corpus.offsets = (np.arange(n_seq + 1) * S).astype(np.int64)
    corpus.S_tot = n_seq * S
    corpus.x_cat = torch.randn(C, n_seq * S, device="cuda", generator=g)
    corpus.u_cat = torch.randn(U, n_seq * S, device="cuda", generator=g)
    B, T = 65536, 512
    idx, ln, st = V.training.sample_triples_fast(corpus.seq_lens, B, 256, 512, np.random.default_rng(0))
    ln[0] = T
    col0 = torch.from_numpy(corpus.offsets[idx] + st).cuda()
    Ld = torch.from_numpy(ln).cuda()
    x = torch.empty(B, C, T, device="cuda"); u = torch.empty(B, U, T, device="cuda")
    from vhmm_b200 import functional as Fn, _lib

    def run_gather():
        _lib.check(V.lib().vhmm_gather_chunks(Fn._ptr(corpus.x_cat), Fn._ptr(corpus.u_cat), corpus.S_tot, C, U, Fn._ptr(col0),
                                              Fn._ptr(Ld), B, T, Fn._ptr(x), Fn._ptr(u), Fn._stream(x.device)), "gather")
    ms = timed(run_gather, iters=10)
    nbytes = 4 * (C + U) * (int(ln.sum()) + B * T)
    gbs = nbytes / (ms * 1e-3) / 1e9
    t0 = time.perf_counter()
    for _ in range(3):
        V.training.sample_triples_fast(corpus.seq_lens, B, 256, 512, np.random.default_rng(0))
    t_fast = (time.perf_counter() - t0) / 3 * 1e3
    t0 = time.perf_counter()
    V.training.sample_triples(corpus.seq_lens, B, 256, 512, random.Random(0))
    t_ref = (time.perf_counter() - t0) * 1e3
    print(json.dumps({"kernel": "vhmm_gather_chunks (f1)", "config": f"B={B} T={T} lengths 256..512, corpus {n_seq}x{S}",
                      "ms": ms, "bytes": nbytes, "steps_per_s": int(ln.sum()) / (ms * 1e-3),
                      "roofline": {"bound": "hbm", "achieved": gbs, "peak": pk, "unit": "GB/s", "frac": gbs / pk, "peak_source": src},
                      "host_sampling_ms": {"reference_order_python_random": t_ref, "vectorised_numpy": t_fast}}), flush=True)
    # ---- f2: cfg-2 fused training step ---------------------------------------------------
    torch.manual_seed(0)
    model = V.VAE_HMM(C, 64, 3, 32, u_dim=U).cuda()
    opt = V.FusedAdam(model, lr=1e-3)
    rng = np.random.default_rng(1)

    def step_cfg2():
        xb, ub, L = corpus.sample_batch_fast(B, 256, 512, rng)
        loss, flat = model.loss_and_flat_grads(xb, ub, L, 1.0)
        opt.step(flat)
        return L
    for _ in range(2):
        step_cfg2()
    torch.cuda.synchronize()
    t0 = time.perf_counter(); nsteps = 0
    for _ in range(5):
        nsteps += int(step_cfg2().sum())
    torch.cuda.synchronize()
    dt = (time.perf_counter() - t0) / 5
    print(json.dumps({"step": "train_cfg2 fused (sample + gather + ELBO fwd+bwd + Adam)", "config": f"B={B} T<=512 (256..512)",
                      "ms_per_step": dt * 1e3, "sequence_steps_per_s": nsteps / 5 / dt}), flush=True)
    del corpus, x, u
    torch.cuda.empty_cache()
    # ---- f2: cfg-1 epoch, host path vs fused path ------------------------------------------
    gcpu = torch.Generator().manual_seed(1234)
    xs = torch.randn(40, C, 400, generator=gcpu); us = torch.randn(40, U, 400, generator=gcpu)
    ds = V.RandomChunkDataset(xs, us, min_len=20, max_len=200)
    dl = torch.utils.data.DataLoader(ds, batch_size=64, collate_fn=V.collate_fn)

    def epoch_host(m, o):
        tot, steps = 0.0, 0
        for xb, ub, L in dl:
            o.zero_grad()
            loss = m.compute_loss(xb, ub, L, 1.0)
            loss.backward()
            o.step()
            tot += loss.item()
            steps += int(L.sum())
        return steps

    def epoch_fused(m, o, loader):
        tot = torch.zeros((), device="cuda", dtype=torch.float64); steps = 0
        for xb, ub, L in loader:
            loss, flat = m.loss_and_flat_grads(xb, ub, L, 1.0)
            o.step(flat)
            tot += loss
            steps += int(L.sum())
        float(tot)
        return steps
    res = {}
    for name in ("host", "fused"):
        torch.manual_seed(0)
        m = V.VAE_HMM(C, 64, 3, 32, u_dim=U).cuda()
        random.seed(0)
        if name == "host":
            o = torch.optim.Adam(m.parameters(), lr=1e-3)
            fn = lambda: epoch_host(m, o)     # noqa: E731
        else:
            o = V.FusedAdam(m, lr=1e-3)
            loader = V.DeviceChunkLoader(V.DeviceCorpus(list(xs), list(us)), 64)
            fn = lambda: epoch_fused(m, o, loader)   # noqa: E731
        fn(); torch.cuda.synchronize()
        t0 = time.perf_counter(); steps = 0
        for _ in range(5):
            steps += fn()
        torch.cuda.synchronize()
        dt = (time.perf_counter() - t0) / 5
        res[name] = {"ms_per_epoch": dt * 1e3, "ms_per_batch": dt * 1e3 / 16, "sequence_steps_per_s": steps / 5 / dt}
    print(json.dumps({"step": "train_cfg1 epoch (16 batches of <=64 items, chunks 20..200)", **res,
                      "speedup_fused_vs_host": res["host"]["ms_per_epoch"] / res["fused"]["ms_per_epoch"]}), flush=True)


if __name__ == "__main__":
    main()
