#!/usr/bin/env python3
"""cfg-1 training (batch 64, chunks of 20..200 steps): a few dozen graph-replayed steps (for ncu launch lists)"""
import os, random, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import vhmm_b200 as V
dev = torch.device("cuda", 0)
g = torch.Generator().manual_seed(1234)
xs = list(torch.randn(40, 5, 400, generator=g)); us = list(torch.randn(40, 4, 400, generator=g))
corpus = V.DeviceCorpus(xs, us, device=dev)
torch.manual_seed(0)
m = V.VAE_HMM(5, 64, 3, 32, u_dim=4).to(dev)
opt = V.FusedAdam(m, lr=1e-3)
loader = V.DeviceChunkLoader(corpus, 64)
random.seed(0)
gs = V.training.GraphedTrainStep(m, corpus, 64, optimizer=opt)
n_ep = int(sys.argv[1]) if len(sys.argv) > 1 else 3
for ep in range(n_ep):
    gs.reset_epoch_totals()
    t0 = time.perf_counter(); nb = 0
    for tr in loader.iter_triples():
        gs.step(*tr, beta=1.0); nb += 1
    loss = gs.epoch_mean_loss()
    torch.cuda.synchronize()
    print("epoch", ep, "loss %.5f" % loss, "%.3f ms per batch" % ((time.perf_counter() - t0) * 1e3 / nb), flush=True)
