#!/usr/bin/env python3
"""GPU-side diagnostic: fused vs layer-wise vs golden (float64 reference) for the ELBO path.
Prints per-tensor norm-wise relative errors and a quick timing; asserts nothing.
    python tools/fused_check.py [--big]"""
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import torch  # noqa: E402

import vhmm_b200 as V  # noqa: E402
from vhmm_b200 import functional as Fn  # noqa: E402
from conftest import load_golden, rel_err  # noqa: E402


def run(model, x, u, L, beta, impl):
    Fn.elbo_select(impl)
    model.zero_grad(set_to_none=True)
    loss = model.compute_loss(x, u, L, beta)
    loss.backward()
    torch.cuda.synchronize()
    g = {k: p.grad.detach().double().cpu() for k, p in model.named_parameters()}
    terms = [float(t) for t in model.loss_terms(x, u, L)]
    return float(loss), terms, g


def main():
    torch.cuda.set_device(0)
    for name in ("base_fresh", "base_ckpt", "base_collate"):
        g = load_golden(name)
        d = g["dims"]
        m = V.VAE_HMM(d["C"], d["H"], d["K"], d["H2"], u_dim=d["U"], trans_hidden=d["TH"])
        m.load_state_dict(g["params"])
        m = m.cuda()
        x, u, L = g["x"].cuda(), g["u"].cuda(), g["lengths"]
        ref = float(g["raw"]["loss64"])
        print(f"== {name}: x{tuple(x.shape)} lengths[{int(L.min())}..{int(L.max())}] beta={g['beta']} ref loss {ref:.9g}")
        res = {}
        for impl in ("layerwise", "fused"):
            try:
                res[impl] = run(m, x, u, L, g["beta"], impl)
            except Exception as e:  # noqa: BLE001
                print(f"   {impl}: FAILED {e}")
                continue
            loss, terms, gr = res[impl]
            print(f"   {impl:9s} loss {loss:.9g} rel {abs(loss-ref)/abs(ref):.2e}  terms {terms}")
        for k in g["grads64"]:
            row = [f"{k:36s}"]
            for impl in ("layerwise", "fused"):
                if impl in res:
                    row.append(f"{impl[:5]} {rel_err(res[impl][2][k], g['grads64'][k]):.2e}")
            print("   " + "  ".join(row))
    # larger random case: fused vs layerwise, ragged lengths, and timing
    torch.manual_seed(0)
    m = V.VAE_HMM(5, 64, 3, 32, u_dim=4).cuda()
    for (B, T, ragged) in ((64, 512, False), (96, 333, True), (4096, 512, False)):
        gen = torch.Generator(device="cuda").manual_seed(7)
        x = torch.randn(B, 5, T, device="cuda", generator=gen)
        u = torch.randn(B, 4, T, device="cuda", generator=gen)
        if ragged:
            L = torch.randint(1, T + 1, (B,))
            L[0] = T
            for b in range(B):
                x[b, :, int(L[b]):] = 0
                u[b, :, int(L[b]):] = 0
        else:
            L = torch.full((B,), T, dtype=torch.int64)
        out = {}
        for impl in ("layerwise", "fused"):
            out[impl] = run(m, x, u, L, 0.7, impl)
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            for _ in range(3):
                run(m, x, u, L, 0.7, impl)
            dt = (time.perf_counter() - t0) / 3
            print(f"== B={B} T={T} ragged={ragged} {impl:9s} loss {out[impl][0]:.9g}  {dt*1e3:.2f} ms/step (incl. loss_terms) -> {B*T/dt:.3e} steps/s")
        la, fu = out["layerwise"], out["fused"]
        print(f"   loss rel diff {abs(la[0]-fu[0])/abs(la[0]):.2e}; terms {la[1]} vs {fu[1]}")
        worst = max((rel_err(fu[2][k], la[2][k]), k) for k in la[2])
        print(f"   worst grad rel diff fused vs layerwise: {worst[0]:.2e} ({worst[1]})")
    Fn.elbo_select("auto")


if __name__ == "__main__":
    main()
