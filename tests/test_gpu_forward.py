"""The fused forward-only tile kernel (csrc/forward_tc5.cuh: encode / decode / forward of the base config in ONE
launch, conv2 layers on tcgen05, no activation in HBM) against the float64 oracle (pinned to the reference by the
golden fixtures), against the layer-wise kernels, and on the shapes the reference's call sites use."""
import numpy as np
import pytest
import torch

from conftest import maxabs_rel

pytestmark = pytest.mark.gpu
TOL = 1e-5
DIMS = (5, 64, 3, 32, 4, 128)


def _model(seed=0, ckpt=False):
    import vhmm_b200 as V
    from oracle import elbo as O
    from conftest import load_golden
    p = load_golden("base_ckpt")["params"] if ckpt else O.init_params(*DIMS, seed=seed)
    m = V.VAE_HMM(5, 64, 3, 32, u_dim=4)
    m.load_state_dict(p)
    return m.cuda(), {k: v.double() for k, v in p.items()}


@pytest.mark.parametrize("B,T", [(1, 1), (2, 3), (3, 57), (2, 120), (2, 121), (5, 124), (3, 240), (2, 241), (7, 512), (2, 1000)])
@pytest.mark.parametrize("ckpt", [False, True])
@pytest.mark.parametrize("tma", [False, True])
def test_fused_forward_vs_float64_oracle(B, T, ckpt, tma):
    """tile boundaries (T = 120, 121, 240, 241), windows shorter than the halo (T = 1, 3), T % 4 != 0 (scalar cp.async)"""
    from vhmm_b200 import functional as Fn
    from oracle import elbo as O
    m, p64 = _model(seed=B + T, ckpt=ckpt)
    assert Fn.elbo_active_impl(m._dims) == "fused"
    Fn.tma_select(tma)      # input windows by tensor-map TMA (where T % 4 == 0) or cp.async
    g = torch.Generator().manual_seed(B * 1000 + T)
    x = torch.randn(B, 5, T, generator=g) * 1.7
    (mu64, lv64), q64 = O.forward(p64, x.double())
    lg64 = O.encode(p64, x.double())
    n0 = Fn._lib.lib().vhmm_launch_count()
    with torch.no_grad():
        (mu, lv), q = m(x.cuda())
    assert Fn._lib.lib().vhmm_launch_count() - n0 == 2          # fold + ONE tile kernel
    with torch.no_grad():
        logits = m.encode(x.cuda())
        q2 = m.regime_probs_meanfield(x.cuda())
        mu2, lv2 = m.decode(q)
    assert maxabs_rel(logits.cpu(), lg64) < TOL
    assert maxabs_rel(q.cpu(), q64) < TOL
    assert maxabs_rel(mu.cpu(), mu64) < TOL
    assert maxabs_rel(lv.cpu(), lv64) < TOL
    assert torch.equal(q, q2)
    Fn.tma_select(False)
    assert torch.equal(mu, mu2) and torch.equal(lv, lv2)        # decode(q) of a probability-valued q == forward()


def test_fused_forward_equals_layerwise_kernels_and_decodes_any_q():
    from vhmm_b200 import functional as Fn
    from oracle import elbo as O
    m, p64 = _model(seed=3)
    g = torch.Generator().manual_seed(5)
    x = torch.randn(6, 5, 333, generator=g)
    qa = torch.randn(6, 3, 333, generator=g) * 4.0               # decode() takes any (B,K,T), not only probabilities
    with torch.no_grad():
        (mu, lv), q = m(x.cuda())
        mua, lva = m.decode(qa.cuda())
        Fn.elbo_select("layerwise")
        try:
            (mu_l, lv_l), q_l = m(x.cuda())
            mua_l, lva_l = m.decode(qa.cuda())
        finally:
            Fn.elbo_select("auto")
    assert maxabs_rel(q.cpu(), q_l.cpu()) < TOL and maxabs_rel(mu.cpu(), mu_l.cpu()) < TOL and maxabs_rel(lv.cpu(), lv_l.cpu()) < TOL
    mua64, lva64 = O.decode(p64, qa.double())
    assert maxabs_rel(mua.cpu(), mua64) < TOL and maxabs_rel(lva.cpu(), lva64) < TOL
    assert maxabs_rel(mua_l.cpu(), mua64) < TOL


def test_fused_forward_large_batch_properties():
    """cfg-2-sized inference (B reduced to 4096, T = 512): q rows sum to 1, batch-permutation equivariance,
    the concatenation of two half batches equals the full batch bit for bit (tiles are independent)."""
    m, _ = _model(seed=9)
    g = torch.Generator(device="cuda").manual_seed(2)
    x = torch.randn(4096, 5, 512, device="cuda", generator=g)
    with torch.no_grad():
        (mu, lv), q = m(x)
        (mu_a, lv_a), q_a = m(x[:2048])
        (mu_b, lv_b), q_b = m(x[2048:])
        perm = torch.randperm(4096, device="cuda", generator=g)
        (mu_p, _), q_p = m(x[perm])
    assert float((q.sum(1) - 1).abs().max()) < 1e-6
    assert torch.equal(torch.cat([mu_a, mu_b]), mu) and torch.equal(torch.cat([q_a, q_b]), q) and torch.equal(torch.cat([lv_a, lv_b]), lv)
    assert torch.equal(mu_p, mu[perm]) and torch.equal(q_p, q[perm])
