"""Parity at BASELINE.json's FULL sizes through a size-independent property: replication.  A batch made of N copies of a small
ragged base batch has the base batch's loss (every term of the ELBO is a mean: VQ_VAE_HMM_fixed.py:120, :131, :135) and the base
batch's gradients, and every copy decodes to the base sequence's path / posterior -- and the base batch is small enough for the
CPU oracle.  So the full-size kernels (all tiles, all CTAs, every trip of the persistent loops) are checked against the oracle:
  configs[1]  ELBO fwd+bwd, B=65536, T=512 (fused tile kernels)          loss + 18 gradients, 1e-5
  configs[2]  Viterbi + posterior decode, B=2^20, T=256, K=3             paths bit-exact, gamma / log-lik 1e-5
  configs[3]  log-likelihood scan, B=64, T=65536                         1e-5 against the float64 forward recursion
  configs[4]  ELBO fwd+bwd of the scaled model, B=32768, T=1024, K=32    loss + 18 gradients, 1e-5 (tensor-core layer kernels)
"""
import numpy as np
import pytest
import torch

from conftest import rel_err
from oracle import elbo as O
from oracle import hmm as H

pytestmark = pytest.mark.gpu
TOL = 1e-5


def _elbo_replicated(dims, B0, T, copies, seed, beta):
    import vhmm_b200 as V
    C, Hd, K, H2, U, TH = dims
    p = O.init_params(C, Hd, K, H2, U, TH, seed=seed)
    x, u, L = O.synth_batch(B0, C, U, T, seed=seed + 1, ragged=True, min_len=T // 3)
    p64 = {k: v.double() for k, v in p.items()}
    ref, _, g64 = O.loss_and_grads(p64, x.double(), u.double(), L, U, beta)
    m = V.VAE_HMM(C, Hd, K, H2, u_dim=U, trans_hidden=TH)
    m.load_state_dict(p)
    m = m.cuda()
    xb = x.cuda().repeat(copies, 1, 1)
    ub = u.cuda().repeat(copies, 1, 1)
    Lb = L.repeat(copies)
    assert xb.shape[0] == B0 * copies
    loss = m.compute_loss(xb, ub, Lb, beta)
    loss.backward()
    torch.cuda.synchronize()
    assert abs(float(loss.detach()) - float(ref)) <= TOL * abs(float(ref)), (float(loss.detach()), float(ref))
    sd = dict(m.named_parameters())
    for k in O.PARAM_ORDER:
        e = rel_err(sd[k].grad.cpu(), g64[k])
        assert e <= TOL, (k, e)
    del xb, ub, m
    from vhmm_b200 import functional as Fn
    Fn.release_workspaces()
    torch.cuda.empty_cache()


def test_cfg2_elbo_full_batch_equals_oracle_of_the_replicated_base():
    """BASELINE.json configs[1]: B = 65536 = 1024 copies of 64 ragged sequences, T = 512 (327,680 tiles over 148 CTAs)."""
    _elbo_replicated((5, 64, 3, 32, 4, 128), B0=64, T=512, copies=1024, seed=11, beta=0.8)


def test_cfg5_elbo_full_batch_equals_oracle_of_the_replicated_base():
    """BASELINE.json configs[4] (scaled model, hidden_dim2 = 128 assumed as in SURVEY 8): B = 32768 = 4096 copies of 8 ragged
    sequences, T = 1024: 16 chunks of 2048 sequences through the tensor-core layer kernels, the fused prior kernel and the
    folded decoder.conv1."""
    _elbo_replicated((5, 256, 32, 128, 16, 128), B0=8, T=1024, copies=4096, seed=21, beta=0.6)


def test_cfg3_decode_full_batch_every_copy_equals_the_oracle():
    """BASELINE.json configs[2]: B = 2^20 = 4096 copies of 256 ragged sequences (quantised scores: exact ties), T = 256, K = 3."""
    from vhmm_b200 import functional as Fn
    B0, T, K, copies = 256, 256, 3, 4096
    r = np.random.default_rng(5)
    lp = r.normal(size=K); lp -= np.log(np.exp(lp).sum())
    la = r.normal(size=(B0, T, K, K)); la -= np.log(np.exp(la).sum(-1, keepdims=True))
    lb = r.normal(size=(B0, T, K)) * 3.0
    lp, la, lb = (np.round(v * 2) / 2 for v in (lp, la, lb))
    L = r.integers(1, T + 1, size=B0); L[0] = T; L[1] = 1
    lp, la, lb, L = lp.astype(np.float32), la.astype(np.float32), lb.astype(np.float32), L.astype(np.int64)
    path64, _ = H.viterbi(lp, la, lb, L, dtype=np.float32)
    g64, _, ll64 = H.posteriors(lp, la, lb, L)
    tlp = torch.from_numpy(lp).cuda()
    tla = torch.from_numpy(la).cuda().repeat(copies, 1, 1, 1)
    tlb = torch.from_numpy(lb).cuda().repeat(copies, 1, 1)
    tL = torch.from_numpy(L).cuda().repeat(copies)
    assert tla.shape[0] == 1 << 20
    path, gamma, ll = Fn.hmm_decode(tlp, tla, tlb, tL)
    torch.cuda.synchronize()
    path = path.view(copies, B0, T)
    assert np.array_equal(path[0].cpu().numpy(), path64)                       # bit-exact against the oracle
    assert bool((path == path[0:1]).all())                                     # ... in every copy
    gamma = gamma.view(copies, B0, T, K)
    assert np.abs(gamma[0].cpu().numpy() - g64).max() <= TOL
    assert bool((gamma == gamma[0:1]).all())                                   # same arithmetic wherever a sequence lands
    ll = ll.view(copies, B0)
    assert np.allclose(ll[0].cpu().numpy(), ll64, rtol=TOL, atol=1e-5)
    assert bool((ll == ll[0:1]).all())
    pv, _ = Fn.viterbi(tlp, tla, tlb, tL)
    assert bool((pv.view(copies, B0, T) == path[0:1]).all())


def test_cfg4_scan_full_size_equals_float64_forward():
    """BASELINE.json configs[3]: B = 64 = 8 copies of 8 sequences, T = 65536 (lengths around the segment boundaries)."""
    from vhmm_b200 import functional as Fn
    B0, T, K, copies = 8, 65536, 3, 8
    g = torch.Generator().manual_seed(9)
    la = torch.log_softmax(torch.randn(B0, T, K, K, generator=g) * 1.5, dim=-1)
    lb = torch.randn(B0, T, K, generator=g) * 2 - 3
    lp = torch.log_softmax(torch.randn(K, generator=g), dim=-1)
    L = torch.tensor([T, T - 1, 40000, 256, 257, 1, 12345, 65280])
    _, ll64 = H.forward(lp.numpy(), la.numpy(), lb.numpy(), L.numpy())
    tla, tlb, tL = la.cuda().repeat(copies, 1, 1, 1), lb.cuda().repeat(copies, 1, 1), L.cuda().repeat(copies)
    ll = Fn.hmm_scan(lp.cuda(), tla, tlb, tL).view(copies, B0)
    assert np.allclose(ll[0].cpu().numpy(), ll64, rtol=TOL, atol=1e-5), (ll[0].cpu().numpy(), ll64)
    assert bool((ll == ll[0:1]).all())
    ll2, alpha = Fn.hmm_scan(lp.cuda(), tla, tlb, tL, want_alpha=True)
    assert np.allclose(ll2.view(copies, B0)[0].cpu().numpy(), ll64, rtol=TOL, atol=1e-5)
