"""ADVICE r1 (per-device state): ONE process that uses two GPUs in turn.  Function attributes (opt-in shared memory), SM counts,
occupancy-derived grids, the tensor-core weight caches and the static scratch of vhmm_viterbi_static are per device; a second
device must get its own.  Skipped on a one-GPU box (run with `gpurun --gpus 2`)."""
import numpy as np
import pytest
import torch

from oracle import elbo as O

pytestmark = pytest.mark.gpu


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs two GPUs in one process")
def test_same_results_on_cuda0_cuda1_cuda0():
    import vhmm_b200 as V
    from vhmm_b200 import functional as Fn, _lib
    p = O.init_params(5, 64, 3, 32, 4, 128, seed=2)
    x, u, L = O.synth_batch(9, 5, 4, 256, seed=3, ragged=True)
    r = np.random.default_rng(1)
    lp = torch.log_softmax(torch.from_numpy(r.normal(size=3).astype(np.float32)), -1)
    la = torch.log_softmax(torch.from_numpy(r.normal(size=(40, 64, 3, 3)).astype(np.float32)), -1)
    lb = torch.from_numpy((r.normal(size=(40, 64, 3)) * 3).astype(np.float32))
    Lh = torch.from_numpy(r.integers(1, 65, size=40).astype(np.int64))
    out = []
    for dev in ("cuda:0", "cuda:1", "cuda:0"):
        d = torch.device(dev)
        m = V.VAE_HMM(5, 64, 3, 32, u_dim=4)
        m.load_state_dict(p)
        m = m.to(d)
        loss = m.compute_loss(x.to(d), u.to(d), L, 0.7)              # fused tile kernels (226 KB of opt-in shared memory)
        loss.backward()
        with torch.no_grad():
            (mu, lv), q = m(x.to(d))                                 # fused forward kernel
        dlp, dla, dlb, dL = lp.to(d), la.to(d), lb.to(d), Lh.to(d)     # (kept alive: the C-ABI call below takes raw pointers)
        path, gamma, ll = Fn.hmm_decode(dlp, dla, dlb, dL)
        # allocation-free Viterbi: the static scratch array has one device address PER device
        with torch.cuda.device(d):
            sp = torch.full((40, 64), 9, dtype=torch.uint8, device=d)
            ss = torch.empty(40, device=d)
            _lib.check(V.lib().vhmm_viterbi_static(Fn._ptr(dlp), Fn._ptr(dla), Fn._ptr(dlb), Fn._ptr(dL), 40, 64, 3,
                                                   Fn._ptr(sp), Fn._ptr(ss), Fn._stream(d)), "vhmm_viterbi_static")
            torch.cuda.synchronize(d)
        assert torch.equal(sp.cpu(), path.cpu()), dev
        grads = torch.cat([t.grad.flatten().cpu() for t in m.parameters()])
        out.append((float(loss.detach()), grads, mu.cpu(), q.cpu(), path.cpu(), gamma.cpu(), ll.cpu()))
    for o in out[1:]:
        assert o[0] == out[0][0]
        for a, b in zip(o[1:], out[0][1:]):
            assert torch.equal(a, b)
    # the layer-wise path with the tensor-core layer kernels (TMA tensor maps, prepared weights) on the second device
    C, Hd, K, H2, U, TH = 5, 256, 32, 128, 16, 128
    p5 = O.init_params(C, Hd, K, H2, U, TH, seed=1)
    x5, u5, L5 = O.synth_batch(3, C, U, 256, seed=4, ragged=True, min_len=100)
    res = []
    for dev in ("cuda:1", "cuda:0"):
        d = torch.device(dev)
        m = V.VAE_HMM(C, Hd, K, H2, u_dim=U, trans_hidden=TH)
        m.load_state_dict(p5)
        m = m.to(d)
        loss = m.compute_loss(x5.to(d), u5.to(d), L5, 0.7)
        loss.backward()
        res.append((float(loss.detach()), torch.cat([t.grad.flatten().cpu() for t in m.parameters()])))
    assert res[0][0] == res[1][0] and torch.equal(res[0][1], res[1][1])
