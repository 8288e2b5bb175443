import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import vhmm_b200.functional as Fn
from oracle import hmm as H
r = np.random.default_rng(9)
for K, T in ((3, 1024), (4, 768), (3, 4096)):
    B = 6
    A = r.random((B, T, K, K)) + 0.05
    A[:3] *= np.triu(np.ones((K, K)))[None, None]
    A[3:] = 1e-6 * A[3:] + np.eye(K)[None, None]
    A /= A.sum(-1, keepdims=True)
    with np.errstate(divide="ignore"):
        la = np.log(A).astype(np.float32)
    lb = (r.normal(size=(B, T, K)) * 6.0 - 8.0).astype(np.float32)
    lp = np.log(np.full(K, 1.0 / K)).astype(np.float32)
    for L in (np.array([T, T - 3, 600, T, 513, T], np.int64), np.full(B, T, np.int64), np.full(B, T - 3, np.int64), np.full(B, T - 4, np.int64)):
        _, ll64 = H.forward(lp, la, lb, L)
        d = [torch.from_numpy(np.ascontiguousarray(v)).cuda() for v in (lp, la, lb, L)]
        ll_seq = Fn.hmm_forward(*d).cpu().numpy()
        ll_a, al = Fn.hmm_forward(*d, want_alpha=True)
        ll_scan = Fn.hmm_scan(*d).cpu().numpy()
        ll_chk = Fn.hmm_scan(*d, want_alpha=True)[0].cpu().numpy()
        ll_post = Fn.hmm_posterior(*d, want_xi=False)[0].cpu().numpy()
        print(K, T, L.tolist())
        print("  ll64 ", np.round(ll64, 2))
        print("  seq  ", np.round(ll_seq, 2)); print("  seq+a", np.round(ll_a.cpu().numpy(), 2))
        print("  scan ", np.round(ll_scan, 2)); print("  chunk", np.round(ll_chk, 2)); print("  post ", np.round(ll_post, 2))
