"""Known-answer tests for the recursion oracle (oracle/hmm.py): brute-force enumeration over
all K^T paths.  'parity unpinned' vs the reference, which has no such code."""
import numpy as np
import pytest

from oracle import hmm as H


def rand_hmm(B, T, K, seed, rng_scale=1.0):
    r = np.random.default_rng(seed)
    lp = r.normal(size=K)
    lp = lp - np.log(np.exp(lp).sum())
    la = r.normal(size=(B, T, K, K)) * rng_scale
    la = la - np.log(np.exp(la).sum(-1, keepdims=True))
    lb = r.normal(size=(B, T, K)) * 3.0
    return lp, la, lb


@pytest.mark.parametrize("K,T", [(2, 1), (2, 5), (3, 6), (3, 8), (4, 5)])
def test_forward_backward_vs_bruteforce(K, T):
    lp, la, lb = rand_hmm(2, T, K, seed=K * 100 + T)
    L = np.array([T, T])
    gamma, xi, ll = H.posteriors(lp, la, lb, L)
    for b in range(2):
        ll_bf, gamma_bf, path_bf, best, n_best = H.brute_force(lp, la[b], lb[b])
        assert abs(ll[b] - ll_bf) < 1e-12 * max(1, abs(ll_bf))
        assert np.allclose(gamma[b], gamma_bf, atol=1e-12)
        path, score = H.viterbi(lp, la[b:b + 1], lb[b:b + 1], L[b:b + 1], dtype=np.float64)
        assert abs(score[0] - best) < 1e-12 * max(1, abs(best))
        if n_best == 1:
            assert (path[0] == path_bf).all()
    assert np.allclose(gamma.sum(-1), 1.0)
    if T > 1:
        assert np.allclose(xi[:, 1:].sum((-1, -2)), 1.0)
        assert np.allclose(xi[:, 1:].sum(-2), gamma[:, 1:])


def test_ragged_equals_truncated():
    K, T = 3, 9
    lp, la, lb = rand_hmm(4, T, K, seed=5)
    L = np.array([9, 4, 1, 7])
    gamma, xi, ll = H.posteriors(lp, la, lb, L)
    path, score = H.viterbi(lp, la, lb, L)
    sc = H.scan_loglik(lp, la, lb, L, chunk=3)
    for b in range(4):
        n = L[b]
        g2, x2, l2 = H.posteriors(lp, la[b:b + 1, :n], lb[b:b + 1, :n], np.array([n]))
        assert abs(ll[b] - l2[0]) < 1e-12
        assert np.allclose(gamma[b, :n], g2[0]) and (gamma[b, n:] == 0).all()
        assert np.allclose(xi[b, :n], x2[0]) and (xi[b, n:] == 0).all()
        p2, s2 = H.viterbi(lp, la[b:b + 1, :n], lb[b:b + 1, :n], np.array([n]))
        assert (path[b, :n] == p2[0]).all() and (path[b, n:] == H.PAD_STATE).all()
        assert abs(sc[b] - ll[b]) < 1e-10


def test_viterbi_lowest_index_tiebreak():
    # all-equal scores: every path ties; the rule gives the all-zero path
    K, T = 3, 6
    lp = np.full(K, -np.log(K))
    la = np.full((1, T, K, K), -np.log(K))
    lb = np.zeros((1, T, K))
    path, _ = H.viterbi(lp, la, lb, np.array([T]))
    assert (path == 0).all()
    # constructed tie at the last step between states 1 and 2 -> state 1; back-pointer tie -> lowest
    lb2 = lb.copy()
    lb2[0, T - 1] = [-5.0, 1.0, 1.0]
    path, _ = H.viterbi(lp, la, lb2, np.array([T]))
    assert path[0, -1] == 1 and (path[0, :-1] == 0).all()


def test_gamma_is_gradient_of_loglik():
    # finite differences: d loglik / d log_b = gamma, d loglik / d log_A = xi
    K, T = 3, 5
    lp, la, lb = rand_hmm(1, T, K, seed=11)
    L = np.array([T])
    gamma, xi, ll = H.posteriors(lp, la, lb, L)
    eps = 1e-6
    for (t, k) in [(0, 1), (2, 0), (4, 2)]:
        lb2 = lb.copy(); lb2[0, t, k] += eps
        _, l2 = H.forward(lp, la, lb2, L)
        assert abs((l2[0] - ll[0]) / eps - gamma[0, t, k]) < 1e-5
    for (t, i, j) in [(1, 0, 2), (3, 1, 1)]:
        la2 = la.copy(); la2[0, t, i, j] += eps
        _, l2 = H.forward(lp, la2, lb, L)
        assert abs((l2[0] - ll[0]) / eps - xi[0, t, i, j]) < 1e-5


def test_scan_matches_sequential():
    lp, la, lb = rand_hmm(3, 37, 3, seed=2)
    L = np.array([37, 20, 1])
    _, ll = H.forward(lp, la, lb, L)
    for chunk in (1, 4, 16):
        assert np.allclose(H.scan_loglik(lp, la, lb, L, chunk=chunk), ll, rtol=1e-12, atol=1e-10)


def test_weather_model_of_reference_hmm_py():
    """The reference's only HMM statement (hmm.py:11-42, a TFP doc snippet): marginal means 9 - 6*2^-t ("approach
    9.0", :38) and log_prob(zeros(7)) (:42) by enumeration over the 2^7 paths."""
    lp, lA, lb0, mu = H.weather_model(7)
    L = np.array([7])
    gamma, _, ll = H.posteriors(lp, lA, lb0, L)
    assert abs(ll[0]) < 1e-14                                    # no evidence: log 1
    means = gamma[0] @ mu
    assert np.allclose(means, 9.0 - 6.0 * 0.5 ** np.arange(7), atol=1e-13)
    assert np.allclose(means, [3.0, 6.0, 7.5, 8.25, 8.625, 8.8125, 8.90625])
    lp, lA, lb, _ = H.weather_model(7, obs=np.zeros(7))
    _, _, ll = H.posteriors(lp, lA, lb, L)
    ll_bf, gamma_bf, path_bf, best, n_best = H.brute_force(lp, lA[0], lb[0])
    assert abs(ll[0] - ll_bf) < 1e-13
    assert abs(ll[0] - (-19.855635)) < 2e-6                      # value of model.log_prob(zeros(7)), checked by enumeration
    path, score = H.viterbi(lp, lA, lb, L, dtype=np.float64)
    assert (path[0] == 0).all() and n_best == 1 and (path_bf == 0).all()   # a week at 0 degrees is a cold week
