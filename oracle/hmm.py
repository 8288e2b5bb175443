"""Oracle for the recursion stages (forward, backward/posteriors, Viterbi, scan, per-state
emissions).  TEST INFRASTRUCTURE ONLY (see oracle/__init__.py).

**parity unpinned**: the reference has no forward/backward/Viterbi/scan code at all
(hmm.py:1-42 is an un-imported TensorFlow-Probability doc snippet).
This file restates SURVEY.md Appendix B, whose conventions are derived from the pinned path:
  * log_A[b,t] is the transition INTO step t, row = previous state, column = next state
    (VQ_VAE_HMM_fixed.py:125-127); log_A[b,0] is unused (:127 slices [:,1:]).
  * step t is valid iff t < lengths[b] (:111); a transition needs both ends valid (:128).
  * the Gaussian log-density uses var = exp(logvar).clamp(min=1e-8) (:118-119), summed over
    channels (not divided by C).
numpy, vectorised over B with a Python loop over T; float64 unless stated.  Its own pins are
the brute-force enumerations at the bottom (tests/test_oracle_hmm.py).
"""
from __future__ import annotations

import itertools
import math

import numpy as np

PAD_STATE = 255  # sentinel written into Viterbi paths for t >= lengths[b] (Appendix B.4)


def _lse(a, axis):
    m = np.max(a, axis=axis, keepdims=True)
    m = np.where(np.isfinite(m), m, 0.0)
    with np.errstate(divide="ignore"):
        return np.squeeze(m, axis) + np.log(np.sum(np.exp(a - m), axis=axis))


def forward(log_pi, log_A, log_b, lengths):
    """Appendix B.2.  Returns (alpha (B,T,K) float64, loglik (B,)).
    alpha[b,t] for t >= L_b repeats alpha[b,L_b-1] (carried unchanged)."""
    log_pi = np.asarray(log_pi, np.float64)
    log_A = np.asarray(log_A, np.float64)
    log_b = np.asarray(log_b, np.float64)
    B, T, K = log_b.shape
    alpha = np.empty((B, T, K))
    alpha[:, 0] = log_pi[None] + log_b[:, 0]
    for t in range(1, T):
        nxt = log_b[:, t] + _lse(alpha[:, t - 1][:, :, None] + log_A[:, t], axis=1)
        live = (t < np.asarray(lengths))[:, None]
        alpha[:, t] = np.where(live, nxt, alpha[:, t - 1])
    return alpha, _lse(alpha[:, T - 1], axis=1)


def backward(log_A, log_b, lengths):
    """Appendix B.3.  beta (B,T,K); beta[b,t]=0 for t >= L_b-1."""
    log_A = np.asarray(log_A, np.float64)
    log_b = np.asarray(log_b, np.float64)
    B, T, K = log_b.shape
    L = np.asarray(lengths)
    beta = np.zeros((B, T, K))
    for t in range(T - 2, -1, -1):
        nxt = _lse(log_A[:, t + 1] + (log_b[:, t + 1] + beta[:, t + 1])[:, None, :], axis=2)
        live = (t + 1 < L)[:, None]
        beta[:, t] = np.where(live, nxt, 0.0)
    return beta


def posteriors(log_pi, log_A, log_b, lengths):
    """gamma (B,T,K), xi (B,T,K,K) (xi[:,0]=0), loglik (B,).  Appendix B.3.
    gamma/xi are also d loglik / d log_b and d loglik / d log_A."""
    alpha, ll = forward(log_pi, log_A, log_b, lengths)
    beta = backward(log_A, log_b, lengths)
    B, T, K = alpha.shape
    L = np.asarray(lengths)
    valid = (np.arange(T)[None, :] < L[:, None])
    gamma = np.exp(alpha + beta - ll[:, None, None]) * valid[:, :, None]
    xi = np.zeros((B, T, K, K))
    la = np.asarray(log_A, np.float64)
    lb = np.asarray(log_b, np.float64)
    for t in range(1, T):
        e = alpha[:, t - 1][:, :, None] + la[:, t] + (lb[:, t] + beta[:, t])[:, None, :] - ll[:, None, None]
        xi[:, t] = np.exp(e) * valid[:, t][:, None, None]
    return gamma, xi, ll


def viterbi(log_pi, log_A, log_b, lengths, dtype=np.float32):
    """Appendix B.4, evaluated in ``dtype`` with the exact operation order the device kernel
    uses, so that paths are comparable bit for bit:
        delta[0,j] = log_pi[j] + log_b[0,j]
        cand(i,j)  = delta[t-1,i] + log_A[t,i,j]            (one rounding)
        best(j)    = max over i ascending, strict '>' update  -> lowest i on ties
        delta[t,j] = log_b[t,j] + best(j)                   (one rounding)
    final state = lowest j attaining max_j delta[L-1,j].
    Returns (path uint8 (B,T) with PAD_STATE for t >= L_b, score (B,) dtype)."""
    lp = np.asarray(log_pi, dtype)
    lA = np.asarray(log_A, dtype)
    lb = np.asarray(log_b, dtype)
    B, T, K = lb.shape
    L = np.asarray(lengths).astype(np.int64)
    delta = (lp[None] + lb[:, 0]).astype(dtype)
    psi = np.zeros((B, T, K), np.uint8)
    for t in range(1, T):
        cand = (delta[:, :, None] + lA[:, t]).astype(dtype)          # (B,i,j)
        best = cand[:, 0, :].copy()
        arg = np.zeros((B, K), np.uint8)
        for i in range(1, K):
            better = cand[:, i, :] > best
            best = np.where(better, cand[:, i, :], best)
            arg = np.where(better, np.uint8(i), arg)
        nd = (lb[:, t] + best).astype(dtype)
        live = (t < L)[:, None]
        delta = np.where(live, nd, delta)
        psi[:, t] = arg
    score = delta.max(axis=1)
    state = np.argmax(delta, axis=1).astype(np.uint8)               # first index on ties
    path = np.full((B, T), PAD_STATE, np.uint8)
    rows = np.arange(B)
    for t in range(T - 1, -1, -1):
        inside = (t < L)
        path[inside, t] = state[inside]
        # step back only for sequences that are inside and t >= 1
        if t >= 1:
            prev = psi[rows, t, state]
            state = np.where(inside, prev, state).astype(np.uint8)
    return path, score


def scan_loglik(log_pi, log_A, log_b, lengths, chunk=4):
    """Appendix B.5: alpha_t = v0 (x) M_1 (x) ... (x) M_t with the log-semiring product,
    associated chunk-wise (chunk-local products, then a carry pass) -- a different
    association order from ``forward``; float64.  Returns loglik (B,)."""
    log_A = np.asarray(log_A, np.float64)
    log_b = np.asarray(log_b, np.float64)
    log_pi = np.asarray(log_pi, np.float64)
    B, T, K = log_b.shape
    L = np.asarray(lengths)
    ident = np.where(np.eye(K, dtype=bool), 0.0, -np.inf)

    def elem(t):
        m = log_A[:, t] + log_b[:, t][:, None, :]
        live = (t < L)[:, None, None]
        return np.where(live, m, ident[None])

    def mul(P, Q):
        return _lse(P[:, :, :, None] + Q[:, None, :, :], axis=2)

    chunks = []
    for s in range(1, T, chunk):
        P = elem(s)
        for t in range(s + 1, min(T, s + chunk)):
            P = mul(P, elem(t))
        chunks.append(P)
    v = log_pi[None] + log_b[:, 0]
    if chunks:
        tot = chunks[-1]
        for P in reversed(chunks[:-1]):   # right-to-left association on purpose
            tot = mul(P, tot)
        v = _lse(v[:, :, None] + tot, axis=1)
    return _lse(v, axis=1)


def emission_table(params, T, decode_fn):
    """Appendix B.1: decoder evaluated on the K constant one-hot sequences of length T.
    ``decode_fn(params, q)`` is oracle.elbo.decode.  Returns (mu (K,C,T), logvar (K,C,T))."""
    import torch
    K = params["decoder.embeddings.weight"].shape[0]
    dt = params["decoder.embeddings.weight"].dtype
    q = torch.zeros(K, K, T, dtype=dt)
    for k in range(K):
        q[k, k, :] = 1.0
    mu, logvar = decode_fn(params, q)
    return mu.detach(), logvar.detach()


def emission_logprob(x, mu_tab, logvar_tab):
    """log_b[b,t,k] = -0.5 sum_c [ log(2 pi v) + (x-mu)^2 / v ],  v = exp(logvar).clamp(1e-8).
    x (B,C,T); tables (K,C,T).  Appendix B.1 / VQ_VAE_HMM_fixed.py:118-119.  float64."""
    x = np.asarray(x, np.float64)
    mu = np.asarray(mu_tab, np.float64)
    lv = np.asarray(logvar_tab, np.float64)
    v = np.maximum(np.exp(lv), 1e-8)
    d = x[:, None, :, :] - mu[None]                                  # (B,K,C,T)
    lb = -0.5 * (np.log(2 * math.pi * v)[None] + d * d / v[None]).sum(axis=2)  # (B,K,T)
    return np.transpose(lb, (0, 2, 1))                               # (B,T,K)


# ---------------------------------------------------------------------------------------
# brute force (the pins of this file): enumerate all K^T state paths of ONE sequence
# ---------------------------------------------------------------------------------------
def brute_force(log_pi, log_A, log_b):
    """log_A (T,K,K), log_b (T,K) for one full-length sequence.
    Returns (loglik, gamma (T,K), a best path (T,), best score, number of maximisers) by
    enumeration.  The Viterbi tie-break does not pick the lexicographically first maximiser
    in general, so callers compare scores always and paths only when n_best == 1."""
    log_pi = np.asarray(log_pi, np.float64)
    log_A = np.asarray(log_A, np.float64)
    log_b = np.asarray(log_b, np.float64)
    T, K = log_b.shape
    scores = []
    paths = list(itertools.product(range(K), repeat=T))
    for z in paths:
        s = log_pi[z[0]] + log_b[0, z[0]]
        for t in range(1, T):
            s += log_A[t, z[t - 1], z[t]] + log_b[t, z[t]]
        scores.append(s)
    scores = np.array(scores)
    m = scores.max()
    ll = m + math.log(np.exp(scores - m).sum())
    w = np.exp(scores - ll)
    gamma = np.zeros((T, K))
    for z, wz in zip(paths, w):
        for t in range(T):
            gamma[t, z[t]] += wz
    n_best = int((scores == m).sum())
    return ll, gamma, np.array(paths[int(np.argmax(scores))]), m, n_best


# ---------------------------------------------------------------------------------------
# the reference tree's one HMM statement: the TFP "weather" model quoted in hmm.py:11-42
# ---------------------------------------------------------------------------------------
def weather_model(T=7, obs=None):
    """hmm.py:11-42 in this file's conventions: K=2 (0 cold, 1 hot), pi=[.8,.2] (:11), A=[[.7,.3],[.2,.8]]
    (:17-18, stationary -> broadcast over t), emissions N(0,5) / N(15,10) (:25), num_steps=7 (:34).
    ``obs`` (T,) gives log_b[t,k] = log N(obs_t; mu_k, sigma_k); obs=None gives log_b = 0 (no evidence), for which
    the posterior gamma is the prior state marginal pi A^t and the log-likelihood is log 1 = 0.
    Returns (log_pi (2,), log_A (1,T,2,2), log_b (1,T,2), state means (2,)).  Closed forms the snippet states:
    E[x_t] = 15 P(hot at t) = 9 - 6 * 2^-t  ("elements approach 9.0", :38); stationary P(hot) = .3/(.3+.2) = .6."""
    pi = np.array([0.8, 0.2])
    A = np.array([[0.7, 0.3], [0.2, 0.8]])
    mu = np.array([0.0, 15.0])
    sd = np.array([5.0, 10.0])
    log_A = np.broadcast_to(np.log(A), (1, T, 2, 2)).copy()
    if obs is None:
        log_b = np.zeros((1, T, 2))
    else:
        o = np.asarray(obs, np.float64)[:, None]
        log_b = (-0.5 * ((o - mu) / sd) ** 2 - np.log(sd) - 0.5 * math.log(2 * math.pi))[None]
    return np.log(pi), log_A, log_b, mu
