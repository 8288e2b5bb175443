"""The sharded ELBO equals the un-sharded ELBO of the concatenated ragged batch: the reference
divides the reconstruction term by the GLOBAL valid count (VQ_VAE_HMM_fixed.py:120) and the prior / entropy terms
by the GLOBAL batch size (:131, :135), so the result is NOT the mean of per-shard compute_loss values.
world_size 2 over gloo on 127.0.0.1, uneven shards (5 + 7 sequences, one empty, one of length 1)."""
import os
import socket
import subprocess
import sys

import numpy as np
import pytest
import torch

from conftest import rel_err

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return str(p)


def _run_world(mode, tmp_path):
    port = _free_port()
    outs = [str(tmp_path / f"rank{r}.npz") for r in range(2)]
    procs = [subprocess.Popen([sys.executable, os.path.join(HERE, "_sharded_worker.py"), mode, str(r), "2", port, outs[r]],
                              stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True) for r in range(2)]
    logs = []
    for p in procs:
        try:
            o, _ = p.communicate(timeout=600)
        except subprocess.TimeoutExpired:
            for q in procs:
                q.kill()
            raise
        logs.append(o)
    for p, o in zip(procs, logs):
        assert p.returncode == 0, o[-3000:]
    return [dict(np.load(o)) for o in outs]


def _check(results, tol):
    from oracle import elbo as O
    from _sharded_worker import make_batch, shard
    x, u, L = make_batch()
    p64 = {k: v.double() for k, v in O.init_params(5, 64, 3, 32, 4, 128, seed=3).items()}
    for tag, beta in (("a", 1.0), ("b", 0.37)):
        ref, _, g = O.loss_and_grads(p64, x.double(), u.double(), L, 4, beta)     # the concatenated batch, one process
        for r in results:                                                       # identical on every rank
            assert abs(float(r[f"loss_{tag}"]) - float(ref)) <= tol * abs(float(ref)), (tag, float(r[f"loss_{tag}"]), float(ref))
            for k in O.PARAM_ORDER:
                assert rel_err(r[f"grad_{tag}:{k}"], g[k]) <= tol, (tag, k)
        # and it is NOT the mean of the per-shard losses (different valid counts): the test would be vacuous otherwise
        per = [float(O.compute_loss(p64, x[lo:hi].double(), u[lo:hi].double(), L[lo:hi], 4, beta)) for lo, hi in (shard(0), shard(1))]
        assert abs(sum(per) / 2 - float(ref)) > 100 * tol * abs(float(ref))
    ref, _, g = O.loss_and_grads(p64, x.double(), u.double(), L, 4, 0.37)
    flat = torch.cat([g[k].flatten() for k in O.PARAM_ORDER])
    for r in results:
        # tag "c": global normalisers from the host (compute_loss(..., global_lengths=L)): one collective, same result as "b"
        assert abs(float(r["loss_c"]) - float(ref)) <= tol * abs(float(ref))
        for k in O.PARAM_ORDER:
            assert rel_err(r[f"grad_c:{k}"], g[k]) <= tol, ("c", k)
            assert np.array_equal(r[f"grad_c:{k}"], r[f"grad_b:{k}"]), ("c == b", k)
        assert abs(float(r["loss_flat"]) - float(ref)) <= tol * abs(float(ref))
        assert rel_err(r["flat"], flat) <= tol
    assert np.array_equal(results[0]["flat"], results[1]["flat"])               # all-reduced: bitwise equal on both ranks


def test_sharded_elbo_equals_unsharded_world2_gloo_cpu(tmp_path):
    """Host-side logic at world_size 2 on the CPU (gloo): the kernel call is replaced by the float64 oracle inside the
    worker processes of THIS TEST (the product itself has no CPU path)."""
    _check(_run_world("cpu", tmp_path), 2e-6)    # fp32 all-reduce payload, hi/lo split of the sums


@pytest.mark.gpu
def test_sharded_elbo_equals_unsharded_world2_real_kernels(tmp_path):
    """Two ranks on one GPU with the real kernels (gloo moves the CUDA tensors)."""
    _check(_run_world("gpu", tmp_path), 1e-5)
