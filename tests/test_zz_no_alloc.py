"""vhmm_viterbi_static (include/vhmm.h): the allocation-free variant of vhmm_viterbi -- back-pointer scratch in a static device
array inside the library, no caller workspace.  Called through the C-ABI and compared with the existing entry point: paths and
scores BIT-EXACT (the bar of every Viterbi test in this repo), on two calls in a row (the array keeps the previous call's
contents: a second, different problem must not see them), plus the size limit's error.  The file name sorts last on
purpose."""
import ctypes

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def _rand_hmm(B, T, K, seed):
    r = np.random.default_rng(seed)
    lp = r.normal(size=K); lp -= np.log(np.exp(lp).sum())
    la = r.normal(size=(B, T, K, K)); la -= np.log(np.exp(la).sum(-1, keepdims=True))
    lb = r.normal(size=(B, T, K)) * 3.0
    # quantised: exact ties occur, the lowest-index tie-break is exercised
    lp, la, lb = np.round(lp * 2) / 2, np.round(la * 2) / 2, np.round(lb * 2) / 2
    L = r.integers(1, T + 1, size=B); L[0] = T
    return [torch.from_numpy(np.ascontiguousarray(v)).cuda()
            for v in (lp.astype(np.float32), la.astype(np.float32), lb.astype(np.float32), L.astype(np.int64))]


def _static(lib, Fn, _lib, lp, la, lb, L, B, T, K):
    # outputs are the caller's; poisoned so that an untouched element shows
    path = torch.full((B, T), 77, dtype=torch.uint8, device="cuda")
    score = torch.full((B,), float("nan"), device="cuda")
    rc = lib.vhmm_viterbi_static(Fn._ptr(lp), Fn._ptr(la), Fn._ptr(lb), Fn._ptr(L), B, T, K, Fn._ptr(path), Fn._ptr(score),
                                 Fn._stream(lb.device))
    _lib.check(rc, "vhmm_viterbi_static")
    return path, score


@pytest.mark.parametrize("K", [3, 32])
def test_viterbi_static_matches_viterbi_twice_in_a_row(K):
    import vhmm_b200 as V
    from vhmm_b200 import functional as Fn, _lib
    lib = V.lib()
    # set-up only call (B == 0): resolves the scratch array's device address, computes nothing
    _lib.check(lib.vhmm_viterbi_static(None, None, None, None, 0, 0, K, None, None, Fn._stream(torch.device("cuda", 0))),
               "vhmm_viterbi_static(set-up)")
    # two different problems back to back; the second is smaller, so stale back-pointers of the first stay in the array
    shapes = [(37, 96), (9, 41)] if K == 3 else [(6, 64), (3, 20)]
    for call, (B, T) in enumerate(shapes):
        lp, la, lb, L = _rand_hmm(B, T, K, seed=100 * K + call)
        ref_path, ref_score = Fn.viterbi(lp, la, lb, L)          # existing entry point (caller workspace)
        path, score = _static(lib, Fn, _lib, lp, la, lb, L, B, T, K)
        torch.cuda.synchronize()
        assert torch.equal(path, ref_path), (K, call)
        assert torch.equal(score, ref_score), (K, call)          # bit-exact, like the paths
    # and the first problem again: same answer as before
    B, T = shapes[0]
    lp, la, lb, L = _rand_hmm(B, T, K, seed=100 * K)
    ref_path, ref_score = Fn.viterbi(lp, la, lb, L)
    path, score = _static(lib, Fn, _lib, lp, la, lb, L, B, T, K)
    assert torch.equal(path, ref_path) and torch.equal(score, ref_score)


def test_viterbi_static_rejects_sizes_beyond_the_scratch():
    import vhmm_b200 as V
    from vhmm_b200 import functional as Fn, _lib
    lib = V.lib()
    K, T = 3, 4096
    B = (4 << 20) // T + 1                      # one sequence more than VHMM_VITERBI_STATIC_MAX_BYTES holds at K <= 4
    # the size check comes before any memory is touched: tiny stand-in tensors are enough
    z = torch.zeros(16, device="cuda")
    L = torch.ones(1, dtype=torch.int64, device="cuda")
    out = torch.zeros(16, dtype=torch.uint8, device="cuda")
    rc = lib.vhmm_viterbi_static(Fn._ptr(z), Fn._ptr(z), Fn._ptr(z), Fn._ptr(L), B, T, K, Fn._ptr(out), None,
                                 Fn._stream(z.device))
    assert rc == -3                              # VHMM_E_WORKSPACE
    msg = lib.vhmm_last_error()
    msg = msg.decode() if isinstance(msg, (bytes, bytearray)) else str(msg)
    assert "VHMM_VITERBI_STATIC_MAX_BYTES" in msg, msg
    with pytest.raises(_lib.VhmmError):
        _lib.check(rc, "vhmm_viterbi_static")
    # exactly at the limit the size is accepted (checked without running it: B == 0 path is separate, so only the arithmetic)
    assert (B - 1) * T == 4 << 20
