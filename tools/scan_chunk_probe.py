import sys, os
sys.path.insert(0, '/root/repo'); sys.path.insert(0, '/root/repo/tools')
import torch
from vhmm_b200 import functional as Fn
from bench_hmm import make_inputs, timed
lp, la, lb, L = make_inputs(64, 65536, 3, seed=1)
print(os.environ.get("VHMM_LIB"), "scan ll ms", timed(lambda: Fn.hmm_scan(lp, la, lb, L), iters=20), "alpha ms", timed(lambda: Fn.hmm_scan(lp, la, lb, L, want_alpha=True), iters=20))
