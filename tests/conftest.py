import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

GOLDEN = os.path.join(ROOT, "tests", "golden")
GOLDEN_CASES = ["base_fresh", "base_ckpt", "base_collate", "smoke_cfg", "scaled_small"]


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


def load_golden(name):
    import torch
    z = np.load(os.path.join(GOLDEN, name + ".npz"))
    d = {k: z[k] for k in z.files}
    params = {k[len("param:"):]: torch.from_numpy(v) for k, v in d.items() if k.startswith("param:")}
    grads = {k[len("grad:"):]: torch.from_numpy(v) for k, v in d.items() if k.startswith("grad:")}
    grads64 = {k[len("grad64:"):]: torch.from_numpy(v) for k, v in d.items() if k.startswith("grad64:")}
    C, H, K, H2, U, TH = [int(v) for v in d["ctor"]]
    return dict(raw=d, params=params, grads=grads, grads64=grads64,
                dims=dict(C=C, H=H, K=K, H2=H2, U=U, TH=TH),
                x=torch.from_numpy(d["x"]), u=torch.from_numpy(d["u"]),
                lengths=torch.from_numpy(d["lengths"]), beta=float(d["beta"]))


def rel_err(a, b):
    """norm-wise relative error |a-b| / max(|b|, tiny) (SURVEY.md 8(d) parity gates)."""
    import torch
    a = torch.as_tensor(a, dtype=torch.float64).flatten()
    b = torch.as_tensor(b, dtype=torch.float64).flatten()
    den = float(b.norm())
    return float((a - b).norm()) / (den if den > 0 else 1.0)


def maxabs_rel(a, b):
    import torch
    a = torch.as_tensor(a, dtype=torch.float64)
    b = torch.as_tensor(b, dtype=torch.float64)
    den = float(b.abs().max())
    return float((a - b).abs().max()) / (den if den > 0 else 1.0)


@pytest.fixture(params=GOLDEN_CASES)
def golden(request):
    return load_golden(request.param)
