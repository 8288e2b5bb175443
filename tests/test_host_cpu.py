"""CPU-side checks of the host package: the C-ABI library builds/loads and exports every symbol
declared in include/vhmm.h (no compute calls without a GPU), the module mirrors the reference
interface, collate semantics."""
import os
import re

import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_library_exports_every_declared_symbol():
    import vhmm_b200 as V
    L = V.lib()
    hdr = open(os.path.join(ROOT, "include", "vhmm.h")).read()
    declared = set(re.findall(r"\b(vhmm_[a-z0-9_]+)\s*\(", hdr))
    assert len(declared) >= 17
    from vhmm_b200 import _lib
    assert declared == set(_lib.PROTOTYPES), declared ^ set(_lib.PROTOTYPES)
    for name in declared:
        assert hasattr(L, name), name
    assert L.vhmm_version() == 113


def test_param_layout_matches_state_dict():
    import vhmm_b200 as V
    from vhmm_b200 import functional as Fn
    m = V.VAE_HMM(5, 64, 3, 32, u_dim=4)
    assert list(m.state_dict().keys()) == list(Fn.PARAM_ORDER)
    off = Fn.param_offsets(m._dims)
    sizes = [m.state_dict()[k].numel() for k in Fn.PARAM_ORDER]
    assert [off[i + 1] - off[i] for i in range(18)] == sizes
    assert Fn.param_count(m._dims) == 34649


def test_same_init_as_reference_golden():
    """Default-initialised weights equal the reference's for the same seed (same RNG order)."""
    import vhmm_b200 as V
    from conftest import load_golden
    g = load_golden("base_fresh")
    torch.manual_seed(0)
    m = V.VAE_HMM(5, 64, 3, 32, u_dim=4)
    for k, v in m.state_dict().items():
        assert torch.equal(v, g["params"][k]), k


def test_no_cpu_fallback_and_reference_errors():
    import vhmm_b200 as V
    m = V.VAE_HMM(5, 8, 3, 4, u_dim=2, trans_hidden=8)
    with pytest.raises(ValueError):
        m.compute_loss(torch.zeros(1, 5, 4), torch.zeros(1, 2, 4), None)
    with pytest.raises(ValueError):
        V.VAE_HMM(5, 8, 3, 4)
    if not torch.cuda.is_available():
        with pytest.raises(V.VhmmError):
            m.encode(torch.zeros(1, 5, 4))


def test_collate_contract():
    import random
    import vhmm_b200 as V
    random.seed(1)
    xs = torch.randn(4, 5, 90)
    us = torch.randn(4, 4, 90)
    ds = V.RandomChunkDataset(xs, us, min_len=20, max_len=50)
    assert len(ds) == 1000
    batch = [ds[i] for i in range(6)]
    x, u, L = V.collate_fn(batch, "cpu")
    assert L.dtype == torch.int64 and L.device.type == "cpu"
    T = int(L.max())
    assert x.shape == (6, 5, T) and u.shape == (6, 4, T)
    for i, (xi, ui, n) in enumerate(batch):
        assert torch.equal(x[i, :, :n], xi) and float(x[i, :, n:].abs().sum()) == 0
        assert torch.equal(u[i, :, :n], ui) and float(u[i, :, n:].abs().sum()) == 0
    import src.utils
    import VQ_VAE_HMM_fixed as R
    assert R.VAE_HMM is V.VAE_HMM and R.collate_fn is V.collate_fn


def test_sample_triples_consumes_random_like_the_reference_dataset():
    """training.sample_triples + plain slicing reproduces, on the CPU, the batches the real reference
    RandomChunkDataset + collate_fn produced (tests/golden/collate_stream.npz): same three
    random.randint calls per item in the same order."""
    import random
    import numpy as np
    from conftest import GOLDEN
    from vhmm_b200.training import sample_triples
    z = np.load(os.path.join(GOLDEN, "collate_stream.npz"))
    g = torch.Generator().manual_seed(77)
    lens = [int(n) for n in z["lens"]]
    xs = [torch.randn(5, n, generator=g) for n in lens]
    us = [torch.randn(4, n, generator=g) for n in lens]
    random.seed(5)
    for i in range(3):
        idx, ln, st = sample_triples(lens, 16, 20, 50)
        assert np.array_equal(ln, z[f"L{i}"])
        T = int(ln.max())
        x = np.zeros((16, 5, T), np.float32); u = np.zeros((16, 4, T), np.float32)
        for b in range(16):
            x[b, :, :ln[b]] = xs[idx[b]][:, st[b]:st[b] + ln[b]].numpy()
            u[b, :, :ln[b]] = us[idx[b]][:, st[b]:st[b] + ln[b]].numpy()
        assert np.array_equal(x, z[f"x{i}"]) and np.array_equal(u, z[f"u{i}"])


# every ``from VQ_VAE_HMM_fixed import`` / ``from src... import`` line of the reference tree (grep of the checkout;
# the files are cited so the list can be re-derived)
REFERENCE_IMPORT_LINES = [
    "from VQ_VAE_HMM_fixed import VAE_HMM",                                                   # api.py:4, app.py:5, inference_api/app.py:10, tests/smoke_test.py:11
    "from VQ_VAE_HMM_fixed import VAE_HMM, RandomChunkDataset, collate_fn",                   # evaluation/evaluate.py:19, training_pipeline/train.py:16
    "from VQ_VAE_HMM_fixed import RandomChunkDataset, collate_fn",                            # data_loader.py:6
    "from VQ_VAE_HMM_fixed import VAE_HMM, train_model, RegimePortfolioOptimizer, sharpe_loss",   # train.py:4
    "from src.models import VAE_HMM",                                                         # examples/train_example.py:5
    "from src.portfolio import RegimePortfolioOptimizer, sharpe_loss",                        # examples/train_example.py:6
    "from src.training import Trainer",                                                       # examples/train_example.py:7
    "from src.utils import RandomChunkDataset, collate_fn",                                   # examples/train_example.py:8
    "from src.models.vae_hmm import VAE_HMM, Encoder, Decoder, Prior",                        # src/__init__.py:11
    "from src.portfolio.optimizers import RegimePortfolioOptimizer, AttentionPortfolioOptimizer, TransformerPortfolioOptimizer, BayesianPortfolioOptimizer",  # src/__init__.py:12
    "from src.portfolio.hedging import RegimeDeltaHedger, DynamicDeltaHedger",                # src/__init__.py:18
    "from src.training.trainer import Trainer, MetaTrainer, WalkForwardTrainer",              # src/__init__.py:19
    "from src.utils.data import RandomChunkDataset, collate_fn, save_checkpoint, load_checkpoint",  # src/utils/__init__.py:3
]
OUT_OF_SCOPE = ("RegimePortfolioOptimizer", "sharpe_loss", "src.portfolio", "MetaTrainer", "WalkForwardTrainer")


def _run_import_lines(lines, pythonpath):
    import subprocess
    import sys
    code = "import vhmm_b200\n" + "\n".join(lines) + "\nprint('IMPORTS_OK')\n"
    env = dict(os.environ, PYTHONPATH=pythonpath, PYTHONDONTWRITEBYTECODE="1")
    env.pop("VHMM_REFERENCE_ROOT", None)
    return subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, env=env, cwd="/tmp")


def test_every_reference_import_site_resolves():
    """With this repo AHEAD of a reference checkout on PYTHONPATH every import line of the reference tree works:
    hot-path names come from this repo, out-of-scope names from the reference's own files."""
    ref = "/root/reference"
    if not os.path.isdir(ref):
        pytest.skip("no reference checkout on this machine (the GPU box)")
    r = _run_import_lines(REFERENCE_IMPORT_LINES + [
        "import vhmm_b200, VQ_VAE_HMM_fixed, src.models, src.training",
        "assert VQ_VAE_HMM_fixed.VAE_HMM is vhmm_b200.VAE_HMM and src.models.VAE_HMM is vhmm_b200.VAE_HMM",
        "assert src.training.Trainer is vhmm_b200.Trainer",
        "assert RegimePortfolioOptimizer.__module__.startswith('src.portfolio')",
    ], ROOT + os.pathsep + ref)
    assert r.returncode == 0 and "IMPORTS_OK" in r.stdout, r.stderr[-2000:]


def test_import_sites_without_reference_checkout():
    """Without a reference checkout the hot-path names still import; out-of-scope names raise ImportError."""
    inscope = [l for l in REFERENCE_IMPORT_LINES if not any(k in l for k in OUT_OF_SCOPE)]
    r = _run_import_lines(inscope, ROOT)
    assert r.returncode == 0 and "IMPORTS_OK" in r.stdout, r.stderr[-2000:]
    rest = [l for l in REFERENCE_IMPORT_LINES if l not in inscope]
    body = ["bad = []"]
    for l in rest:
        body += ["try:", "    " + l, "    bad.append(%r)" % l, "except ImportError:", "    pass"]
    body.append("assert not bad, bad")
    r = _run_import_lines(body, ROOT)
    assert r.returncode == 0 and "IMPORTS_OK" in r.stdout, r.stderr[-2000:]


def test_round2_host_helpers():
    """host-side helpers added in round 2: global normalisers from the whole batch's lengths (one collective per sharded
    step), the chunk-growth rule of the streamed ELBO, NUMA binding (never raises)."""
    import torch
    from vhmm_b200 import parallel, functional as Fn
    L = torch.tensor([5, 0, 12, 7, 300, -3])
    n = parallel.global_norms_from_lengths(L, C=5, T=10, device="cpu")
    assert n.dtype == torch.float64 and n.tolist() == [float((5 + 0 + 10 + 7 + 10 + 0) * 5), 6.0]
    assert parallel.global_norms_from_lengths(torch.zeros(3, dtype=torch.int64), 5, 10, "cpu").tolist() == [1.0, 3.0]   # clamp(min=1)
    assert Fn._stream_growth() == 2.0                       # single process: PCIe outruns the kernels 2x
    os.environ["VHMM_STREAM_GROWTH"] = "1.25"
    try:
        assert Fn._stream_growth() == 1.25
    finally:
        del os.environ["VHMM_STREAM_GROWTH"]
    assert parallel.bind_to_gpu_numa_node(0) in (None, 0, 1, 2, 3, 4, 5, 6, 7)   # no GPU here: None, and no exception
