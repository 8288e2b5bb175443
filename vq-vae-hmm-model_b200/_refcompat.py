"""Delegation of the reference's OUT-OF-SCOPE names (portfolio optimisers, Sharpe losses, hedgers,
back-testers, MetaTrainer ...; SURVEY.md section 2 rows marked out of scope) to the reference's own
files.  The drop-in modules of this repo (``VQ_VAE_HMM_fixed``, ``src``) shadow the reference modules of
the same name, so call sites such as ``from VQ_VAE_HMM_fixed import VAE_HMM, train_model,
RegimePortfolioOptimizer, sharpe_loss`` (train.py:4) also ask them for names this repo does not
(and must not) re-implement.  Those names are looked up lazily in the reference checkout that sits
BEHIND this repo on ``sys.path`` (``PYTHONPATH=<this repo>:<reference>``) or under
``$VHMM_REFERENCE_ROOT``; without one the import fails with an ImportError that says so."""
from __future__ import annotations

import importlib.util
import os
import sys

_REPO_ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
_loaded = {}


def reference_roots():
    """Directories that hold a reference checkout (identified by its VQ_VAE_HMM_fixed.py), this repo excluded."""
    cand = []
    env = os.environ.get("VHMM_REFERENCE_ROOT")
    if env:
        cand.append(env)
    cand += [p or os.getcwd() for p in sys.path]
    out = []
    for p in cand:
        try:
            ap = os.path.abspath(p)
        except Exception:
            continue
        if ap == _REPO_ROOT or ap in out:
            continue
        if os.path.isfile(os.path.join(ap, "VQ_VAE_HMM_fixed.py")) and os.path.isdir(os.path.join(ap, "src")):
            out.append(ap)
    return out


def reference_file(relpath: str):
    for r in reference_roots():
        f = os.path.join(r, relpath)
        if os.path.exists(f):
            return f
    return None


def reference_module(relpath: str):
    """The reference file ``relpath`` executed as a private module (never registered under its public name)."""
    f = reference_file(relpath)
    if f is None:
        return None
    m = _loaded.get(f)
    if m is None:
        name = "_vhmm_reference_" + relpath.replace(os.sep, "_").replace(".py", "")
        spec = importlib.util.spec_from_file_location(name, f)
        m = importlib.util.module_from_spec(spec)
        old = sys.dont_write_bytecode
        sys.dont_write_bytecode = True      # the reference checkout may be read-only
        try:
            spec.loader.exec_module(m)
        finally:
            sys.dont_write_bytecode = old
        _loaded[f] = m
    return m


def delegated(public_module: str, relpath: str, name: str):
    m = reference_module(relpath)
    if m is None or not hasattr(m, name):
        raise ImportError(
            f"cannot import name '{name}' from '{public_module}': it is outside the B200 hot path and is served by "
            f"the reference's own {relpath}; put the reference checkout on PYTHONPATH behind this repo "
            f"(or set VHMM_REFERENCE_ROOT)", name=public_module)
    return getattr(m, name)
