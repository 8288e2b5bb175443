#!/usr/bin/env python3
"""Recipe that places the UNMODIFIED reference module of the path into the git-ignored ``oracle/_ref/``.

The reference is a flat set of Python scripts; the whole pinned path (Encoder, Prior, Decoder,
VAE_HMM.compute_loss, collate_fn, train_model) lives in ONE file, /root/reference/VQ_VAE_HMM_fixed.py,
whose only dependency is torch.  ``oracle/_ref/`` is listed in .gitignore (reference sources never enter
this repo's history) but not in .gpurunignore, so the copy travels to the GPU box next to the built
``libvhmm.so`` and ``bench.py --impl reference`` can time the reference ITSELF on the box's host cores
(``cpu_baseline.kind = "reference"``).  Without /root/reference (the GPU box) this script does nothing and
the arm falls back to whatever ``oracle/_ref/`` already holds, else to the pinned port (oracle/elbo.py).

TEST / BENCH INFRASTRUCTURE ONLY: the copy is loaded under a private module name (never as
``VQ_VAE_HMM_fixed``, which is this repo's drop-in) by bench.py's reference arm and by tests/.
"""
import hashlib
import os
import shutil
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
REF_ROOT = os.environ.get("VHMM_REFERENCE_ROOT", "/root/reference")
FILES = ("VQ_VAE_HMM_fixed.py",)
OUT = os.path.join(HERE, "_ref")


def build(verbose: bool = True) -> bool:
    """Copy the reference files; returns True when oracle/_ref/ is populated afterwards."""
    if os.path.isdir(REF_ROOT):
        os.makedirs(OUT, exist_ok=True)
        for f in FILES:
            src, dst = os.path.join(REF_ROOT, f), os.path.join(OUT, f)
            shutil.copyfile(src, dst)
            if verbose:
                h = hashlib.sha256(open(dst, "rb").read()).hexdigest()[:16]
                print(f"oracle/_ref/{f}  <- {src}  sha256:{h}")
    return all(os.path.exists(os.path.join(OUT, f)) for f in FILES)


def load():
    """The reference module from oracle/_ref/ under a private name, or None when it is absent."""
    import importlib.util
    f = os.path.join(OUT, FILES[0])
    if not os.path.exists(f):
        return None
    name = "_vhmm_oracle_ref_module"
    if name in sys.modules:
        return sys.modules[name]
    spec = importlib.util.spec_from_file_location(name, f)
    m = importlib.util.module_from_spec(spec)
    old = sys.dont_write_bytecode
    sys.dont_write_bytecode = True
    try:
        spec.loader.exec_module(m)
    finally:
        sys.dont_write_bytecode = old
    sys.modules[name] = m
    return m


if __name__ == "__main__":
    ok = build()
    print("oracle/_ref populated" if ok else "oracle/_ref NOT populated (no reference checkout here)")
