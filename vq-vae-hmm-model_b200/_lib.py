"""ctypes binding of libvhmm.so (include/vhmm.h) + the in-tree nvcc build.

The library is plain CUDA-runtime C ABI (no torch types in any signature).  It is built
in-tree (``libvhmm.so`` next to this file) so that it travels to the GPU box with the repo
snapshot and shows up in the loaded-library list of every process that uses the product.
There is no CPU fallback: if the library cannot be loaded, every op raises.
"""
from __future__ import annotations

import ctypes
import os
import shutil
import subprocess
import threading

_HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(_HERE, "csrc")
LIB_PATH = os.environ.get("VHMM_LIB") or os.path.join(_HERE, "libvhmm.so")   # VHMM_LIB: an instrumented build (tools/)
SOURCES = ("api.cu", "elbo_layerwise.cu", "hmm.cu", "elbo_fused.cu", "batch_train.cu", "hmm_inputs.cu")
NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
              "-Xcompiler", "-fPIC", "-shared"]

# op codes (include/vhmm.h)
OP_ENCODER_FWD, OP_PRIOR_FWD, OP_DECODER_FWD, OP_ELBO = 1, 2, 3, 4
OP_HMM_FORWARD, OP_HMM_POSTERIOR, OP_VITERBI, OP_HMM_DECODE, OP_HMM_SCAN, OP_EMISSION_TABLE = 5, 6, 7, 8, 9, 10
OP_HMM_INPUTS = 11
OP_HMM_PARAM_GRADS, OP_TRANSITION_PROBS, OP_FORWARD, OP_TRAIN_STEP = 12, 13, 14, 15
# vhmm_train_step state slots (include/vhmm.h)
TS_NORM0, TS_NORM1, TS_BETA, TS_STEP, TS_LR, TS_BETA1, TS_BETA2, TS_LOSS, TS_LOSS_SUM, TS_BATCHES, TS_STEP_SIZE, TS_BC2_SQRT, TS_SUMS, TS_COUNT = \
    0, 1, 2, 3, 4, 5, 6, 7, 8, 9, 10, 11, 12, 16
NUM_PARAMS = 18
PAD_STATE = 255
ELBO_AUTO, ELBO_LAYERWISE, ELBO_FUSED = 0, 1, 2
DECODER_AUTO, DECODER_MMA, DECODER_TC5 = 0, 1, 2


class Dims(ctypes.Structure):
    _fields_ = [("C", ctypes.c_int32), ("H", ctypes.c_int32), ("K", ctypes.c_int32),
                ("H2", ctypes.c_int32), ("U", ctypes.c_int32), ("TH", ctypes.c_int32)]


class Params(ctypes.Structure):
    _fields_ = [("p", ctypes.c_void_p * NUM_PARAMS)]


class VhmmError(RuntimeError):
    pass


def _sources():
    return [os.path.join(CSRC, s) for s in SOURCES if os.path.exists(os.path.join(CSRC, s))]


def needs_build() -> bool:
    if os.environ.get("VHMM_LIB"):
        return False
    if not os.path.exists(LIB_PATH):
        return True
    t = os.path.getmtime(LIB_PATH)
    deps = [os.path.join(CSRC, f) for f in os.listdir(CSRC)] + [os.path.join(_HERE, "..", "include", "vhmm.h")]
    return any(os.path.getmtime(d) > t for d in deps if os.path.exists(d))


def build(force: bool = False, verbose: bool = False) -> str:
    """Compile every CUDA source for sm_100a into libvhmm.so (nvcc cross-compiles without a GPU)."""
    if not force and not needs_build():
        return LIB_PATH
    nvcc = shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
    if not os.path.exists(nvcc):
        raise VhmmError("nvcc not found: cannot build libvhmm.so")
    tmp = LIB_PATH + ".tmp.%d" % os.getpid()
    cmd = [nvcc] + NVCC_FLAGS + (["-Xptxas", "-v"] if verbose else []) + ["-o", tmp] + _sources()
    res = subprocess.run(cmd, capture_output=True, text=True)
    if res.returncode != 0:
        raise VhmmError("nvcc failed:\n" + res.stdout + res.stderr)
    os.replace(tmp, LIB_PATH)
    if verbose:
        print(res.stderr)
    return LIB_PATH


_lock = threading.Lock()
_lib = None

_P, _I64, _SZ, _DBL = ctypes.c_void_p, ctypes.c_int64, ctypes.c_size_t, ctypes.c_double
_DP, _PP = ctypes.POINTER(Dims), ctypes.POINTER(Params)

# name -> (restype, argtypes); must list every symbol declared in include/vhmm.h
PROTOTYPES = {
    "vhmm_version": (ctypes.c_int, []),
    "vhmm_last_error": (ctypes.c_char_p, []),
    "vhmm_launch_count": (_I64, []),
    "vhmm_param_count": (_I64, [_DP]),
    "vhmm_param_offsets": (ctypes.c_int, [_DP, ctypes.POINTER(_I64)]),
    "vhmm_workspace_bytes": (_SZ, [ctypes.c_int, _DP, _I64, _I64]),
    "vhmm_encoder_fwd": (ctypes.c_int, [_DP, _PP, _P, _I64, _I64, _P, _P, _P, _SZ, _P]),
    "vhmm_forward": (ctypes.c_int, [_DP, _PP, _P, _I64, _I64, _P, _P, _P, _P, _P, _SZ, _P]),
    "vhmm_prior_fwd": (ctypes.c_int, [_DP, _PP, _P, _I64, _I64, _P, _P, _P, _SZ, _P]),
    "vhmm_decoder_fwd": (ctypes.c_int, [_DP, _PP, _P, _I64, _I64, _P, _P, _P, _SZ, _P]),
    "vhmm_elbo": (ctypes.c_int, [_DP, _PP, _P, _P, _P, _I64, _I64, _P, _DBL, _P, _P, _P, _SZ, _P]),
    "vhmm_elbo_select": (ctypes.c_int, [ctypes.c_int]),
    "vhmm_elbo_active_impl": (ctypes.c_int, [_DP]),
    "vhmm_elbo_decoder_select": (ctypes.c_int, [ctypes.c_int]),
    "vhmm_elbo_decoder_active": (ctypes.c_int, []),
    "vhmm_tma_select": (ctypes.c_int, [ctypes.c_int]),
    "vhmm_elbo_timing": (ctypes.c_int, [ctypes.c_int]),
    "vhmm_elbo_timing_read": (ctypes.c_int, [ctypes.POINTER(ctypes.c_double), ctypes.POINTER(ctypes.c_int)]),
    "vhmm_conv_select": (ctypes.c_int, [ctypes.c_int]),
    "vhmm_prior_select": (ctypes.c_int, [ctypes.c_int]),
    "vhmm_emb_fold_select": (ctypes.c_int, [ctypes.c_int]),
    "vhmm_conv1d_workspace_bytes": (_SZ, [_I64, _I64, _I64]),
    "vhmm_conv1d_fwd": (ctypes.c_int, [_P, _P, _P, _I64, _I64, _I64, _I64, _I64, ctypes.c_int, _P, _P, _SZ, _P]),
    "vhmm_argmax_regimes": (ctypes.c_int, [_P, _I64, _I64, _I64, _P, _P]),
    "vhmm_gather_chunks": (ctypes.c_int, [_P, _P, _I64, _I64, _I64, _P, _P, _I64, _I64, _P, _P, _P]),
    "vhmm_train_step": (ctypes.c_int, [_DP, _PP, _P, _P, _I64, _P, _P, _I64, _I64, _P, _P, _P, _DBL, _DBL, _DBL, _DBL, _P, _P, _P, _P,
                                       _SZ, _P]),
    "vhmm_adam_step": (ctypes.c_int, [_DP, _PP, _P, _P, _P, _DBL, _DBL, _DBL, _DBL, _I64, _P, _DBL, _P, _P, _P]),
    "vhmm_emission_table": (ctypes.c_int, [_DP, _PP, _I64, _P, _P, _P, _SZ, _P]),
    "vhmm_emission_logprob": (ctypes.c_int, [_P, _P, _P, _I64, _I64, _I64, _I64, _P, _P]),
    "vhmm_hmm_inputs_supported": (ctypes.c_int, [_DP, _I64]),
    "vhmm_hmm_inputs": (ctypes.c_int, [_DP, _PP, _P, _P, _I64, _I64, _P, _P, _P, _P, _P, _SZ, _P]),
    "vhmm_hmm_loglik_fused": (ctypes.c_int, [_DP, _PP, _P, _P, _P, _P, _I64, _I64, _P, _P, _P, _P, _SZ, _P]),
    "vhmm_transition_probs": (ctypes.c_int, [_DP, _PP, _P, _I64, _I64, _P, _P, _SZ, _P]),
    "vhmm_hmm_param_grads": (ctypes.c_int, [_DP, _PP, _P, _P, _P, _P, _P, _P, _P, _P, _I64, _I64, _P, _P, _SZ, _P]),
    "vhmm_hmm_forward": (ctypes.c_int, [_P, _P, _P, _P, _I64, _I64, _I64, _P, _P, _P]),
    "vhmm_hmm_posterior": (ctypes.c_int, [_P, _P, _P, _P, _I64, _I64, _I64, _P, _P, _P, _P, _SZ, _P]),
    "vhmm_hmm_posterior_select": (ctypes.c_int, [ctypes.c_int]),
    "vhmm_hmm_numerics_select": (ctypes.c_int, [ctypes.c_int]),
    "vhmm_viterbi": (ctypes.c_int, [_P, _P, _P, _P, _I64, _I64, _I64, _P, _P, _P, _SZ, _P]),
    "vhmm_viterbi_static": (ctypes.c_int, [_P, _P, _P, _P, _I64, _I64, _I64, _P, _P, _P]),
    "vhmm_hmm_decode_select": (ctypes.c_int, [ctypes.c_int]),
    "vhmm_hmm_decode": (ctypes.c_int, [_P, _P, _P, _P, _I64, _I64, _I64, _P, _P, _P, _P, _SZ, _P]),
    "vhmm_hmm_scan": (ctypes.c_int, [_P, _P, _P, _P, _I64, _I64, _I64, _P, _P, _P, _SZ, _P]),
}


def lib() -> ctypes.CDLL:
    """Load (building first if the sources are newer) and return the bound library."""
    global _lib
    if _lib is not None:
        return _lib
    with _lock:
        if _lib is not None:
            return _lib
        if needs_build():
            build()
        try:
            L = ctypes.CDLL(LIB_PATH)
        except OSError as e:  # no fallback: fail loudly
            raise VhmmError(f"cannot load {LIB_PATH}: {e}") from e
        for name, (res, args) in PROTOTYPES.items():
            fn = getattr(L, name)
            fn.restype = res
            fn.argtypes = args
        _lib = L
    return _lib


def check(rc: int, what: str):
    if rc != 0:
        msg = lib().vhmm_last_error().decode("utf-8", "replace")
        kind = "invalid argument" if rc < 0 else "CUDA error"
        raise VhmmError(f"{what} failed ({kind} {rc}): {msg}")
