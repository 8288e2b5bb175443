"""Tensor-level wrappers around the C ABI: pointer plumbing, workspace reuse, and the
torch.autograd.Function objects (ELBO; exact-HMM log-likelihood whose backward is the
beta-recursion kernel).  PyTorch is used for device memory, streams and autograd bookkeeping
only -- every number is produced by libvhmm.so.  No CPU fallback: non-CUDA tensors raise.
"""
from __future__ import annotations

import ctypes
import os
from typing import Dict, Optional, Sequence, Tuple

import torch

from . import _lib
from ._lib import Dims, Params, VhmmError, check

PARAM_ORDER = (
    "encoder.conv1.weight", "encoder.conv1.bias",
    "encoder.conv2.weight", "encoder.conv2.bias",
    "encoder.to_logits.weight", "encoder.to_logits.bias",
    "prior.log_prior",
    "prior.transition_net.0.weight", "prior.transition_net.0.bias",
    "prior.transition_net.2.weight", "prior.transition_net.2.bias",
    "decoder.embeddings.weight",
    "decoder.conv1.weight", "decoder.conv1.bias",
    "decoder.conv2.weight", "decoder.conv2.bias",
    "decoder.to_params.weight", "decoder.to_params.bias",
)

_workspaces: Dict[Tuple[int, int], torch.Tensor] = {}
ELBO_WORKSPACE_CAP = 12 << 30  # bytes: upper bound for the layer-wise path's activation stash


def _require_cuda(*tensors):
    for t in tensors:
        if t is not None and not t.is_cuda:
            raise VhmmError("vhmm ops run on CUDA tensors only (no CPU fallback); got a "
                            f"{t.device} tensor")


def _f32c(t: torch.Tensor) -> torch.Tensor:
    if t.dtype != torch.float32:
        t = t.float()
    return t if t.is_contiguous() else t.contiguous()


def _stream(dev) -> ctypes.c_void_p:
    return ctypes.c_void_p(torch.cuda.current_stream(dev).cuda_stream)


def _ptr(t: Optional[torch.Tensor]):
    return ctypes.c_void_p(0 if t is None else t.data_ptr())


def workspace(dev: torch.device, nbytes: int) -> torch.Tensor:
    """Per-(device, stream) scratch buffer, grown on demand and reused across calls."""
    key = (dev.index if dev.index is not None else torch.cuda.current_device(),
           torch.cuda.current_stream(dev).cuda_stream)
    nbytes = max(int(nbytes), 256)
    ws = _workspaces.get(key)
    if ws is None or ws.numel() < nbytes:
        _workspaces.pop(key, None)
        ws = None
        ws = torch.empty(nbytes, dtype=torch.uint8, device=dev)
        _workspaces[key] = ws
    return ws


def release_workspaces():
    _workspaces.clear()


def make_dims(C, H, K, H2, U, TH) -> Dims:
    return Dims(int(C), int(H), int(K), int(H2), int(U), int(TH))


def pack_params(tensors: Sequence[torch.Tensor]) -> Tuple[Params, list]:
    """18 tensors in PARAM_ORDER -> (struct of device pointers, keep-alive list)."""
    if len(tensors) != _lib.NUM_PARAMS:
        raise VhmmError(f"expected {_lib.NUM_PARAMS} parameter tensors, got {len(tensors)}")
    keep = [_f32c(t.detach()) for t in tensors]
    _require_cuda(*keep)
    P = Params()
    for i, t in enumerate(keep):
        P.p[i] = t.data_ptr()
    return P, keep


def param_count(d: Dims) -> int:
    return int(_lib.lib().vhmm_param_count(ctypes.byref(d)))


def param_offsets(d: Dims):
    out = (ctypes.c_int64 * (_lib.NUM_PARAMS + 1))()
    check(_lib.lib().vhmm_param_offsets(ctypes.byref(d), out), "vhmm_param_offsets")
    return list(out)


def _ws_bytes(op, d, B, T) -> int:
    return int(_lib.lib().vhmm_workspace_bytes(op, ctypes.byref(d) if d is not None else None, B, T))


# ---------------------------------------------------------------------------------------
# pinned path
# ---------------------------------------------------------------------------------------
def encoder_fwd(d: Dims, params, x, want_q=False, want_logits=True):
    x = _f32c(x); _require_cuda(x)
    B, C, T = x.shape
    if C != d.C:
        raise VhmmError(f"encode: expected {d.C} input channels, got {C}")
    P, keep = pack_params(params)
    logits = torch.empty(B, d.K, T, device=x.device, dtype=torch.float32) if want_logits else None
    q = torch.empty(B, d.K, T, device=x.device, dtype=torch.float32) if want_q else None
    with torch.cuda.device(x.device):
        ws = workspace(x.device, _ws_bytes(_lib.OP_ENCODER_FWD, d, B, T))
        check(_lib.lib().vhmm_encoder_fwd(ctypes.byref(d), ctypes.byref(P), _ptr(x), B, T, _ptr(logits), _ptr(q),
                                          _ptr(ws), ws.numel(), _stream(x.device)), "vhmm_encoder_fwd")
    return logits, q


def forward_fwd(d: Dims, params, x, want_logits=False):
    """VAE_HMM.forward (VQ_VAE_HMM_fixed.py:139-143) in one library call: ((mu, logvar), q[, logits])."""
    x = _f32c(x); _require_cuda(x)
    B, C, T = x.shape
    if C != d.C:
        raise VhmmError(f"forward: expected {d.C} input channels, got {C}")
    P, keep = pack_params(params)
    q = torch.empty(B, d.K, T, device=x.device, dtype=torch.float32)
    logits = torch.empty(B, d.K, T, device=x.device, dtype=torch.float32) if want_logits else None
    mu = torch.empty(B, d.C, T, device=x.device, dtype=torch.float32)
    logvar = torch.empty(B, d.C, T, device=x.device, dtype=torch.float32)
    with torch.cuda.device(x.device):
        ws = workspace(x.device, _ws_bytes(_lib.OP_FORWARD, d, B, T))
        check(_lib.lib().vhmm_forward(ctypes.byref(d), ctypes.byref(P), _ptr(x), B, T, _ptr(logits), _ptr(q), _ptr(mu),
                                      _ptr(logvar), _ptr(ws), ws.numel(), _stream(x.device)), "vhmm_forward")
    return (mu, logvar), q, logits


def prior_fwd(d: Dims, params, u):
    """u in the (B,U,T) layout."""
    u = _f32c(u); _require_cuda(u)
    B, U, T = u.shape
    if U != d.U:
        raise VhmmError(f"prior: expected u_dim {d.U}, got {U}")
    P, keep = pack_params(params)
    log_pi = torch.empty(d.K, device=u.device, dtype=torch.float32)
    log_A = torch.empty(B, T, d.K, d.K, device=u.device, dtype=torch.float32)
    with torch.cuda.device(u.device):
        ws = workspace(u.device, _ws_bytes(_lib.OP_PRIOR_FWD, d, B, T))
        check(_lib.lib().vhmm_prior_fwd(ctypes.byref(d), ctypes.byref(P), _ptr(u), B, T, _ptr(log_pi), _ptr(log_A),
                                        _ptr(ws), ws.numel(), _stream(u.device)), "vhmm_prior_fwd")
    return log_pi, log_A


def decoder_fwd(d: Dims, params, q):
    q = _f32c(q); _require_cuda(q)
    B, K, T = q.shape
    if K != d.K:
        raise VhmmError(f"decode: expected K={d.K} rows, got {K}")
    P, keep = pack_params(params)
    mu = torch.empty(B, d.C, T, device=q.device, dtype=torch.float32)
    logvar = torch.empty(B, d.C, T, device=q.device, dtype=torch.float32)
    with torch.cuda.device(q.device):
        ws = workspace(q.device, _ws_bytes(_lib.OP_DECODER_FWD, d, B, T))
        check(_lib.lib().vhmm_decoder_fwd(ctypes.byref(d), ctypes.byref(P), _ptr(q), B, T, _ptr(mu), _ptr(logvar),
                                          _ptr(ws), ws.numel(), _stream(q.device)), "vhmm_decoder_fwd")
    return mu, logvar


def conv_select(impl: str = "auto"):
    """Layer kernels of the layer-wise path: "auto" | "ffma" | "tc5" (include/vhmm.h: vhmm_conv_select)."""
    check(_lib.lib().vhmm_conv_select({"auto": 0, "ffma": 1, "tc5": 2}[impl]), "vhmm_conv_select")


def conv1d(x, weight, bias=None, relu=False):
    """One Conv1d layer (kernel width 1 or 3, padding kw // 2) through the layer kernels of libvhmm.so."""
    x = _f32c(x); weight = _f32c(weight); _require_cuda(x, weight)
    B, Cin, T = x.shape
    Cout, Cin2, kw = weight.shape
    if Cin2 != Cin:
        raise VhmmError(f"conv1d: weight {tuple(weight.shape)} does not match x {tuple(x.shape)}")
    bias = _f32c(bias) if bias is not None else None
    y = torch.empty(B, Cout, T, device=x.device, dtype=torch.float32)
    with torch.cuda.device(x.device):
        ws = workspace(x.device, int(_lib.lib().vhmm_conv1d_workspace_bytes(Cin, Cout, kw)))
        check(_lib.lib().vhmm_conv1d_fwd(_ptr(x), _ptr(weight), _ptr(bias), B, Cin, Cout, T, kw, int(bool(relu)), _ptr(y),
                                         _ptr(ws), ws.numel(), _stream(x.device)), "vhmm_conv1d_fwd")
    return y


def elbo_norms(lengths_dev: torch.Tensor, B: int, C: int, T: int) -> torch.Tensor:
    """device double[2] = {max(N_valid*C, 1), B}: the reference's normalisers
    (VQ_VAE_HMM_fixed.py:120, :131, :135) for THIS process's sequences."""
    nv = lengths_dev.clamp(min=0, max=T).sum().to(torch.float64)
    return torch.stack([(nv * C).clamp(min=1.0), torch.tensor(float(B), dtype=torch.float64, device=lengths_dev.device)])


_stream_state = {}


def _stream_growth() -> float:
    e = os.environ.get("VHMM_STREAM_GROWTH")
    if e:
        return max(1.0, float(e))
    try:
        import torch.distributed as dist
        if dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1:
            return 1.4
    except Exception:
        pass
    return 2.0


def elbo_raw_streamed(d: Dims, params, x, u, lengths_dev, norms, beta: float, want_grads: bool, chunk: int = 0):
    """vhmm_elbo over a HOST-resident batch (pinned memory recommended): the batch is cut into
    chunks of sequences that are copied host->device on a side stream while the previous chunk
    computes (double-buffered staging, events both ways).  Sequences are independent and the
    normalisers are global, so the un-normalised sums and the gradients of the chunks add up to
    exactly the single-call result (up to fp32 summation order).  Chunks run in a fixed order."""
    dev = lengths_dev.device
    B, C, T = x.shape
    U = u.shape[1]
    # chunk schedule.  Explicit ``chunk``: uniform.  Default: geometric -- B/32, B/32, B/16, B/8, then B/4 --
    # so that only a small first chunk's copy is exposed (the kernels are ~4x slower per byte than the copy,
    # a chunk twice the size of its predecessor still arrives in time) while the per-call cost (kernel
    # prologues, the 64 MB partial reduction of the finish kernel: ~0.19 ms) is paid 7 times, not 16-32.
    # Measured at cfg 2 (tools/stream_chunk_probe.py): uniform 8 / 16 / 32 chunks = 97.0 / 97.8 / 99.2 ms.
    if chunk > 0:
        chunk = min(chunk, B)
        sizes = [min(chunk, B - b0) for b0 in range(0, B, chunk)]
    else:
        # A chunk `growth` times the size of its predecessor arrives in time only if the copy runs `growth` times faster per
        # byte than the kernels (14 GB/s at cfg 2).  One process per box: PCIe gives ~50 GB/s -> growth 2.  Several ranks
        # streaming from the same host memory (torch.distributed world > 1) share its bandwidth (8 ranks: 93.6 ms per step
        # with growth 2 against 85.2 resident) -> growth 1.4, three more chunks.  VHMM_STREAM_GROWTH overrides.
        growth = _stream_growth()
        cap = max(1024, -(-B // (4 if growth >= 2.0 else 6)))
        cur = max(1024, -(-B // 32))
        sizes, left, first = [], B, True
        while left > 0:
            n = min(cur, left)
            sizes.append(n)
            left -= n
            if first:
                first = False          # the first size is used twice
            else:
                cur = min(cap, int(cur * growth))
    big = max(sizes)
    key = (dev.index, big, C, U, T)
    st = _stream_state.get(key)
    if st is None:
        _stream_state.clear()           # one staging set at a time (it is hundreds of MB)
        st = {"copy": torch.cuda.Stream(device=dev),
              "buf": [(torch.empty(big, C, T, device=dev), torch.empty(big, U, T, device=dev)) for _ in range(2)],
              "copied": [torch.cuda.Event(), torch.cuda.Event()], "done": [torch.cuda.Event(), torch.cuda.Event()]}
        _stream_state[key] = st
    cs = torch.cuda.current_stream(dev)
    copy_s = st["copy"]
    copy_s.wait_stream(cs)              # staging buffers may still be read by earlier work on the compute stream
    sums_t = torch.zeros(4, device=dev, dtype=torch.float64)
    grads_t = torch.zeros(param_count(d), device=dev, dtype=torch.float32) if want_grads else None
    b0 = 0
    for i, n in enumerate(sizes):
        b1, k = b0 + n, i % 2
        xb, ub = st["buf"][k]
        with torch.cuda.stream(copy_s):
            if i >= 2:
                copy_s.wait_event(st["done"][k])
            xb[:n].copy_(x[b0:b1], non_blocking=True)
            ub[:n].copy_(u[b0:b1], non_blocking=True)
            st["copied"][k].record(copy_s)
        cs.wait_event(st["copied"][k])
        sums, grads = elbo_raw(d, params, xb[:n], ub[:n], lengths_dev[b0:b1], norms, beta, want_grads)
        sums_t += sums
        if want_grads:
            grads_t += grads
        st["done"][k].record(cs)
        b0 = b1
    return sums_t, grads_t


def elbo_raw(d: Dims, params, x, u, lengths_dev, norms, beta: float, want_grads: bool):
    """One call of vhmm_elbo.  Returns (sums double[4] on device, flat grads or None).
    Host-resident x/u (CPU tensors, the model on a GPU) are streamed: elbo_raw_streamed."""
    if x.device.type == "cpu" and u.device.type == "cpu" and lengths_dev.is_cuda:
        return elbo_raw_streamed(d, params, _f32c(x), _f32c(u), lengths_dev, norms, beta, want_grads)
    x = _f32c(x); u = _f32c(u); _require_cuda(x, u, lengths_dev, norms)
    B, C, T = x.shape
    if C != d.C or u.shape != (B, d.U, T):
        raise VhmmError(f"compute_loss: shape mismatch x{tuple(x.shape)} u{tuple(u.shape)} for dims C={d.C} U={d.U}")
    if lengths_dev.dtype != torch.int64 or lengths_dev.numel() != B:
        raise VhmmError("compute_loss: lengths must be int64 of shape (B,)")
    P, keep = pack_params(params)
    sums = torch.empty(4, device=x.device, dtype=torch.float64)
    grads = torch.empty(param_count(d), device=x.device, dtype=torch.float32) if want_grads else None
    with torch.cuda.device(x.device):
        need = _ws_bytes(_lib.OP_ELBO, d, B, T)
        if _lib.lib().vhmm_elbo_active_impl(ctypes.byref(d)) == _lib.ELBO_FUSED:
            ws = workspace(x.device, need)               # exact: no activation stash
        else:                                            # layer-wise: enough for min(B,1024) sequences,
            per_seq = max(1, (need // max(1, min(B, 1024))))   # more lets it chunk the batch less
            want = min(ELBO_WORKSPACE_CAP, per_seq * B + (64 << 20))
            ws = workspace(x.device, max(need, want))
        check(_lib.lib().vhmm_elbo(ctypes.byref(d), ctypes.byref(P), _ptr(x), _ptr(u), _ptr(lengths_dev), B, T,
                                   _ptr(norms), float(beta), _ptr(sums), _ptr(grads), _ptr(ws), ws.numel(),
                                   _stream(x.device)), "vhmm_elbo")
    return sums, grads


def elbo_select(impl: str = "auto"):
    """Choose the vhmm_elbo implementation: "auto" | "layerwise" | "fused" (include/vhmm.h)."""
    code = {"auto": _lib.ELBO_AUTO, "layerwise": _lib.ELBO_LAYERWISE, "fused": _lib.ELBO_FUSED}[impl]
    check(_lib.lib().vhmm_elbo_select(code), "vhmm_elbo_select")


def elbo_decoder_select(variant: str = "auto"):
    """Decoder kernel of the fused path: "auto" | "mma" (mma.sync 3xTF32) | "tc5" (tcgen05 2xFP16)."""
    code = {"auto": _lib.DECODER_AUTO, "mma": _lib.DECODER_MMA, "tc5": _lib.DECODER_TC5}[variant]
    check(_lib.lib().vhmm_elbo_decoder_select(code), "vhmm_elbo_decoder_select")


def prior_select(mode: str = "auto"):
    """Transition prior of the layer-wise path at K = 32: "auto" (fused kernel, operand-form gradient) | "fused_f32" |
    "layers" (include/vhmm.h: vhmm_prior_select)."""
    check(_lib.lib().vhmm_prior_select({"auto": 0, "fused_f32": 1, "layers": 2}[mode]), "vhmm_prior_select")


def emb_fold_select(on: bool):
    """decoder.embeddings folded into decoder.conv1 in the layer-wise ELBO (include/vhmm.h: vhmm_emb_fold_select)."""
    check(_lib.lib().vhmm_emb_fold_select(1 if on else 0), "vhmm_emb_fold_select")


def tma_select(on: bool):
    """Input windows of the tile kernels by tensor-map TMA (True) or cp.async (False, default).  include/vhmm.h."""
    check(_lib.lib().vhmm_tma_select(1 if on else 0), "vhmm_tma_select")


def hmm_posterior_select(variant: str = "auto"):
    """Posterior kernels for K <= 4: "auto" (one-pass warp kernel when it applies) | "sweep"
    (forward + backward sweep, one thread per sequence).  include/vhmm.h."""
    code = {"auto": 0, "sweep": 1}[variant]
    check(_lib.lib().vhmm_hmm_posterior_select(code), "vhmm_hmm_posterior_select")


def hmm_numerics_select(mode: str = "scaled"):
    """Sum-product numerics of the exact-HMM stages: "scaled" (default, tuned fp32 scaled-linear kernels) | "logspace"
    (logsumexp recursions: any dynamic range, e.g. structural zeros under sharp emissions).  include/vhmm.h."""
    check(_lib.lib().vhmm_hmm_numerics_select({"scaled": 0, "logspace": 1}[mode]), "vhmm_hmm_numerics_select")


def elbo_decoder_active() -> str:
    return {_lib.DECODER_MMA: "mma", _lib.DECODER_TC5: "tc5"}[int(_lib.lib().vhmm_elbo_decoder_active())]


def elbo_active_impl(d: Dims) -> str:
    return {_lib.ELBO_LAYERWISE: "layerwise", _lib.ELBO_FUSED: "fused"}[int(_lib.lib().vhmm_elbo_active_impl(ctypes.byref(d)))]


def loss_from_sums(sums: torch.Tensor, norms: torch.Tensor, beta: float) -> torch.Tensor:
    """recon + beta*(prior_loss - entropy) from un-normalised sums (VQ_VAE_HMM_fixed.py:137)."""
    recon = sums[0] / norms[0]
    prior_loss = -sums[1] / norms[1]
    ent = sums[2] / norms[1]
    return (recon + beta * (prior_loss - ent)).to(torch.float32)


class ElboFunction(torch.autograd.Function):
    """loss = ELBO(x,u,lengths,beta; 18 params).  The gradient is produced by the same
    library call as the loss (fused forward+backward); backward() only scales it."""

    @staticmethod
    def forward(ctx, d, x, u, lengths_dev, beta, norms, *params):
        want = any(ctx.needs_input_grad)   # (grad mode is always off inside forward)
        sums, grads = elbo_raw(d, params, x, u, lengths_dev, norms, beta, want)
        ctx.has_grads = grads is not None
        ctx.shapes = [tuple(p.shape) for p in params]
        ctx.offsets = param_offsets(d)
        ctx.save_for_backward(grads if grads is not None else sums)
        ctx.mark_non_differentiable(sums)
        return loss_from_sums(sums, norms, beta), sums

    @staticmethod
    def backward(ctx, gout, _gsums):
        if not ctx.has_grads:
            raise VhmmError("ELBO gradients were not computed (forward ran without grad)")
        (flat,) = ctx.saved_tensors
        flat = flat * gout.to(flat.dtype)
        outs = []
        for i, shp in enumerate(ctx.shapes):
            outs.append(flat[ctx.offsets[i]:ctx.offsets[i + 1]].view(shp))
        return (None, None, None, None, None, None, *outs)


def argmax_regimes(q):
    q = _f32c(q); _require_cuda(q)
    B, K, T = q.shape
    out = torch.empty(B, T, device=q.device, dtype=torch.int64)
    with torch.cuda.device(q.device):
        check(_lib.lib().vhmm_argmax_regimes(_ptr(q), B, K, T, _ptr(out), _stream(q.device)), "vhmm_argmax_regimes")
    return out


# ---------------------------------------------------------------------------------------
# exact-HMM stages (extensions)
# ---------------------------------------------------------------------------------------
def emission_table(d: Dims, params, T: int, device):
    P, keep = pack_params(params)
    mu = torch.empty(d.K, d.C, T, device=device, dtype=torch.float32)
    lv = torch.empty(d.K, d.C, T, device=device, dtype=torch.float32)
    with torch.cuda.device(device):
        ws = workspace(mu.device, _ws_bytes(_lib.OP_EMISSION_TABLE, d, d.K, T))
        check(_lib.lib().vhmm_emission_table(ctypes.byref(d), ctypes.byref(P), T, _ptr(mu), _ptr(lv), _ptr(ws),
                                             ws.numel(), _stream(mu.device)), "vhmm_emission_table")
    return mu, lv


def emission_logprob(x, mu_tab, logvar_tab):
    x = _f32c(x); mu_tab = _f32c(mu_tab); logvar_tab = _f32c(logvar_tab); _require_cuda(x, mu_tab, logvar_tab)
    B, C, T = x.shape
    K = mu_tab.shape[0]
    out = torch.empty(B, T, K, device=x.device, dtype=torch.float32)
    with torch.cuda.device(x.device):
        check(_lib.lib().vhmm_emission_logprob(_ptr(x), _ptr(mu_tab), _ptr(logvar_tab), B, C, T, K, _ptr(out),
                                               _stream(x.device)), "vhmm_emission_logprob")
    return out


def hmm_inputs_fused(d: Dims, params, x, u, mu_tab, logvar_tab):
    """(log_A (B,T,K,K), log_b (B,T,K)) from raw x, u in one kernel (vhmm_hmm_inputs), or None when the
    dims / T are outside its instantiation (the caller then uses prior_fwd + emission_logprob)."""
    x = _f32c(x); u = _f32c(u); _require_cuda(x, u, mu_tab, logvar_tab)
    B, C, T = x.shape
    if not _lib.lib().vhmm_hmm_inputs_supported(ctypes.byref(d), T) or u.shape != (B, d.U, T):
        return None
    log_A = torch.empty(B, T, d.K, d.K, device=x.device, dtype=torch.float32)
    log_b = torch.empty(B, T, d.K, device=x.device, dtype=torch.float32)
    P, keep = pack_params(params)
    with torch.cuda.device(x.device):
        ws = workspace(x.device, _ws_bytes(_lib.OP_HMM_INPUTS, d, B, T))
        check(_lib.lib().vhmm_hmm_inputs(ctypes.byref(d), ctypes.byref(P), _ptr(x), _ptr(u), B, T, _ptr(_f32c(mu_tab)),
                                         _ptr(_f32c(logvar_tab)), _ptr(log_A), _ptr(log_b), _ptr(ws), ws.numel(),
                                         _stream(x.device)), "vhmm_hmm_inputs")
    return log_A, log_b


def hmm_loglik_fused(d: Dims, params, log_pi, x, u, lengths, mu_tab, logvar_tab):
    """loglik (B) straight from raw x, u: transition softmax + emission log-probabilities + forward recursion
    in ONE kernel (vhmm_hmm_loglik_fused), or None when the dims / T are outside its instantiation."""
    x = _f32c(x); u = _f32c(u); _require_cuda(x, u, mu_tab, logvar_tab)
    B, C, T = x.shape
    if not _lib.lib().vhmm_hmm_inputs_supported(ctypes.byref(d), T) or u.shape != (B, d.U, T):
        return None
    log_pi = _f32c(log_pi.to(x.device))
    if lengths is not None:
        lengths = lengths.to(device=x.device, dtype=torch.int64).contiguous()
    out = torch.empty(B, device=x.device, dtype=torch.float32)
    P, keep = pack_params(params)
    with torch.cuda.device(x.device):
        ws = workspace(x.device, _ws_bytes(_lib.OP_HMM_INPUTS, d, B, T))
        check(_lib.lib().vhmm_hmm_loglik_fused(ctypes.byref(d), ctypes.byref(P), _ptr(log_pi), _ptr(x), _ptr(u),
                                               _ptr(lengths) if lengths is not None else None, B, T,
                                               _ptr(_f32c(mu_tab)), _ptr(_f32c(logvar_tab)), _ptr(out), _ptr(ws),
                                               ws.numel(), _stream(x.device)), "vhmm_hmm_loglik_fused")
    return out


def _hmm_inputs(log_pi, log_A, log_b, lengths):
    log_pi = _f32c(log_pi); log_A = _f32c(log_A); log_b = _f32c(log_b)
    _require_cuda(log_pi, log_A, log_b)
    B, T, K = log_b.shape
    if log_A.shape != (B, T, K, K) or log_pi.shape != (K,):
        raise VhmmError(f"hmm: shape mismatch log_pi{tuple(log_pi.shape)} log_A{tuple(log_A.shape)} log_b{tuple(log_b.shape)}")
    if lengths is None:
        lengths = torch.full((B,), T, dtype=torch.int64, device=log_b.device)
    lengths = lengths.to(device=log_b.device, dtype=torch.int64).contiguous()
    return log_pi, log_A, log_b, lengths, B, T, K


def hmm_forward(log_pi, log_A, log_b, lengths=None, want_alpha=False, time_parallel=None):
    """vhmm_hmm_forward: the forward recursion as ONE pass (no workspace).  Sequences up to ~580 steps (K <= 4) run
    time-parallel on one warp each; longer ones one thread per sequence -- which leaves FEW LONG sequences (T >= 1024,
    fewer than 2368 of them: B = 64, T = 65536 takes 43 ms) on a handful of threads, so those go to ``hmm_scan`` (same
    results and padding conventions, time-parallel over 256-step segments, 0.05 - 0.12 ms) unless ``time_parallel=False``."""
    if time_parallel is None:
        time_parallel = log_b.dim() == 3 and log_b.shape[1] >= 1024 and log_b.shape[0] < 2368 and log_b.shape[2] <= 4 \
            and (log_b.shape[1] * log_b.shape[2]) % 4 == 0
        # 4 < K <= 32: segment products + carry (+ the forward kernel on 128-step rows for alpha): csrc/hmm_wide_scan.cuh
        time_parallel = time_parallel or (log_b.dim() == 3 and 4 < log_b.shape[2] <= 32
                                          and log_b.shape[1] >= 512 and log_b.shape[0] <= 256)
    if time_parallel:
        return hmm_scan(log_pi, log_A, log_b, lengths, want_alpha=want_alpha)
    log_pi, log_A, log_b, lengths, B, T, K = _hmm_inputs(log_pi, log_A, log_b, lengths)
    ll = torch.empty(B, device=log_b.device, dtype=torch.float32)
    alpha = torch.empty(B, T, K, device=log_b.device, dtype=torch.float32) if want_alpha else None
    with torch.cuda.device(log_b.device):
        check(_lib.lib().vhmm_hmm_forward(_ptr(log_pi), _ptr(log_A), _ptr(log_b), _ptr(lengths), B, T, K, _ptr(ll),
                                          _ptr(alpha), _stream(log_b.device)), "vhmm_hmm_forward")
    return (ll, alpha) if want_alpha else ll


def hmm_posterior(log_pi, log_A, log_b, lengths=None, want_xi=True):
    log_pi, log_A, log_b, lengths, B, T, K = _hmm_inputs(log_pi, log_A, log_b, lengths)
    dev = log_b.device
    ll = torch.empty(B, device=dev, dtype=torch.float32)
    gamma = torch.empty(B, T, K, device=dev, dtype=torch.float32)
    xi = torch.empty(B, T, K, K, device=dev, dtype=torch.float32) if want_xi else None
    dk = Dims(1, 1, K, 1, 1, 1)
    with torch.cuda.device(dev):
        ws = workspace(dev, _ws_bytes(_lib.OP_HMM_POSTERIOR, dk, B, T))
        check(_lib.lib().vhmm_hmm_posterior(_ptr(log_pi), _ptr(log_A), _ptr(log_b), _ptr(lengths), B, T, K, _ptr(ll),
                                            _ptr(gamma), _ptr(xi), _ptr(ws), ws.numel(), _stream(dev)),
              "vhmm_hmm_posterior")
    return ll, gamma, xi


def viterbi(log_pi, log_A, log_b, lengths=None):
    log_pi, log_A, log_b, lengths, B, T, K = _hmm_inputs(log_pi, log_A, log_b, lengths)
    dev = log_b.device
    path = torch.empty(B, T, device=dev, dtype=torch.uint8)
    score = torch.empty(B, device=dev, dtype=torch.float32)
    dk = Dims(1, 1, K, 1, 1, 1)
    with torch.cuda.device(dev):
        ws = workspace(dev, _ws_bytes(_lib.OP_VITERBI, dk, B, T))
        check(_lib.lib().vhmm_viterbi(_ptr(log_pi), _ptr(log_A), _ptr(log_b), _ptr(lengths), B, T, K, _ptr(path),
                                      _ptr(score), _ptr(ws), ws.numel(), _stream(dev)), "vhmm_viterbi")
    return path, score


def hmm_decode_select(mode: str = "auto"):
    """cfg-3 decode: "auto" (Viterbi sweep + one-pass posterior) | "one" (one kernel, inputs read once; measured slower).
    include/vhmm.h."""
    check(_lib.lib().vhmm_hmm_decode_select({"auto": 0, "one": 1}[mode]), "vhmm_hmm_decode_select")


def hmm_decode(log_pi, log_A, log_b, lengths=None):
    """Viterbi path + posterior gamma + loglik from one forward sweep (cfg 3)."""
    log_pi, log_A, log_b, lengths, B, T, K = _hmm_inputs(log_pi, log_A, log_b, lengths)
    dev = log_b.device
    path = torch.empty(B, T, device=dev, dtype=torch.uint8)
    gamma = torch.empty(B, T, K, device=dev, dtype=torch.float32)
    ll = torch.empty(B, device=dev, dtype=torch.float32)
    dk = Dims(1, 1, K, 1, 1, 1)
    with torch.cuda.device(dev):
        ws = workspace(dev, _ws_bytes(_lib.OP_HMM_DECODE, dk, B, T))
        check(_lib.lib().vhmm_hmm_decode(_ptr(log_pi), _ptr(log_A), _ptr(log_b), _ptr(lengths), B, T, K, _ptr(path),
                                         _ptr(gamma), _ptr(ll), _ptr(ws), ws.numel(), _stream(dev)), "vhmm_hmm_decode")
    return path, gamma, ll


def hmm_scan(log_pi, log_A, log_b, lengths=None, want_alpha=False):
    log_pi, log_A, log_b, lengths, B, T, K = _hmm_inputs(log_pi, log_A, log_b, lengths)
    dev = log_b.device
    ll = torch.empty(B, device=dev, dtype=torch.float32)
    alpha = torch.empty(B, T, K, device=dev, dtype=torch.float32) if want_alpha else None
    dk = Dims(1, 1, K, 1, 1, 1)
    with torch.cuda.device(dev):
        ws = workspace(dev, _ws_bytes(_lib.OP_HMM_SCAN, dk, B, T))
        check(_lib.lib().vhmm_hmm_scan(_ptr(log_pi), _ptr(log_A), _ptr(log_b), _ptr(lengths), B, T, K, _ptr(ll),
                                       _ptr(alpha), _ptr(ws), ws.numel(), _stream(dev)), "vhmm_hmm_scan")
    return (ll, alpha) if want_alpha else ll


class HmmLogLikFunction(torch.autograd.Function):
    """loglik[b] of the exact HMM (B.2); backward is the beta-recursion kernel (B.3):
    d loglik / d log_b = gamma, d loglik / d log_A = xi, d loglik / d log_pi = sum_b gamma[b,0]."""

    @staticmethod
    def forward(ctx, log_pi, log_A, log_b, lengths):
        if any(ctx.needs_input_grad):
            ll, gamma, xi = hmm_posterior(log_pi, log_A, log_b, lengths, want_xi=True)
            ctx.save_for_backward(gamma, xi)
            ctx.ok = True
        else:
            ll = hmm_forward(log_pi, log_A, log_b, lengths)
            ctx.ok = False
        return ll

    @staticmethod
    def backward(ctx, g):
        if not ctx.ok:
            raise VhmmError("hmm log-likelihood gradients were not computed")
        gamma, xi = ctx.saved_tensors
        g = g.to(gamma.dtype)
        g_b = gamma * g[:, None, None]
        g_A = xi * g[:, None, None, None]
        g_pi = g_b[:, 0, :].sum(dim=0)
        return g_pi, g_A, g_b, None


def transition_probs(d: Dims, params, u):
    """A = exp(log_A) (B,T,K,K), rows summing to 1: the `transition_matrix` / `trans_matrix` / `transition_probs`
    argument of the reference's consumers (regime_utilities.py:42-53, delta_hedger.py:167-183,
    loss_functions.py:128-147), which index it with [:, -1].  u in the (B,U,T) layout."""
    u = _f32c(u); _require_cuda(u)
    B, U, T = u.shape
    if U != d.U:
        raise VhmmError(f"transition_probs: expected u_dim {d.U}, got {U}")
    P, keep = pack_params(params)
    out = torch.empty(B, T, d.K, d.K, device=u.device, dtype=torch.float32)
    with torch.cuda.device(u.device):
        ws = workspace(u.device, _ws_bytes(_lib.OP_TRANSITION_PROBS, d, B, T))
        check(_lib.lib().vhmm_transition_probs(ctypes.byref(d), ctypes.byref(P), _ptr(u), B, T, _ptr(out), _ptr(ws),
                                               ws.numel(), _stream(u.device)), "vhmm_transition_probs")
    return out


HMM_GRAD_WORKSPACE_CAP = 4 << 30   # bytes: the chunked transition-MLP backward takes what it gets


def hmm_param_grads(d: Dims, params, x, u, log_A, gamma, xi, w, mu_tab, logvar_tab):
    """Flat d(sum_b w[b] loglik[b]) / d(18 parameters) from the posterior pair (vhmm_hmm_param_grads)."""
    x = _f32c(x); u = _f32c(u); w = _f32c(w)
    _require_cuda(x, u, log_A, gamma, xi, w, mu_tab, logvar_tab)
    B, C, T = x.shape
    P, keep = pack_params(params)
    grads = torch.empty(param_count(d), device=x.device, dtype=torch.float32)
    with torch.cuda.device(x.device):
        need = _ws_bytes(_lib.OP_HMM_PARAM_GRADS, d, B, T)
        per_seq = T * (2 * d.TH + d.K * d.K) * 4
        ws = workspace(x.device, max(need, min(HMM_GRAD_WORKSPACE_CAP, need + per_seq * B)))
        check(_lib.lib().vhmm_hmm_param_grads(ctypes.byref(d), ctypes.byref(P), _ptr(x), _ptr(u), _ptr(_f32c(log_A)),
                                              _ptr(_f32c(gamma)), _ptr(_f32c(xi)), _ptr(w), _ptr(_f32c(mu_tab)),
                                              _ptr(_f32c(logvar_tab)), B, T, _ptr(grads), _ptr(ws), ws.numel(),
                                              _stream(x.device)), "vhmm_hmm_param_grads")
    return grads


def hmm_inputs_any(d: Dims, params, x, u, mu_tab, logvar_tab):
    """(log_A, log_b) from raw (x, u (B,U,T)): the fused front end when the dims allow, else Prior + emission kernels."""
    fused = hmm_inputs_fused(d, params, x, u, mu_tab, logvar_tab)
    if fused is not None:
        return fused
    _, log_A = prior_fwd(d, params, u)
    return log_A, emission_logprob(x, mu_tab, logvar_tab)


class HmmParamLogLikFunction(torch.autograd.Function):
    """loglik (B) = log p(x_b | u_b) of the exact HMM as a function of the 18 PARAMETERS: forward =
    front end + forward/backward recursion; backward = the beta-recursion gradient (gamma, xi: already produced by the
    forward pass) chained through the row log-softmax, the transition MLP, the emission table / decoder and log_pi to
    the parameters (vhmm_hmm_param_grads).  This is north_star's "torch.autograd.Function whose backward is the
    beta-recursion gradient kernel"; model.hmm_nll(...).backward() and train_model-style loops go through it."""

    @staticmethod
    def forward(ctx, d, x, u, lengths, *params):
        x = _f32c(x); u = _f32c(u)
        B, C, T = x.shape
        want = any(ctx.needs_input_grad)
        mu_tab, lv_tab = emission_table(d, params, T, x.device)
        log_pi = torch.log_softmax(params[6].detach().float(), dim=0)        # VQ_VAE_HMM_fixed.py:54 (K values)
        log_A, log_b = hmm_inputs_any(d, params, x, u, mu_tab, lv_tab)
        ctx.want = want
        if want:
            ll, gamma, xi = hmm_posterior(log_pi, log_A, log_b, lengths, want_xi=True)
            del log_b
            ctx.d = d
            ctx.params = [p.detach() for p in params]
            ctx.shapes = [tuple(p.shape) for p in params]
            ctx.offsets = param_offsets(d)
            ctx.save_for_backward(x, u, log_A, gamma, xi, mu_tab, lv_tab)
        else:
            ll = hmm_forward(log_pi, log_A, log_b, lengths)
        return ll

    @staticmethod
    def backward(ctx, g):
        if not ctx.want:
            raise VhmmError("hmm log-likelihood gradients were not computed (forward ran without grad)")
        x, u, log_A, gamma, xi, mu_tab, lv_tab = ctx.saved_tensors
        flat = hmm_param_grads(ctx.d, ctx.params, x, u, log_A, gamma, xi, g.to(torch.float32).contiguous(), mu_tab, lv_tab)
        outs = [flat[ctx.offsets[i]:ctx.offsets[i + 1]].view(shp) for i, shp in enumerate(ctx.shapes)]
        return (None, None, None, None, *outs)
