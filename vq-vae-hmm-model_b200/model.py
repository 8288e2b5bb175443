"""Host-side mirror of the reference model API (VQ_VAE_HMM_fixed.py:31-143; duplicate at
src/models/vae_hmm.py:8-126): same class names, constructor, methods, exceptions and the same 18
``state_dict`` keys / shapes / initialisation order, so reference checkpoints
(models/vae_hmm.pt) and ``torch.save / load_state_dict`` call sites work unchanged.

The nn.Conv1d / nn.Linear / nn.Embedding sub-modules are kept only as parameter containers
(default torch init, identical RNG consumption to the reference).  Their ``forward`` is never
used: every method below calls the sm_100a kernels in libvhmm.so through functional.py.
"""
from __future__ import annotations

import torch
import torch.nn as nn

from . import functional as Fn
from ._lib import VhmmError


class Encoder(nn.Module):
    """VQ_VAE_HMM_fixed.py:31-41."""

    def __init__(self, input_dim, hidden_dim, hidden_dim2, K):
        super().__init__()
        self.conv1 = nn.Conv1d(input_dim, hidden_dim, 3, padding=1)
        self.conv2 = nn.Conv1d(hidden_dim, hidden_dim2, 3, padding=1)
        self.to_logits = nn.Conv1d(hidden_dim2, K, 1)

    def forward(self, x):
        owner = self.__dict__.get("_owner")
        if owner is None:
            raise VhmmError("Encoder must be used through its VAE_HMM (model.encode / model.encoder)")
        return owner.encode(x)


class Prior(nn.Module):
    """VQ_VAE_HMM_fixed.py:43-71."""

    def __init__(self, K, u_dim=None, trans_hidden=128):
        super().__init__()
        self.K = K
        self.u_dim = u_dim
        self.log_prior = nn.Parameter(torch.zeros(K))
        if u_dim is None:
            raise ValueError('Stationary transitions not implemented')
        self.transition_net = nn.Sequential(
            nn.Linear(u_dim, trans_hidden),
            nn.ReLU(),
            nn.Linear(trans_hidden, K * K),
        )

    def forward(self, u=None):
        if u is None:
            raise ValueError('u required for non-stationary transitions')
        owner = self.__dict__.get("_owner")
        if owner is None:
            raise VhmmError("Prior must be used through its VAE_HMM (model.prior)")
        # layout rule of VQ_VAE_HMM_fixed.py:64-65: (B,U,T) is permuted only when dim 1 == u_dim;
        # the kernels consume the (B,U,T) layout, so a (B,T,U) input is permuted the other way.
        if u.dim() != 3:
            raise VhmmError(f"prior: expected a 3-d u, got {tuple(u.shape)}")
        if u.shape[1] != self.u_dim:
            if u.shape[2] != self.u_dim:
                raise VhmmError(f"prior: no axis of u{tuple(u.shape)} equals u_dim={self.u_dim}")
            u = u.permute(0, 2, 1)
        return Fn.prior_fwd(owner._dims, owner._params(), u)


class Decoder(nn.Module):
    """VQ_VAE_HMM_fixed.py:73-90."""

    def __init__(self, K, latent_dim, hidden_dim, output_dim):
        super().__init__()
        self.embeddings = nn.Embedding(K, latent_dim)
        self.conv1 = nn.Conv1d(latent_dim, hidden_dim, 3, padding=1)
        self.conv2 = nn.Conv1d(hidden_dim, hidden_dim, 3, padding=1)
        self.to_params = nn.Conv1d(hidden_dim, output_dim * 2, 1)

    def forward(self, q):
        owner = self.__dict__.get("_owner")
        if owner is None:
            raise VhmmError("Decoder must be used through its VAE_HMM (model.decode / model.decoder)")
        return owner.decode(q)


class VAE_HMM(nn.Module):
    """Drop-in for VQ_VAE_HMM_fixed.VAE_HMM (:92-143) on B200.

    ``compute_loss`` is differentiable w.r.t. the 18 parameters (fused forward+backward kernel
    behind a torch.autograd.Function).  ``encode/decode/prior/forward`` are inference ops and
    return tensors without grad_fn.  Additions (absent from the reference):
    ``hmm_inputs``, ``hmm_log_likelihood``, ``hmm_nll`` (differentiable w.r.t. the 18 parameters),
    ``get_regime_probs`` (the method api.py:18 calls), ``transition_probs``, ``viterbi``, ``decode_regimes``,
    ``argmax_regimes``.
    """

    def __init__(self, input_dim, hidden_dim, K, hidden_dim2, u_dim=None, trans_hidden=128):
        super().__init__()
        self.K = K
        self.encoder = Encoder(input_dim, hidden_dim, hidden_dim2, K)
        self.prior = Prior(K, u_dim, trans_hidden)
        self.decoder = Decoder(K, hidden_dim, hidden_dim, input_dim)
        self._dims = Fn.make_dims(input_dim, hidden_dim, K, hidden_dim2, u_dim, trans_hidden)
        # back-reference for the sub-modules' forward(); object.__setattr__ keeps it out of
        # nn.Module's registries (no parameter duplication, state_dict unchanged)
        for sub in (self.encoder, self.prior, self.decoder):
            object.__setattr__(sub, "_owner", self)
        self.shard_group = None  # set by parallel helpers: ELBO normalisers / grads become global

    # -- plumbing -------------------------------------------------------------------------
    def _params(self):
        sd = dict(self.named_parameters())
        return [sd[k] for k in Fn.PARAM_ORDER]

    # -- reference API --------------------------------------------------------------------
    def encode(self, x):
        logits, _ = Fn.encoder_fwd(self._dims, self._params(), x)
        return logits

    def decode(self, q):
        return Fn.decoder_fwd(self._dims, self._params(), q)

    def compute_loss(self, x, u=None, lengths=None, beta=1.0, global_lengths=None):
        """VQ_VAE_HMM_fixed.py:106-137.  `global_lengths` (sharded models only, optional): the lengths of the WHOLE batch over
        all ranks -- the global normalisers then come from the host and the step has one collective instead of two."""
        B, C, T = x.shape
        if lengths is None:
            raise ValueError('lengths required')
        if u is None:
            raise ValueError('u required for non-stationary transitions')
        if u.dim() == 3 and u.shape[1] != self._dims.U and u.shape[2] == self._dims.U:
            u = u.permute(0, 2, 1)
        # x, u may live in (pinned) host memory while the model is on a GPU: the batch is then streamed
        # to the device chunk by chunk underneath the kernels (functional.elbo_raw_streamed)
        dev = x.device if x.is_cuda else self.prior.log_prior.device
        lengths_dev = lengths.to(device=dev, dtype=torch.int64)
        norms = Fn.elbo_norms(lengths_dev, B, C, T)
        if self.shard_group is not None:
            from . import parallel
            ng = None if global_lengths is None else parallel.global_norms_from_lengths(global_lengths, C, T, dev)
            return parallel.sharded_elbo(self, x, u, lengths_dev, float(beta), norms, self.shard_group, ng)
        loss, _ = Fn.ElboFunction.apply(self._dims, x, u, lengths_dev, float(beta), norms, *self._params())
        return loss

    def loss_and_flat_grads(self, x, u, lengths, beta=1.0, global_lengths=None):
        """compute_loss and its gradient without an autograd graph: (loss 0-d fp32 on the device,
        flat fp32 gradient buffer in state_dict order).  This is what vhmm_elbo produces in one call;
        training.FusedAdam consumes the buffer directly."""
        B, C, T = x.shape
        if lengths is None:
            raise ValueError('lengths required')
        if u is None:
            raise ValueError('u required for non-stationary transitions')
        if u.dim() == 3 and u.shape[1] != self._dims.U and u.shape[2] == self._dims.U:
            u = u.permute(0, 2, 1)
        dev = x.device if x.is_cuda else self.prior.log_prior.device
        lengths_dev = lengths.to(device=dev, dtype=torch.int64)
        norms = Fn.elbo_norms(lengths_dev, B, C, T)
        with torch.no_grad():
            if self.shard_group is not None:
                from . import parallel
                ng = None if global_lengths is None else parallel.global_norms_from_lengths(global_lengths, C, T, dev)
                return parallel.sharded_loss_and_flat_grads(self, x, u, lengths_dev, float(beta), norms, self.shard_group, ng)
            sums, grads = Fn.elbo_raw(self._dims, self._params(), x, u, lengths_dev, norms, float(beta), True)
            return Fn.loss_from_sums(sums, norms, float(beta)), grads

    def loss_terms(self, x, u, lengths):
        """(recon_loss, prior_loss, entropy) as normalised by the reference (:120, :131, :135)."""
        B, C, T = x.shape
        lengths_dev = lengths.to(device=x.device, dtype=torch.int64)
        norms = Fn.elbo_norms(lengths_dev, B, C, T)
        sums, _ = Fn.elbo_raw(self._dims, self._params(), x, u, lengths_dev, norms, 1.0, False)
        return sums[0] / norms[0], -sums[1] / norms[1], sums[2] / norms[1]

    def forward(self, x):
        (mu, logvar), q, _ = Fn.forward_fwd(self._dims, self._params(), x)   # one fused tile kernel for the base dims
        return (mu, logvar), q

    # -- regime decoding used all over the reference (backtesting.py:155 etc.) -------------
    def regime_probs_meanfield(self, x):
        _, q = Fn.encoder_fwd(self._dims, self._params(), x, want_q=True, want_logits=False)
        return q

    def argmax_regimes(self, x):
        return Fn.argmax_regimes(self.regime_probs_meanfield(x))

    # -- exact-HMM extensions ----------------------------------------
    def hmm_inputs(self, x, u):
        """(log_pi (K), log_A (B,T,K,K), log_b (B,T,K)) with the per-state emission table of B.1."""
        B, C, T = x.shape
        mu_tab, lv_tab = Fn.emission_table(self._dims, self._params(), T, x.device)
        if u.dim() == 3 and u.shape[1] == self._dims.U:
            fused = Fn.hmm_inputs_fused(self._dims, self._params(), x, u, mu_tab, lv_tab)   # one FFMA kernel
            if fused is not None:
                log_pi = torch.log_softmax(self.prior.log_prior.detach().float(), dim=0)    # :54 (K values)
                return log_pi, fused[0], fused[1]
        log_pi, log_A = self.prior(u)
        log_b = Fn.emission_logprob(x, mu_tab, lv_tab)
        return log_pi, log_A, log_b

    def _u_layout(self, u):
        """(B,U,T) view of u under the permute rule of VQ_VAE_HMM_fixed.py:64-65"""
        if u is None:
            raise ValueError('u required for non-stationary transitions')
        if u.dim() != 3:
            raise VhmmError(f"expected a 3-d u, got {tuple(u.shape)}")
        if u.shape[1] != self._dims.U:
            if u.shape[2] != self._dims.U:
                raise VhmmError(f"no axis of u{tuple(u.shape)} equals u_dim={self._dims.U}")
            u = u.permute(0, 2, 1)
        return u

    def hmm_log_likelihood(self, x, u, lengths=None):
        """log p(x | u) per sequence (B.2).  With autograd enabled and parameters that require grad the result is
        differentiable w.r.t. the 18 parameters: backward = the beta-recursion gradient kernel chained to the
        parameters (functional.HmmParamLogLikFunction).  Without grad: one fused kernel from raw (x, u) when the dims
        allow (nothing materialised), else front end + forward recursion."""
        u = self._u_layout(u)
        if torch.is_grad_enabled() and any(p.requires_grad for p in self.parameters()) and x.numel() > 0:
            if lengths is not None:
                lengths = lengths.to(device=x.device, dtype=torch.int64)
            return Fn.HmmParamLogLikFunction.apply(self._dims, x, u, lengths, *self._params())
        # FEW LONG sequences: the fused kernel walks a sequence with one warp, serially in time -- below ~2,400 sequences it
        # cannot fill the machine, and at T = 65536 a warp needs milliseconds.  Front end (parallel over every (b, t)) + the
        # time-parallel scan instead (B = 64, T = 65536: tools/loglik_long_time.py).
        few_long = x.numel() > 0 and x.shape[2] >= 1024 and x.shape[0] < 2368
        if x.numel() > 0 and not few_long:
            mu_tab, lv_tab = Fn.emission_table(self._dims, self._params(), x.shape[2], x.device)
            log_pi = torch.log_softmax(self.prior.log_prior.detach().float(), dim=0)
            ll = Fn.hmm_loglik_fused(self._dims, self._params(), log_pi, x, u, lengths, mu_tab, lv_tab)
            if ll is not None:
                return ll
        log_pi, log_A, log_b = self.hmm_inputs(x, u)
        if few_long:
            return Fn.hmm_scan(log_pi, log_A, log_b, lengths)
        return Fn.hmm_forward(log_pi, log_A, log_b, lengths)

    def hmm_nll(self, x, u, lengths=None):
        """-mean_b loglik[b] (B.2): the exact-HMM training loss.  ``model.hmm_nll(x, u, lengths).backward()`` leaves
        the gradients of all 18 parameters in ``.grad`` (the six encoder tensors get zeros: the exact likelihood
        does not use the encoder), so the reference's loop (VQ_VAE_HMM_fixed.py:153-157) can train on it."""
        return -self.hmm_log_likelihood(x, u, lengths).mean()

    def transition_probs(self, u):
        """exp(log_A) as (B,T,K,K) -- the transition tensor regime_utilities.py:42-53, delta_hedger.py:167-183 and
        loss_functions.py:128-147 take next to the (B,K,T) regime probabilities (they read [:, -1])."""
        return Fn.transition_probs(self._dims, self._params(), self._u_layout(u))

    def get_regime_probs(self, x, u, lengths=None):
        """Smoothed posteriors gamma as (B,K,T) -- the layout every reference consumer of
        ``softmax(encode(x))`` expects (api.py:18 calls this method; it exists nowhere upstream)."""
        log_pi, log_A, log_b = self.hmm_inputs(x, u)
        _, gamma, _ = Fn.hmm_posterior(log_pi, log_A, log_b, lengths, want_xi=False)
        return gamma.permute(0, 2, 1)

    def viterbi(self, x, u, lengths=None):
        """uint8 MAP state path (B,T) (255 beyond each length) and its score (B.4)."""
        log_pi, log_A, log_b = self.hmm_inputs(x, u)
        return Fn.viterbi(log_pi, log_A, log_b, lengths)

    def decode_regimes(self, x, u, lengths=None):
        """(path uint8 (B,T), gamma (B,T,K), loglik (B)) from one pass over the inputs."""
        log_pi, log_A, log_b = self.hmm_inputs(x, u)
        return Fn.hmm_decode(log_pi, log_A, log_b, lengths)
