"""The callers either side of the ELBO path, host side.

f1  ``DeviceCorpus`` keeps the training sequences resident in HBM and builds a zero-padded batch
    with ONE gather kernel (vhmm_gather_chunks) instead of the reference's per-item slicing and 2B
    slice copies (RandomChunkDataset.__getitem__ VQ_VAE_HMM_fixed.py:20-29, collate_fn :164-179).
    ``sample_triples(..., rng=random)`` draws (sequence, length, start) with exactly the three
    ``random.randint`` calls per item the reference makes, in the same order, so that a seeded run
    produces bit-identical batches; ``sample_triples_fast`` draws the same distribution vectorised.

f2  ``FusedAdam`` / ``Trainer`` / ``train_model_fused``: the optimiser tail of train_model
    (:145-162) and src/training/trainer.py:9-43 (Adam, optional clip_grad_norm_(1.0), beta warm-up)
    with the flat gradient buffer of vhmm_elbo consumed directly by vhmm_adam_step -- no autograd
    graph, no per-tensor optimiser kernels, and no ``loss.item()`` sync per batch (the epoch loss is
    accumulated on the device and read once per epoch).
"""
from __future__ import annotations

import ctypes
import random
import random as _pyrandom
from typing import Optional, Sequence

import numpy as np
import torch

from . import _lib
from . import functional as Fn
from ._lib import VhmmError, check


# ---------------------------------------------------------------------------------------------
# f1: device-side batching
# ---------------------------------------------------------------------------------------------
def sample_triples(seq_lens: Sequence[int], n_items: int, min_len: int = 20, max_len: int = 200, rng=random):
    """(seq_idx, chunk_len, start) int64 arrays of n_items draws, consuming ``rng`` exactly like
    n_items calls of RandomChunkDataset.__getitem__ (VQ_VAE_HMM_fixed.py:21-27)."""
    idx = np.empty(n_items, np.int64)
    ln = np.empty(n_items, np.int64)
    st = np.empty(n_items, np.int64)
    n_seq = len(seq_lens)
    grb = getattr(rng, "getrandbits", None)
    if grb is not None and (rng is _pyrandom or isinstance(rng, _pyrandom.Random)):
        # Same draws as rng.randint(a, b), three call layers cheaper (the step is launch-bound: at batch 64 the 192
        # randint calls were as long as the GPU work).  CPython's randint(a, b) is a + _randbelow(b - a + 1) with
        # _randbelow(n): k = n.bit_length(); r = getrandbits(k) until r < n  (Lib/random.py, unchanged since 3.2);
        # tests/golden/collate_stream.npz pins the equivalence to batches the real reference dataset produced.
        def below(n):
            k = n.bit_length()
            r = grb(k)
            while r >= n:
                r = grb(k)
            return r
        li, ll, ls = [0] * n_items, [0] * n_items, [0] * n_items
        for k in range(n_items):
            i = below(n_seq)
            S = int(seq_lens[i])
            hi = max_len if max_len < S else S
            n = min_len + below(hi - min_len + 1)
            li[k], ll[k], ls[k] = i, n, below(S - n + 1)
        return np.array(li, np.int64), np.array(ll, np.int64), np.array(ls, np.int64)
    for k in range(n_items):
        i = rng.randint(0, n_seq - 1)
        S = int(seq_lens[i])
        n = rng.randint(min_len, min(max_len, S))
        s = rng.randint(0, S - n)
        idx[k], ln[k], st[k] = i, n, s
    return idx, ln, st


def sample_triples_fast(seq_lens: Sequence[int], n_items: int, min_len: int = 20, max_len: int = 200,
                        generator: Optional[np.random.Generator] = None):
    """Same distribution as ``sample_triples`` (uniform sequence, uniform length in
    [min_len, min(max_len, S)], uniform start), vectorised with numpy: for large batches where the
    3B Python ``random`` calls of the reference would dominate the step."""
    g = generator or np.random.default_rng()
    sl = np.asarray(seq_lens, np.int64)
    idx = g.integers(0, len(sl), size=n_items)
    S = sl[idx]
    hi = np.minimum(max_len, S)
    ln = min_len + np.floor(g.random(n_items) * (hi - min_len + 1)).astype(np.int64)
    ln = np.minimum(ln, hi)
    st = np.floor(g.random(n_items) * (S - ln + 1)).astype(np.int64)
    st = np.minimum(st, S - ln)
    return idx.astype(np.int64), ln, st


class DeviceCorpus:
    """Training sequences resident on the GPU: x_cat (C, S_tot), u_cat (U, S_tot)."""

    def __init__(self, x_sequences, u_sequences, device=None):
        dev = torch.device(device) if device is not None else torch.device("cuda", torch.cuda.current_device())
        if dev.type != "cuda":
            raise VhmmError("DeviceCorpus: a CUDA device is required (no CPU fallback)")
        xs = [torch.as_tensor(x, dtype=torch.float32) for x in x_sequences]
        us = [torch.as_tensor(u, dtype=torch.float32) for u in u_sequences]
        if len(xs) != len(us) or not xs:
            raise VhmmError("DeviceCorpus: need as many u sequences as x sequences (and at least one)")
        self.seq_lens = [int(x.shape[1]) for x in xs]
        for x, u in zip(xs, us):
            if x.shape[1] != u.shape[1]:
                raise VhmmError("DeviceCorpus: x and u sequences differ in length")
        self.C, self.U = int(xs[0].shape[0]), int(us[0].shape[0])
        self.x_cat = torch.cat(xs, dim=1).contiguous().to(dev)
        self.u_cat = torch.cat(us, dim=1).contiguous().to(dev)
        self.offsets = np.concatenate([[0], np.cumsum(self.seq_lens)]).astype(np.int64)
        self.S_tot = int(self.offsets[-1])
        self.device = dev

    def gather(self, seq_idx, chunk_len, start, T: Optional[int] = None):
        """The batch collate_fn would build from the items (x_seqs[i][:, s:s+n], u_seqs[i][:, s:s+n], n):
        x (B,C,T), u (B,U,T) zero padded on the device, lengths int64 on the CPU (:165)."""
        seq_idx = np.asarray(seq_idx, np.int64); chunk_len = np.asarray(chunk_len, np.int64); start = np.asarray(start, np.int64)
        B = int(seq_idx.shape[0])
        sl = np.asarray(self.seq_lens, np.int64)[seq_idx]
        if B == 0 or (chunk_len < 0).any() or (start < 0).any() or (start + chunk_len > sl).any():
            raise VhmmError("DeviceCorpus.gather: empty batch or chunk outside its sequence")
        lengths = torch.from_numpy(chunk_len.copy())
        Tm = int(chunk_len.max()) if T is None else int(T)
        col0 = torch.from_numpy(self.offsets[seq_idx] + start)
        both = torch.stack([col0, lengths]).to(self.device, non_blocking=True)   # one H2D copy
        x = torch.empty(B, self.C, Tm, device=self.device)
        u = torch.empty(B, self.U, Tm, device=self.device)
        with torch.cuda.device(self.device):
            check(_lib.lib().vhmm_gather_chunks(Fn._ptr(self.x_cat), Fn._ptr(self.u_cat), self.S_tot, self.C, self.U,
                                                Fn._ptr(both[0]), Fn._ptr(both[1]), B, Tm, Fn._ptr(x), Fn._ptr(u),
                                                Fn._stream(self.device)), "vhmm_gather_chunks")
        return x, u, lengths

    def sample_batch(self, batch_size: int, min_len: int = 20, max_len: int = 200, rng=random):
        """One batch exactly as DataLoader(RandomChunkDataset(...), batch_size, collate_fn=collate_fn)
        yields it for the same ``random`` state."""
        return self.gather(*sample_triples(self.seq_lens, batch_size, min_len, max_len, rng))

    def sample_batch_fast(self, batch_size: int, min_len: int = 20, max_len: int = 200, generator=None):
        return self.gather(*sample_triples_fast(self.seq_lens, batch_size, min_len, max_len, generator))


class DeviceChunkLoader:
    """Iterable with the DataLoader contract of the reference's training loop: ``len(loader)``
    batches per epoch of ``batch_size`` items out of the fixed 1000 of RandomChunkDataset.__len__
    (:17-18), each built on the device."""

    def __init__(self, corpus: DeviceCorpus, batch_size: int, min_len: int = 20, max_len: int = 200,
                 items_per_epoch: int = 1000, rng=random, fast: bool = False, generator=None):
        self.corpus, self.batch_size, self.min_len, self.max_len = corpus, batch_size, min_len, max_len
        self.items, self.rng, self.fast, self.generator = items_per_epoch, rng, fast, generator

    def __len__(self):
        return -(-self.items // self.batch_size)

    def iter_triples(self):
        """the (seq_idx, chunk_len, start) draws of one epoch, consuming the RNG exactly like __iter__"""
        left = self.items
        while left > 0:
            n = min(self.batch_size, left)
            left -= n
            if self.fast:
                yield sample_triples_fast(self.corpus.seq_lens, n, self.min_len, self.max_len, self.generator)
            else:
                yield sample_triples(self.corpus.seq_lens, n, self.min_len, self.max_len, self.rng)

    def __iter__(self):
        for triples in self.iter_triples():
            yield self.corpus.gather(*triples)


# ---------------------------------------------------------------------------------------------
# f2: fused optimiser tail
# ---------------------------------------------------------------------------------------------
class FusedAdam:
    """torch.optim.Adam(model.parameters(), lr) semantics (betas (0.9, 0.999), eps 1e-8, no weight
    decay, no amsgrad; VQ_VAE_HMM_fixed.py:146) over the flat gradient buffer of vhmm_elbo:
    one vhmm_adam_step call per update (two launches with clipping).  ``state_dict`` /
    ``load_state_dict`` use torch.optim.Adam's format so that checkpoints written by
    save_checkpoint (src/utils/data.py:47-53) interchange."""

    def __init__(self, model, lr: float = 1e-3, betas=(0.9, 0.999), eps: float = 1e-8):
        self.model, self.lr, self.betas, self.eps = model, float(lr), (float(betas[0]), float(betas[1])), float(eps)
        self.d = model._dims
        self.params = model._params()
        dev = self.params[0].device
        if dev.type != "cuda":
            raise VhmmError("FusedAdam: the model must be on a CUDA device (no CPU fallback)")
        for p in self.params:
            if p.dtype != torch.float32 or not p.is_contiguous():
                raise VhmmError("FusedAdam: parameters must be contiguous fp32")
        n = Fn.param_count(self.d)
        self.exp_avg = torch.zeros(n, device=dev)
        self.exp_avg_sq = torch.zeros(n, device=dev)
        self.scratch = torch.zeros(1, device=dev, dtype=torch.float64)
        self.grad_norm = torch.zeros(1, device=dev)
        self.step_count = 0
        self.table = _lib.Params()     # vhmm_params_mut: same layout, mutable targets
        for i, p in enumerate(self.params):
            self.table.p[i] = p.data_ptr()

    def step(self, flat_grads: torch.Tensor, grad_scale: Optional[torch.Tensor] = None, max_norm: float = 0.0):
        """One Adam update from the flat gradient buffer (state_dict order).  ``max_norm`` > 0
        applies clip_grad_norm_(parameters, max_norm) first (src/training/trainer.py:31)."""
        if flat_grads.numel() != self.exp_avg.numel() or flat_grads.dtype != torch.float32 or not flat_grads.is_cuda:
            raise VhmmError("FusedAdam.step: flat fp32 CUDA gradient buffer of param_count elements expected")
        self._check_params()
        self.step_count += 1
        dev = flat_grads.device
        with torch.cuda.device(dev):
            check(_lib.lib().vhmm_adam_step(ctypes.byref(self.d), ctypes.byref(self.table), Fn._ptr(flat_grads),
                                            Fn._ptr(self.exp_avg), Fn._ptr(self.exp_avg_sq), self.lr, self.betas[0],
                                            self.betas[1], self.eps, self.step_count, Fn._ptr(grad_scale),
                                            float(max_norm), Fn._ptr(self.grad_norm), Fn._ptr(self.scratch),
                                            Fn._stream(dev)), "vhmm_adam_step")

    def _check_params(self):
        """The pointer table was taken at construction: a later model.to(...) / .half() / param.data = ... would leave
        the kernel writing into freed memory.  Fail loudly instead."""
        for i, p in enumerate(self.params):
            if p.data_ptr() != self.table.p[i] or p.dtype != torch.float32 or not p.is_contiguous() or \
                    p.device != self.exp_avg.device:
                raise VhmmError("FusedAdam: a parameter tensor was moved, re-typed or re-assigned after the optimiser was "
                                "built; construct a new FusedAdam for the new tensors")

    def zero_grad(self, set_to_none: bool = True):   # API parity; the flat buffer is rewritten every step
        self.model.zero_grad(set_to_none=set_to_none)

    def state_dict(self):
        off = Fn.param_offsets(self.d)
        state = {}
        for i, p in enumerate(self.model.parameters()):
            j = next(k for k, q in enumerate(self.params) if q is p)
            state[i] = {"step": torch.tensor(float(self.step_count)),
                        "exp_avg": self.exp_avg[off[j]:off[j + 1]].view_as(p).clone(),
                        "exp_avg_sq": self.exp_avg_sq[off[j]:off[j + 1]].view_as(p).clone()}
        group = {"lr": self.lr, "betas": self.betas, "eps": self.eps, "weight_decay": 0, "amsgrad": False,
                 "maximize": False, "foreach": None, "capturable": False, "differentiable": False, "fused": None,
                 "decoupled_weight_decay": False, "params": list(range(len(state)))}
        return {"state": state, "param_groups": [group]}

    def load_state_dict(self, sd):
        off = Fn.param_offsets(self.d)
        g = sd["param_groups"][0]
        self.lr, self.betas, self.eps = float(g["lr"]), (float(g["betas"][0]), float(g["betas"][1])), float(g["eps"])
        for i, p in enumerate(self.model.parameters()):
            st = sd["state"].get(i)
            if st is None:
                continue
            j = next(k for k, q in enumerate(self.params) if q is p)
            self.exp_avg[off[j]:off[j + 1]].copy_(st["exp_avg"].reshape(-1))
            self.exp_avg_sq[off[j]:off[j + 1]].copy_(st["exp_avg_sq"].reshape(-1))
            self.step_count = int(float(st["step"]))


class GraphedTrainStep:
    """The small-batch training step (BASELINE.json configs[0]: batch 64, chunks of 20..200 steps) as a CUDA graph.

    One step = batch assembly from the resident corpus + normalisers + ELBO forward/backward + loss + optional
    clip_grad_norm_ + Adam = ``vhmm_train_step``, a single library call whose step-dependent scalars (KL weight,
    learning rate, step counter, bias corrections, normalisers, loss totals) live in a device ``state`` vector.  The call
    is captured ONCE per padded length T (the batch is padded to its own maximum length exactly as collate_fn does,
    VQ_VAE_HMM_fixed.py:166 -- padding further would change the result, SURVEY.md finding 4) and replayed: per step
    the host draws the (sequence, length, start) triples, issues ONE pinned H2D copy of (col0, lengths) and ONE graph
    launch.  No ATen kernel, no allocation, no sync inside an epoch; the epoch loss is read from ``state`` once.

    Semantics: torch.optim.Adam(lr, betas, eps) over the 18 tensors (no weight decay / amsgrad), optional
    clip_grad_norm_(max_norm) first, exactly like FusedAdam.step (pinned to the reference's own loop by
    tests/golden/train_adam.npz / train_clip.npz)."""

    RING = 4   # pinned staging buffers in flight

    def __init__(self, model, corpus: DeviceCorpus, batch_size: int, lr: float = 1e-3, betas=(0.9, 0.999), eps: float = 1e-8,
                 max_norm: float = 0.0, optimizer: Optional[FusedAdam] = None):
        self.model, self.corpus, self.B = model, corpus, int(batch_size)
        self.opt = optimizer if optimizer is not None else FusedAdam(model, lr=lr, betas=betas, eps=eps)
        self.max_norm = float(max_norm)
        self.d = model._dims
        dev = corpus.device
        if _lib.lib().vhmm_elbo_active_impl(ctypes.byref(self.d)) != _lib.ELBO_FUSED:
            raise VhmmError("GraphedTrainStep: needs the model dimensions of the fused ELBO path")
        if self.opt.exp_avg.device != dev:
            raise VhmmError("GraphedTrainStep: model and corpus live on different devices")
        self.dev = dev
        self.state = torch.zeros(_lib.TS_COUNT, device=dev, dtype=torch.float64)
        self._host_state = torch.zeros(_lib.TS_COUNT, dtype=torch.float64).pin_memory()
        self.both = torch.zeros(2, self.B, device=dev, dtype=torch.int64)           # [col0; lengths], graph input
        self.stage = [torch.zeros(2, self.B, dtype=torch.int64).pin_memory() for _ in range(self.RING)]
        self.stage_ev = [None] * self.RING
        self.k = 0
        self.graphs = {}
        self.grads = torch.empty(Fn.param_count(self.d), device=dev)
        self.stream = torch.cuda.Stream(device=dev)
        self._beta = None
        self._push_constants(beta=1.0)

    # -- device state -------------------------------------------------------------------------------
    def _push_constants(self, beta: float):
        """(re)write the host-controlled slots: KL weight, learning rate, Adam betas, step counter."""
        with torch.cuda.stream(self.stream):
            st = self.state
            st[_lib.TS_BETA] = float(beta)
            st[_lib.TS_LR] = self.opt.lr
            st[_lib.TS_BETA1] = self.opt.betas[0]
            st[_lib.TS_BETA2] = self.opt.betas[1]
            st[_lib.TS_STEP] = float(self.opt.step_count)
        self._beta = float(beta)

    def set_beta(self, beta: float):
        if self._beta != float(beta):
            with torch.cuda.stream(self.stream):
                self.state[_lib.TS_BETA] = float(beta)
            self._beta = float(beta)

    def reset_epoch_totals(self):
        with torch.cuda.stream(self.stream):
            self.state[_lib.TS_LOSS_SUM:_lib.TS_BATCHES + 1] = 0.0

    def epoch_mean_loss(self) -> float:
        """mean batch loss since reset_epoch_totals() -- the one device->host sync of an epoch"""
        self.stream.synchronize()
        st = self.state.cpu()
        return float(st[_lib.TS_LOSS_SUM]) / max(float(st[_lib.TS_BATCHES]), 1.0)

    def last_loss(self) -> float:
        self.stream.synchronize()
        return float(self.state[_lib.TS_LOSS])

    # -- capture / replay ---------------------------------------------------------------------------
    def _call(self, B: int, T: int, x_buf, u_buf, ws):
        c = self.corpus
        check(_lib.lib().vhmm_train_step(ctypes.byref(self.d), ctypes.byref(self.opt.table), Fn._ptr(c.x_cat), Fn._ptr(c.u_cat),
                                         c.S_tot, Fn._ptr(self.both[0]), Fn._ptr(self.both[1]), B, T, Fn._ptr(self.state),
                                         Fn._ptr(self.opt.exp_avg), Fn._ptr(self.opt.exp_avg_sq), self.opt.betas[0],
                                         self.opt.betas[1], self.opt.eps, self.max_norm, Fn._ptr(x_buf), Fn._ptr(u_buf),
                                         Fn._ptr(self.grads), Fn._ptr(ws), ws.numel(),
                                         ctypes.c_void_p(torch.cuda.current_stream(self.dev).cuda_stream)), "vhmm_train_step")

    def _graph(self, B: int, T: int):
        key = (B, T)
        g = self.graphs.get(key)
        if g is None:
            c = self.corpus
            x_buf = torch.empty(B, c.C, T, device=self.dev)
            u_buf = torch.empty(B, c.U, T, device=self.dev)
            ws = torch.empty(max(Fn._ws_bytes(_lib.OP_TRAIN_STEP, self.d, B, T), 256), dtype=torch.uint8, device=self.dev)
            graph = torch.cuda.CUDAGraph()
            self.stream.synchronize()
            with torch.cuda.device(self.dev), torch.cuda.graph(graph, stream=self.stream):
                self._call(B, T, x_buf, u_buf, ws)
            g = (graph, x_buf, u_buf, ws)
            self.graphs[key] = g
        return g

    def step(self, seq_idx, chunk_len, start, beta: float = 1.0):
        """one training step on the items (sequence, length, start) -- what DeviceCorpus.gather + loss_and_flat_grads +
        FusedAdam.step do, replayed from the graph of this batch's (B, T)"""
        c = self.corpus
        seq_idx = np.asarray(seq_idx, np.int64); chunk_len = np.asarray(chunk_len, np.int64); start = np.asarray(start, np.int64)
        B = int(seq_idx.shape[0])
        if B == 0 or B > self.B:
            raise VhmmError(f"GraphedTrainStep.step: batch of {B} items (capacity {self.B})")
        sl = np.asarray(c.seq_lens, np.int64)[seq_idx]
        if (chunk_len < 0).any() or (start < 0).any() or (start + chunk_len > sl).any():
            raise VhmmError("GraphedTrainStep.step: chunk outside its sequence")
        T = int(chunk_len.max())
        self.opt._check_params()
        self.set_beta(beta)
        graph = self._graph(B, T)[0]
        k = self.k
        self.k = (k + 1) % self.RING
        if self.stage_ev[k] is not None:
            self.stage_ev[k].synchronize()          # the copy that last read this pinned buffer has run
        hb = self.stage[k]
        hb[0, :B] = torch.from_numpy(c.offsets[seq_idx] + start)
        hb[1, :B] = torch.from_numpy(chunk_len)
        with torch.cuda.stream(self.stream):
            self.both[:, :B].copy_(hb[:, :B], non_blocking=True)
            ev = torch.cuda.Event()
            ev.record(self.stream)
            self.stage_ev[k] = ev
            graph.replay()
        self.opt.step_count += 1
        return T


class Trainer:
    """src/training/trainer.py:9-43 with the fused step: same constructor, ``train_epoch`` (returns the
    mean batch loss, clips the gradient norm to 1.0 like :31) and ``train`` (beta warm-up :38).
    A custom ``loss_fn`` runs under autograd + torch.optim.Adam as in the reference -- it must be built on
    ``model.compute_loss`` / ``model.hmm_nll`` / ``model.hmm_log_likelihood``, the differentiable entry points
    (encode / decode / forward return inference tensors without grad_fn)."""

    def __init__(self, model, lr: float = 1e-3, device: str = 'cuda', max_norm: float = 1.0):
        self.model = model.to(device)
        self.device = device
        self.max_norm = max_norm
        self.opt = FusedAdam(self.model, lr=lr)
        self._torch_opt = None
        self._graphed = None

    def train_epoch(self, dataloader, loss_fn=None, beta: float = 1.0) -> float:
        self.model.train()
        if loss_fn is None and isinstance(dataloader, DeviceChunkLoader) and \
                _lib.lib().vhmm_elbo_active_impl(ctypes.byref(self.model._dims)) == _lib.ELBO_FUSED:
            # launch-bound small-batch regime: one pinned H2D copy + one CUDA-graph launch per step
            gs = self._graphed
            if gs is None or gs.corpus is not dataloader.corpus or gs.B < dataloader.batch_size:
                gs = self._graphed = GraphedTrainStep(self.model, dataloader.corpus, dataloader.batch_size,
                                                      max_norm=self.max_norm, optimizer=self.opt)
            gs.reset_epoch_totals()
            for triples in dataloader.iter_triples():
                gs.step(*triples, beta=beta)
            return gs.epoch_mean_loss()
        total = torch.zeros((), device=self.opt.exp_avg.device, dtype=torch.float64)
        nb = 0
        for batch in dataloader:
            x, u, lengths = batch
            if loss_fn is None:
                loss, flat = self.model.loss_and_flat_grads(x, u, lengths, beta)
                self.opt.step(flat, max_norm=self.max_norm)
            else:
                if self._torch_opt is None:
                    self._torch_opt = torch.optim.Adam(self.model.parameters(), lr=self.opt.lr)
                x, u, lengths = [b.to(self.device) for b in batch]
                self._torch_opt.zero_grad()
                loss = loss_fn(self.model, x, u, lengths)
                loss.backward()
                torch.nn.utils.clip_grad_norm_(self.model.parameters(), self.max_norm)
                self._torch_opt.step()
            total += loss.detach().to(torch.float64)
            nb += 1
        return float(total) / max(nb, 1)           # the only device->host sync of the epoch

    def train(self, dataloader, num_epochs: int = 100, use_beta_warmup: bool = True) -> None:
        for ep in range(num_epochs):
            beta = min(1.0, 2.0 * (ep + 1) / num_epochs) if use_beta_warmup else 1.0
            avg_loss = self.train_epoch(dataloader, beta=beta)
            print(f"Epoch {ep+1}/{num_epochs}, Loss: {avg_loss:.4f}, Beta: {beta:.2f}")


def train_model_fused(model, dataloader, num_epochs=10, lr=1e-3):
    """train_model (VQ_VAE_HMM_fixed.py:145-162) with the fused step: Adam(lr), beta warm-up
    min(1, 2(ep+1)/E), no gradient clipping, the same line printed per epoch."""
    opt = FusedAdam(model, lr=lr)
    model.train()
    if isinstance(dataloader, DeviceChunkLoader) and \
            _lib.lib().vhmm_elbo_active_impl(ctypes.byref(model._dims)) == _lib.ELBO_FUSED:
        # launch-bound regime: every step is one pinned H2D copy + one CUDA-graph launch (GraphedTrainStep)
        gs = GraphedTrainStep(model, dataloader.corpus, dataloader.batch_size, optimizer=opt)
        for ep in range(num_epochs):
            beta = min(1.0, 2.0 * (ep + 1) / num_epochs)
            gs.reset_epoch_totals()
            for triples in dataloader.iter_triples():
                gs.step(*triples, beta=beta)
            print(f"Epoch {ep+1}/{num_epochs}, Loss: {gs.epoch_mean_loss():.4f}")
        return model
    for ep in range(num_epochs):
        beta = min(1.0, 2.0 * (ep + 1) / num_epochs)
        total = torch.zeros((), device=opt.exp_avg.device, dtype=torch.float64)
        nb = 0
        for x, u, lengths in dataloader:
            loss, flat = model.loss_and_flat_grads(x, u, lengths, beta)
            opt.step(flat)
            total += loss.to(torch.float64)
            nb += 1
        print(f"Epoch {ep+1}/{num_epochs}, Loss: {float(total)/max(nb, 1):.4f}")
    return model
