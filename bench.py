#!/usr/bin/env python3
"""bench.py -- sequence-steps/s of the VAE_HMM ELBO forward+backward (BASELINE.json metric).

    python bench.py --gpus N --steps K --warmup W            # this repo (CUDA, sm_100a)
    python bench.py --impl reference --steps K --warmup W    # the reference's CPU PyTorch path
                                                             # (oracle port) on the host cores
Under torchrun (N>1) every rank runs the same per-GPU workload (weak scaling: the batch is
sharded, B_local sequences per GPU) and the step contains the one NCCL all-reduce of
[gradients || loss sums]; rank 0 prints ONE JSON line.

Workload = BASELINE.json configs[1]: VAE_HMM(input_dim=5, hidden_dim=64, K=3, hidden_dim2=32,
u_dim=4), synthetic B=65536 T=512 fp32, one step = compute_loss(...).backward() over the batch.
`value` is measured with inputs resident in HBM; `e2e` through the public API from pinned HOST
buffers (H2D of x,u,lengths and D2H of the loss inside the timed region).
"""
import argparse
import json
import os
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

MODEL = dict(C=5, H=64, K=3, H2=32, U=4, TH=128)
FLOP_PER_STEP = 205740          # algorithmic fwd+bwd FLOPs per sequence-step (base)
BYTES_PER_STEP = 72             # x,u read once forward and once backward
TRAFFIC_FILE = "r2_decoder_traffic.json"   # latest committed ncu --set full capture of the decoder tile kernel
DEC_FLOP_PER_STEP = 3 * 50816   # decoder forward 50,816 FLOP/step; fwd+bwd = 3x
FP32_PEAK_TFLOPS = 148 * 128 * 2 * 1.965e9 / 1e12   # FFMA peak at max SM clock (derived)


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return dict(hbm=d["hbm_gbs"], bf16=d.get("bf16_tflops_sustained", d["bf16_tflops"]), src="measured")
    return dict(hbm=6650.0, bf16=1400.0, src="fallback")


class ClockSampler:
    """samples SM clock / throttle reasons of one GPU during the timed region"""

    def __init__(self, index, interval=0.05):
        self.index, self.samples, self.reasons, self.max_mhz = index, [], set(), None
        self.interval = interval
        self._stop = threading.Event()
        self._thr = None
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
        except Exception:
            self.nv = None

    def _run(self):
        nv = self.nv
        names = {"hw_slowdown": 0x8, "sw_power_cap": 0x4, "hw_thermal_slowdown": 0x40,
                 "sw_thermal_slowdown": 0x20, "hw_power_brake": 0x80}
        while not self._stop.is_set():
            try:
                self.samples.append(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM))
                try:
                    r = nv.nvmlDeviceGetCurrentClocksEventReasons(self.h)
                except Exception:
                    r = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                for k, bit in names.items():
                    if r & bit:
                        self.reasons.add(k)
            except Exception:
                pass
            time.sleep(self.interval)

    def start(self):
        if self.nv is not None:
            self._thr = threading.Thread(target=self._run, daemon=True)
            self._thr.start()

    def stop(self):
        self._stop.set()
        if self._thr is not None:
            self._thr.join()
        s = sorted(self.samples)
        return {"sm_mhz": (s[len(s) // 2] if s else None), "sm_max_mhz": self.max_mhz,
                "reasons": sorted(self.reasons), "samples": len(s)}


def cpu_reference_step_rate(B_cpu, T, steps, warmup, threads=None):
    """The reference's CPU PyTorch path timed on the host cores: compute_loss + backward on a bounded sample of
    the workload.  kind "reference": the UNMODIFIED reference module (oracle/_ref/VQ_VAE_HMM_fixed.py, placed
    there by oracle/build_ref.py; it travels to the GPU box like the built library); kind "port": the pinned
    op-for-op restatement oracle/elbo.py when the copy is absent.  Returns (steps/s, s/step, threads, kind)."""
    import torch
    from oracle import elbo as O
    from oracle import build_ref
    n = threads or os.cpu_count() or 1
    torch.set_num_threads(n)
    x, u, L = O.synth_batch(B_cpu, MODEL["C"], MODEL["U"], T, seed=1234, ragged=False)
    R = build_ref.load()
    if R is not None:
        torch.manual_seed(0)
        m = R.VAE_HMM(MODEL["C"], MODEL["H"], MODEL["K"], MODEL["H2"], u_dim=MODEL["U"])   # stays on the CPU
        m.train()

        def step():
            m.zero_grad(set_to_none=True)
            loss = m.compute_loss(x, u, L, 1.0)      # VQ_VAE_HMM_fixed.py:106-137
            loss.backward()                          # :156
            return loss
        kind = "reference"
    else:
        p = O.init_params(MODEL["C"], MODEL["H"], MODEL["K"], MODEL["H2"], MODEL["U"], MODEL["TH"], seed=0)

        def step():
            return O.loss_and_grads(p, x, u, L, MODEL["U"], 1.0)
        kind = "port"
    for _ in range(warmup):
        step()
    t0 = time.perf_counter()
    for _ in range(steps):
        step()
    dt = (time.perf_counter() - t0) / steps
    return B_cpu * T / dt, dt, n, kind


CPU_KIND_TEXT = {"reference": "the unmodified reference module (oracle/_ref/VQ_VAE_HMM_fixed.py) on the CPU, torch fp32",
                 "port": "oracle port (oracle/elbo.py, pinned bit for bit to the reference) on the CPU, torch fp32"}


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    B_cpu, T = args.cpu_batch, args.T
    rate, dt, n, kind = cpu_reference_step_rate(B_cpu, T, args.steps, args.warmup)
    sample = f"B={B_cpu} of {args.B} sequences, T={T}, compute_loss+backward, {CPU_KIND_TEXT[kind]}"
    line = {
        "impl": "reference", "metric": "sequence-steps/sec for ELBO fwd+bwd", "value": rate,
        "unit": "sequence-steps/s", "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": dt * 1e3, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f32", "data": "synthetic",
        "config": {"workload": f"cfg2: VAE_HMM(5,64,K=3,32,u_dim=4) ELBO fwd+bwd B={args.B} T={T} (CPU sample B={B_cpu})"},
        "cpu_baseline": {"value": rate, "unit": "sequence-steps/s", "cores": n, "kind": kind, "sample": sample},
        "e2e": {"value": rate, "unit": "sequence-steps/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


# ---------------------------------------------------------------------------------------------------------
# stage lines: the HBM-bound exact-HMM kernels (and the front ends that feed them) on BASELINE.json configs
# [2] (B=2^20, T=256, K=3), [3] (B=64, T=65536) and a K=32 case of [4], each timed with CUDA events on the
# launch stream with its own NVML clock record.  Algorithmic bytes per step: SURVEY.md 8(d).
# ---------------------------------------------------------------------------------------------------------
def timed_stage(torch, fn, local, min_ms=250.0, warmup=3, max_iters=400):
    for _ in range(warmup):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(); fn(); e1.record()
    torch.cuda.synchronize()
    one = max(e0.elapsed_time(e1), 1e-3)
    iters = int(min(max_iters, max(5, min_ms / one)))
    sampler = ClockSampler(local, interval=0.01)
    sampler.start()
    e0.record()
    for _ in range(iters):
        fn()
    e1.record()
    torch.cuda.synchronize()
    clocks = sampler.stop()
    return e0.elapsed_time(e1) / iters, iters, clocks


def stage_line(name, cfg, ms, iters, clocks, steps, bytes_per_step, pk, flop_per_step=None, note=None):
    gbs = bytes_per_step * steps / (ms * 1e-3) / 1e9
    d = {"stage": name, "config": cfg, "ms": ms, "iters": iters, "steps": steps, "steps_per_s": steps / (ms * 1e-3),
         "bytes_per_step": bytes_per_step, "bound": "hbm", "achieved_gbs": gbs, "peak_gbs": pk["hbm"],
         "frac": gbs / pk["hbm"], "clocks": clocks}
    if flop_per_step:
        tf = flop_per_step * steps / (ms * 1e-3) / 1e12
        d.update({"bound": "fp32-ffma", "flop_per_step": flop_per_step, "achieved_tflops": tf,
                  "frac_of_fp32_ffma_peak": tf / FP32_PEAK_TFLOPS, "frac_hbm": d["frac"]})
        d["frac"] = tf / FP32_PEAK_TFLOPS
    if note:
        d["note"] = note
    return d


def synth_hmm_inputs(torch, dev, B, T, K, seed):
    """row-normalised log_A (B,T,K,K), Gaussian-like log_b (B,T,K), log_pi -- generated in slabs"""
    g = torch.Generator(device=dev).manual_seed(seed)
    log_pi = torch.log_softmax(torch.randn(K, device=dev, generator=g), 0)
    log_A = torch.empty(B, T, K, K, device=dev)
    slab = max(1, (1 << 26) // (T * K * K))
    for b0 in range(0, B, slab):
        n = min(slab, B - b0)
        log_A[b0:b0 + n] = torch.log_softmax(torch.randn(n, T, K, K, device=dev, generator=g), -1)
    log_b = -0.5 * torch.randn(B, T, K, device=dev, generator=g) ** 2 - 1.0
    return log_pi, log_A, log_b


def run_stages(torch, V, Fn, model, dev, local, pk, quick=False):
    """rank 0, single GPU.  Returns the list of stage lines."""
    out = []
    mn = 60.0 if quick else 250.0
    with torch.no_grad():
        # ---- cfg 2 shape, inference: the fused forward-only tile kernel (encode / forward of the reference API)
        B, T = 65536, 512
        cfg = f"cfg2 inference: B={B} T={T}, VAE_HMM.forward / encode through the public API"
        g = torch.Generator(device=dev).manual_seed(99)
        x = torch.randn(B, MODEL["C"], T, device=dev, generator=g)
        ms, it, ck = timed_stage(torch, lambda: model(x), local, mn)
        out.append(stage_line("forward (x -> q, mu, logvar: ONE fused tile kernel, conv2 layers on tcgen05, no activation in HBM)",
                              cfg, ms, it, ck, B * T, 20 + 12 + 40, pk, flop_per_step=68580,
                              note="68,580 algorithmic FLOP per step (SURVEY 8d fwd); compute-bound: judged against the FP32 FFMA peak"))
        ms, it, ck = timed_stage(torch, lambda: model.regime_probs_meanfield(x), local, mn)
        out.append(stage_line("encode + softmax (x -> q: the regime-probability call of every reference consumer)",
                              cfg, ms, it, ck, B * T, 20 + 12, pk, flop_per_step=14400))
        del x
        torch.cuda.empty_cache()
        # ---- cfg 3: B = 2^20, T = 256, K = 3; inputs produced by the pipeline from seeded x, u
        B, T, K = (1 << 20), 256, 3
        cfg = f"cfg3: B={B} T={T} K={K}, lengths=T, log_A/log_b from the model's front end on seeded x,u"
        g = torch.Generator(device=dev).manual_seed(4321)
        x = torch.randn(B, MODEL["C"], T, device=dev, generator=g)
        u = torch.randn(B, MODEL["U"], T, device=dev, generator=g)
        L = torch.full((B,), T, dtype=torch.int64, device=dev)
        steps = B * T
        ms, it, ck = timed_stage(torch, lambda: model.hmm_inputs(x, u), local, mn)
        out.append(stage_line("hmm_inputs (x,u -> log_A, log_b: transition MLP + row log-softmax + emission log-prob, one kernel)",
                              cfg, ms, it, ck, steps, 36 + 48, pk, flop_per_step=3328 + 90,
                              note="compute-bound front end: 3,328 FLOP of the 4->128->9 MLP per 84 B"))
        ms, it, ck = timed_stage(torch, lambda: model.hmm_log_likelihood(x, u, L), local, mn)
        out.append(stage_line("hmm_loglik_fused (raw x,u -> log p(x|u): softmax + emission + forward recursion, one kernel)",
                              cfg, ms, it, ck, steps, 36, pk, flop_per_step=3328 + 90,
                              note="north-star kernel #2; nothing materialised, HBM traffic = x,u only"))
        ms, it, ck = timed_stage(torch, lambda: model.decode_regimes(x, u, L), local, mn)
        out.append(stage_line("decode_regimes from raw x,u (front end + Viterbi + gamma)", cfg, ms, it, ck, steps, 36 + 13, pk,
                              flop_per_step=3328 + 90,
                              note="three kernels: the FP32-bound front end (22.6 of the ~29.5 ms, materialises log_A / log_b), the "
                                   "Viterbi sweep and the one-pass posterior (HBM-bound, see hmm_decode below)"))
        log_pi, log_A, log_b = model.hmm_inputs(x, u)
        del x, u
        inb = 4 * (K * K + K)
        for name, bps, fn in (
                ("hmm_forward (log-likelihood)", inb, lambda: Fn.hmm_forward(log_pi, log_A, log_b, L)),
                ("viterbi (uint8 path)", inb + 1, lambda: Fn.viterbi(log_pi, log_A, log_b, L)),
                ("hmm_posterior (gamma)", inb + 4 * K, lambda: Fn.hmm_posterior(log_pi, log_A, log_b, L, want_xi=False)),
                ("hmm_posterior (gamma + xi = beta-recursion gradient)", 2 * inb, lambda: Fn.hmm_posterior(log_pi, log_A, log_b, L, want_xi=True)),
                ("hmm_decode (Viterbi path + gamma)", inb + 4 * K + 1, lambda: Fn.hmm_decode(log_pi, log_A, log_b, L))):
            ms, it, ck = timed_stage(torch, fn, local, mn)
            out.append(stage_line(name, cfg, ms, it, ck, steps, bps, pk))
        del log_A, log_b
        Fn.release_workspaces()
        torch.cuda.empty_cache()
        # ---- cfg 4: B = 64, T = 65536 (long sequences, small batch): log-semiring scan over T
        B, T = 64, 65536
        cfg = f"cfg4: B={B} T={T} K={K} (201 MB job: launch/latency-bound)"
        log_pi, log_A, log_b = synth_hmm_inputs(torch, dev, B, T, K, 11)
        L = torch.full((B,), T, dtype=torch.int64, device=dev)
        ms, it, ck = timed_stage(torch, lambda: Fn.hmm_scan(log_pi, log_A, log_b, L), local, mn, max_iters=4000)
        out.append(stage_line("hmm_scan (log-likelihood)", cfg, ms, it, ck, B * T, inb, pk,
                              note="201 MB of inputs fit the 126 MB L2 only partly; repeated launches re-read from HBM"))
        ms, it, ck = timed_stage(torch, lambda: Fn.hmm_scan(log_pi, log_A, log_b, L, want_alpha=True), local, mn, max_iters=2000)
        out.append(stage_line("hmm_scan (all-prefix alpha)", cfg, ms, it, ck, B * T, inb + 4 * K, pk,
                              note="segment products + carry scan + alpha sweep per segment: the inputs are read twice"))
        ms, it, ck = timed_stage(torch, lambda: Fn.hmm_posterior(log_pi, log_A, log_b, L, want_xi=False), local, mn, max_iters=2000)
        out.append(stage_line("hmm_posterior (gamma) of long sequences", cfg, ms, it, ck, B * T, inb + 4 * K, pk,
                              note="time-parallel: segment products, forward + backward carry scans, one-pass posterior kernel on "
                                   "256-step rows (two reads of the inputs); the thread-per-sequence sweeps took 94 ms here"))
        ms, it, ck = timed_stage(torch, lambda: Fn.viterbi(log_pi, log_A, log_b, L), local, mn, max_iters=500)
        out.append(stage_line("viterbi of long sequences", cfg, ms, it, ck, B * T, inb + 1, pk,
                              note="bit-exact max-plus is serial in time: one warp per sequence, chain of add / 3-input max / add "
                                   "per step, back-pointers and back-track off the chain (issue-bound at ~45 cycles per step, "
                                   "not HBM-bound); the thread-per-sequence sweep took 31.6 ms here"))
        del log_A, log_b
        # ---- cfg 5 recursion shapes: K = 32 (warp-per-sequence kernels), T = 1024
        B, T, K = 8192, 1024, 32
        cfg = f"cfg5 recursion: B={B} T={T} K={K} (a quarter of the cfg-5 batch: the materialised log_A is 4,224 B/step = 35 GB here)"
        log_pi, log_A, log_b = synth_hmm_inputs(torch, dev, B, T, K, 12)
        L = torch.full((B,), T, dtype=torch.int64, device=dev)
        inb = 4 * (K * K + K)
        for name, bps, fn in (
                ("hmm_forward K=32", inb, lambda: Fn.hmm_forward(log_pi, log_A, log_b, L)),
                ("viterbi K=32", inb + 1, lambda: Fn.viterbi(log_pi, log_A, log_b, L)),
                ("hmm_posterior K=32 (gamma)", inb + 4 * K, lambda: Fn.hmm_posterior(log_pi, log_A, log_b, L, want_xi=False))):
            ms, it, ck = timed_stage(torch, fn, local, mn)
            out.append(stage_line(name, cfg, ms, it, ck, B * T, bps, pk))
        del log_A, log_b
        Fn.release_workspaces()
        torch.cuda.empty_cache()
        # ---- K = 32 on FEW LONG sequences: segment products + carries (K^3 work per step over B*T/128 warps)
        B, T = 4, 16384
        cfg = f"K=32 few long: B={B} T={T} K={K} (277 MB of inputs)"
        log_pi, log_A, log_b = synth_hmm_inputs(torch, dev, B, T, K, 13)
        L = torch.full((B,), T, dtype=torch.int64, device=dev)
        for name, bps, fn, note in (
                ("hmm_scan K=32 of long sequences (log-likelihood)", inb, lambda: Fn.hmm_scan(log_pi, log_A, log_b, L),
                 "K x K segment products (one warp per 128 steps) + per-sequence carry; the warp-per-sequence forward took 18.8 ms"),
                ("hmm_posterior K=32 of long sequences (gamma)", inb + 4 * K, lambda: Fn.hmm_posterior(log_pi, log_A, log_b, L, want_xi=False),
                 "segment products, forward and backward carries, then the forward / backward kernels on 128-step rows (three reads of the inputs)"),
                ("viterbi K=32 of long sequences", inb + 1, lambda: Fn.viterbi(log_pi, log_A, log_b, L),
                 "serial in time (bit-exact max-plus): one warp per sequence, bulk-copy input ring, (value, index) tree; issue-bound")):
            ms, it, ck = timed_stage(torch, fn, local, mn, max_iters=500)
            out.append(stage_line(name, cfg, ms, it, ck, B * T, bps, pk, note=note))
        del log_A, log_b
        Fn.release_workspaces()
        torch.cuda.empty_cache()
    return out


def run_train_cfg1(torch, V, dev):
    """BASELINE.json configs[0] as a TRAINING loop (train_model, VQ_VAE_HMM_fixed.py:145-162): batch 64, chunks of
    20..200 steps of 40 base sequences, 1000 items per epoch (16 batches).  Wall clock per batch over whole epochs
    (host sampling included, one sync per epoch): eager fused step vs the CUDA-graph replay (GraphedTrainStep)."""
    import random
    gcpu = torch.Generator().manual_seed(1234)
    xs = list(torch.randn(40, MODEL["C"], 400, generator=gcpu))
    us = list(torch.randn(40, MODEL["U"], 400, generator=gcpu))
    corpus = V.DeviceCorpus(xs, us, device=dev)
    res = {}
    for mode in ("eager", "graph"):
        torch.manual_seed(0)
        m = V.VAE_HMM(MODEL["C"], MODEL["H"], MODEL["K"], MODEL["H2"], u_dim=MODEL["U"]).to(dev)
        opt = V.FusedAdam(m, lr=1e-3)
        loader = V.DeviceChunkLoader(corpus, 64)
        random.seed(0)
        gs = V.training.GraphedTrainStep(m, corpus, 64, optimizer=opt) if mode == "graph" else None

        def epoch():
            steps = 0
            if gs is None:
                tot = torch.zeros((), device=dev, dtype=torch.float64)
                for xb, ub, L in loader:
                    loss, flat = m.loss_and_flat_grads(xb, ub, L, 1.0)
                    opt.step(flat)
                    tot += loss
                    steps += int(L.sum())
                float(tot)
            else:
                gs.reset_epoch_totals()
                for tr in loader.iter_triples():
                    gs.step(*tr, beta=1.0)
                    steps += int(tr[1].sum())
                gs.epoch_mean_loss()
            return steps
        for _ in range(12 if mode == "graph" else 2):     # graph: the first epochs capture one graph per distinct padded length
            epoch()
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        steps, n_ep, per_epoch = 0, 10, []
        for _ in range(n_ep):
            te = time.perf_counter()
            steps += epoch()                                  # (ends with a device->host read of the epoch loss: synchronised)
            per_epoch.append(time.perf_counter() - te)
        torch.cuda.synchronize()
        dt = (time.perf_counter() - t0) / n_ep
        res[mode] = {"ms_per_epoch": dt * 1e3, "ms_per_batch": dt * 1e3 / len(loader), "sequence_steps_per_s": steps / n_ep / dt,
                     # the loop is host-bound (a Python call per batch): the quietest epoch shows the step without host jitter
                     "ms_per_batch_best_epoch": min(per_epoch) * 1e3 / len(loader)}
        if gs is not None:
            res[mode]["graphs_captured"] = len(gs.graphs)
    res["workload"] = "cfg1 training: batch 64, chunks 20..200 of 40x400-step sequences, 16 batches/epoch, Adam; wall clock incl. host sampling"
    return res


def run_cfg5(torch, V, Fn, dev, local, pk):
    """BASELINE.json configs[4]: scaled model K=32, hidden_dim=256 (hidden_dim2=128, u_dim=16), B=32768, T=1024, ELBO fwd+bwd on
    the layer-wise path with the tcgen05 layer kernels (csrc/conv_tc5.cuh, prior_tc5.cuh) at the FULL batch; the FP32-FFMA layer
    kernels of round 1 on a B=4096 slice (the layer-wise path processes the batch in chunks of 2048 sequences anyway; at
    0.7 s per slice the full batch would take a minute)."""
    T = 1024
    FLOP = 3872256          # algorithmic fwd+bwd FLOP per step of the scaled (un-folded) model
    torch.manual_seed(0)
    m = V.VAE_HMM(5, 256, 32, 128, u_dim=16).to(dev)
    res = {}
    for conv, B in (("tc5", 32768), ("ffma", 4096)):
        g = torch.Generator(device=dev).manual_seed(1)
        x = torch.randn(B, 5, T, device=dev, generator=g)
        u = torch.randn(B, 16, T, device=dev, generator=g)
        L = torch.full((B,), T, dtype=torch.int64)
        Fn.conv_select(conv)
        try:
            def step():
                m.zero_grad(set_to_none=True)
                loss = m.compute_loss(x, u, L, 1.0)
                loss.backward()
                return loss
            ms, it, ck = timed_stage(torch, step, local, min_ms=400.0, warmup=1, max_iters=3)
        finally:
            Fn.conv_select("auto")
        tf = B * T * FLOP / (ms * 1e-3) / 1e12
        res[conv] = {"B": B, "ms_per_step": ms, "iters": it, "sequence_steps_per_s": B * T / (ms * 1e-3), "achieved_tflops": tf,
                     "frac_of_bf16_sustained": tf / pk["bf16"], "frac_of_3pass_cap": 3 * tf / pk["bf16"],
                     "s_per_step_at_B32768": ms * (32768 / B) / 1e3, "clocks": ck}
        del x, u
        Fn.release_workspaces()
        torch.cuda.empty_cache()
    del m
    return {"workload": f"cfg5: VAE_HMM(5,256,K=32,128,u_dim=16) ELBO fwd+bwd, B=32768 T={T}, layer-wise path "
                        "(tc5: the full batch; ffma: a B=4096 slice, scaled)",
            "flop_per_step": FLOP, "layer_kernels": res,
            "note": "tc5 = tcgen05 GEMM-convolutions (2 x fp16 split, fp32 TMEM accumulators: at most 1/3 of the bf16 peak), fused "
                    "transition-prior kernel, decoder.embeddings folded into decoder.conv1; ffma = the FP32 layer kernels of round 1"}


def run_decode_sharded(torch, dist, V, Fn, model, dev, local, world, rank, pk):
    """BASELINE.json configs[2]: Viterbi + posterior decoding of B = 2^20 sequences (T = 256, K = 3) batch-sharded
    over the ranks -- pure replicas over disjoint shards, no collective in the data path.  Strong
    scaling: the 2^20 sequences are split; max-over-ranks device time; steps/s over all ranks."""
    B_total, T, K = (1 << 20), 256, 3
    Bl = B_total // world
    g = torch.Generator(device=dev).manual_seed(4321 + rank)
    res = {}
    with torch.no_grad():
        x = torch.randn(Bl, MODEL["C"], T, device=dev, generator=g)
        u = torch.randn(Bl, MODEL["U"], T, device=dev, generator=g)
        L = torch.full((Bl,), T, dtype=torch.int64, device=dev)
        log_pi, log_A, log_b = model.hmm_inputs(x, u)
        for key, bps, fn in (("decode_materialised", 61, lambda: Fn.hmm_decode(log_pi, log_A, log_b, L)),
                             ("decode_from_raw", 49, lambda: model.decode_regimes(x, u, L))):
            for _ in range(3):
                fn()
            torch.cuda.synchronize()
            if world > 1:
                dist.barrier()
            torch.cuda.synchronize()
            sampler = ClockSampler(local, interval=0.01)
            sampler.start()
            iters = 20
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(iters):
                fn()
            e1.record()
            torch.cuda.synchronize()
            clocks = sampler.stop()
            ms = e0.elapsed_time(e1) / iters
            if world > 1:
                t = torch.tensor([ms], device=dev, dtype=torch.float64)
                dist.all_reduce(t, op=dist.ReduceOp.MAX)
                ms = float(t)
            steps = B_total * T
            res[key] = {"ms": ms, "iters": iters, "steps_per_s": steps / (ms * 1e-3), "bytes_per_step": bps,
                        "achieved_gbs_per_gpu": bps * (steps / world) / (ms * 1e-3) / 1e9,
                        "frac_hbm_per_gpu": bps * (steps / world) / (ms * 1e-3) / 1e9 / pk["hbm"], "clocks": clocks}
        del log_A, log_b, x, u
    Fn.release_workspaces()
    torch.cuda.empty_cache()
    return {"workload": f"cfg3: Viterbi path + posterior gamma, B={B_total} total ({Bl} per GPU) T={T} K={K}, "
                        f"batch-sharded x{world}, no collective", "n_gpus": world, "scaling": "strong", **res}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--B", type=int, default=65536, help="sequences per GPU")
    ap.add_argument("--T", type=int, default=512)
    ap.add_argument("--cpu-batch", type=int, default=4096, help="sequences in the CPU baseline sample (~1.7 s per step on 16 cores)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-stages", action="store_true", help="skip the exact-HMM stage lines / cfg-3 decode block")
    ap.add_argument("--quick-stages", action="store_true", help="shorter stage timings")
    ap.add_argument("--scaling", default="weak", choices=["weak", "strong"],
                    help="weak: --B sequences per GPU (default, the contract's per-N lines); strong: --B sequences in total")
    args = ap.parse_args()
    if args.impl == "reference":
        return run_reference(args)

    import torch
    import torch.distributed as dist
    import vhmm_b200 as V
    from vhmm_b200 import functional as Fn

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a GPU (no CPU fallback); use --impl reference for the CPU arm")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    B, T, W, K = args.B, args.T, max(args.warmup, 3), args.steps   # timing rules: at least 3 warm-up steps ("warmup" in the line is the value used)
    if args.scaling == "strong":
        B = B // world

    # several ranks on one box: each pins itself to its GPU's NUMA node before the pinned host buffers are allocated
    numa_node = V.parallel.bind_to_gpu_numa_node(local) if world > 1 else None
    torch.manual_seed(0)
    model = V.VAE_HMM(MODEL["C"], MODEL["H"], MODEL["K"], MODEL["H2"], u_dim=MODEL["U"]).to(dev)
    if world > 1:
        V.parallel.enable_sharding(model)
    g = torch.Generator(device=dev).manual_seed(1234 + rank)
    x = torch.randn(B, MODEL["C"], T, device=dev, generator=g)
    u = torch.randn(B, MODEL["U"], T, device=dev, generator=g)
    lengths_cpu = torch.full((B,), T, dtype=torch.int64)
    lengths = lengths_cpu.to(dev)
    xh = torch.empty(x.shape, pin_memory=True); xh.copy_(x)
    uh = torch.empty(u.shape, pin_memory=True); uh.copy_(u)
    lh = lengths_cpu.pin_memory()
    beta = 1.0
    # N > 1: every rank knows the lengths of the WHOLE batch (here: all T), so the global normalisers come from the host and
    # the step has ONE collective (the all-reduce of [gradients || loss sums]) -- parallel.global_norms_from_lengths
    glen = {"global_lengths": torch.full((B * world,), T, dtype=torch.int64)} if world > 1 else {}

    def step_resident():
        model.zero_grad(set_to_none=True)
        loss = model.compute_loss(x, u, lengths, beta, **glen)
        loss.backward()
        return loss

    def step_e2e():
        # the public API call with HOST buffers (pinned): compute_loss streams the batch to the device
        # chunk by chunk underneath the kernels (functional.elbo_raw_streamed); lengths are a CPU
        # tensor exactly as the reference's collate_fn produces them
        model.zero_grad(set_to_none=True)
        loss = model.compute_loss(xh, uh, lh, beta, **glen)
        loss.backward()
        return loss.item()            # D2H of the step's result, every step

    def sync_all():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(fn, steps):
        sync_all()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(steps):
            out = fn()
        e1.record()
        sync_all()
        ms = e0.elapsed_time(e1) / steps
        if world > 1:
            t = torch.tensor([ms], device=dev, dtype=torch.float64)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ms = float(t)
        return ms, out

    for _ in range(W):
        step_resident()
    launches0 = V.lib().vhmm_launch_count()
    sampler = ClockSampler(local)
    sampler.start()
    import ctypes
    V.lib().vhmm_elbo_timing(1)
    ms, loss = timed(step_resident, K)
    clocks = sampler.stop()
    kt = (ctypes.c_double * 3)()
    kn = ctypes.c_int(0)
    V.lib().vhmm_elbo_timing_read(kt, ctypes.byref(kn))
    V.lib().vhmm_elbo_timing(0)
    ktimes = (kt[0], kt[1], kt[2], kn.value) if kn.value > 0 else None
    launches = (V.lib().vhmm_launch_count() - launches0) // K
    value = world * B * T / (ms * 1e-3)

    for _ in range(2):
        step_e2e()
    ms_e2e, _ = timed(step_e2e, K)
    e2e_value = world * B * T / (ms_e2e * 1e-3)
    h2d = xh.numel() * 4 + uh.numel() * 4 + lh.numel() * 8
    loss_val = float(loss.detach())

    # strong-scaling companion of the weak line (N > 1): the metric config's B sequences IN TOTAL, B/N per GPU --
    # the per-launch prologue (weight staging, partial reduction) and the all-reduce weigh N times more here
    strong = None
    if world > 1 and args.scaling == "weak":
        Bs = B // world
        xs, us, ls = x[:Bs], u[:Bs], lengths[:Bs]

        def step_strong():
            model.zero_grad(set_to_none=True)
            ll = model.compute_loss(xs, us, ls, beta, global_lengths=torch.full((Bs * world,), T, dtype=torch.int64))
            ll.backward()
            return ll
        for _ in range(3):
            step_strong()
        ms_s, _ = timed(step_strong, K)
        strong = {"scaling": "strong", "B_total": Bs * world, "B_per_gpu": Bs, "ms_per_step": ms_s,
                  "value": world * Bs * T / (ms_s * 1e-3), "unit": "sequence-steps/s"}
        xs = us = ls = None
    xh = uh = None
    Fn._stream_state.clear()

    pk = peaks()
    per_gpu_steps = B * T
    tfl = per_gpu_steps * FLOP_PER_STEP / (ms * 1e-3) / 1e12
    gbs = per_gpu_steps * BYTES_PER_STEP / (ms * 1e-3) / 1e9
    # dominant kernel: the decoder tile kernel (forward + backward of decoder.{conv1,conv2,to_params}
    # and the Gaussian NLL), timed with CUDA events on its own stream inside the timed region
    step_ctx = {"achieved_tflops": tfl, "flop_per_step": FLOP_PER_STEP,
                "fp32_ffma_peak_tflops": FP32_PEAK_TFLOPS, "frac_of_fp32_ffma_peak": tfl / FP32_PEAK_TFLOPS,
                "hbm_gbs_on_algorithmic_bytes": gbs, "bytes_per_step": BYTES_PER_STEP,
                "frac_of_measured_hbm": gbs / pk["hbm"]}
    if ktimes is not None and ktimes[3] > 0:
        n = ktimes[3]
        kms = [t / n for t in ktimes[:3]]
        dec_tfl = per_gpu_steps * DEC_FLOP_PER_STEP / (kms[1] * 1e-3) / 1e12
        dec = Fn.elbo_decoder_active()
        traffic = traffic_src = None   # dram bytes per launch of the active decoder kernel: NOT measured in this run,
        # read back from the committed `ncu --set full` capture of the same kernel and shape
        tp = os.path.join(ROOT, "profiles", TRAFFIC_FILE if dec == "tc5" else "r1_fused_decoder_traffic.json")
        if os.path.exists(tp):
            tj = json.load(open(tp))
            if tj.get("B") == B and tj.get("T") == T:
                traffic = tj["dram_bytes_per_launch"]
                traffic_src = "from committed ncu capture profiles/" + os.path.basename(tp) + " (not measured in this run)"
        roofline = {
            "kernel": ("vhmm::fz::decoder_tc5_kernel (decoder fwd+bwd + NLL; 64x64 k=3 GEMMs on tcgen05, 2xFP16 split)"
                       if dec == "tc5" else "vhmm::fz::decoder_kernel (decoder fwd+bwd + NLL, 3xTF32 mma.sync)"),
            "bound": "tensor", "achieved": dec_tfl, "peak": pk["bf16"], "unit": "TFLOP/s",
            "frac": dec_tfl / pk["bf16"], "traffic": traffic, "traffic_source": traffic_src,
            "peak_source": pk["src"] + " bf16 dense cuBLAS, sustained (kernel timed inside a long step)",
            "algorithmic_flop_per_step": DEC_FLOP_PER_STEP, "ms_per_launch": kms[1],
            "share_of_step": kms[1] / ms,
            "kernels_ms": {"enc_prior": kms[0], "decoder": kms[1], "enc_bwd": kms[2]},
            "note": ("fp32-parity path: every product of the three big GEMMs is 3 fp16 tcgen05 MMAs (h*h, h*m, m*h) with "
                     "fp32 accumulate in TMEM, so at most 1/3 of the bf16 peak is reachable; the skinny layers run as FP32 FFMA, the remaining small ones on mma.sync"
                     if dec == "tc5" else
                     "fp32-parity path: every product is 3 TF32 MMAs (hi*hi, lo*hi, hi*lo); legacy mma.sync TF32 "
                     "peaks at 512 MAC/clk/SM on B200 (tools/microbench.cu), i.e. 93 TFLOP/s of fp32-equivalent work"),
            "step": step_ctx,
        }
    else:
        roofline = {"kernel": "ELBO fwd+bwd step (layer-wise path, all launches)", "bound": "tensor",
                    "achieved": tfl, "peak": pk["bf16"], "unit": "TFLOP/s", "frac": tfl / pk["bf16"],
                    "traffic": None, "peak_source": pk["src"] + " bf16 dense cuBLAS, sustained", "step": step_ctx}
    line = {
        "metric": "sequence-steps/sec for ELBO fwd+bwd", "value": value, "unit": "sequence-steps/s",
        "n_gpus": world, "steps": K, "warmup": W, "ms_per_step": ms, "higher_is_better": True,
        "scaling": args.scaling, "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": f"cfg2: VAE_HMM(5,64,K=3,32,u_dim=4) ELBO fwd+bwd, B={B} per GPU x {world} GPU, T={T}, lengths=T",
                   "warmup_note": "at least 3 warm-up steps are always run (timing rules); `warmup` is the count used",
                   "parallelism": f"batch-sharded x{world}, 1 NCCL all-reduce/step" if world > 1 else "single GPU",
                   "l2": f"inputs ({(B * T * 36) / 1e9:.2f} GB/GPU) larger than the 126 MB L2; no flush needed",
                   "elbo_impl": Fn.elbo_active_impl(model._dims), "decoder_kernel": Fn.elbo_decoder_active()},
        "e2e": {"value": e2e_value, "unit": "sequence-steps/s", "ms_per_step": ms_e2e,
                "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": 4, "rank0_numa_node": numa_node},
        "gpu_launches": int(launches), "clocks": clocks, "roofline": roofline, "loss": loss_val,
    }
    if strong is not None:
        line["strong_scaling"] = strong
    x = u = None                      # free the cfg-2 batch before the stage workloads (cfg 3 needs ~30 GB)
    model.zero_grad(set_to_none=True)
    Fn.release_workspaces()
    torch.cuda.empty_cache()
    if not args.no_stages:
        if world == 1:
            line["stages"] = run_stages(torch, V, Fn, model, dev, local, pk, quick=args.quick_stages)
            line["train_cfg1"] = run_train_cfg1(torch, V, dev)
            line["cfg5"] = run_cfg5(torch, V, Fn, dev, local, pk)
        line["decode_cfg3"] = run_decode_sharded(torch, dist, V, Fn, model, dev, local, world, rank, pk)
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        rate, dt, n, kind = cpu_reference_step_rate(args.cpu_batch, T, 5, 1)   # ~10 s of CPU work
        line["cpu_baseline"] = {"value": rate, "unit": "sequence-steps/s", "cores": n, "kind": kind,
                                "sample": f"B={args.cpu_batch} of {B} sequences, T={T}, 5 timed iterations of compute_loss+backward: {CPU_KIND_TEXT[kind]}"}
    if rank == 0:
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
