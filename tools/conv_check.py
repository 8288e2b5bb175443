#!/usr/bin/env python3
"""conv1d layer kernels: tcgen05 build vs FFMA build vs torch float64, and their times at cfg-5 layer shapes."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch, torch.nn.functional as F
from vhmm_b200 import functional as Fn

def timed(fn, iters=5):
    for _ in range(2): fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters

g = torch.Generator(device="cuda").manual_seed(0)
for (B, Cin, Cout, T, kw) in ((3, 16, 128, 64, 3), (2, 32, 128, 200, 1), (5, 256, 256, 260, 3), (4, 128, 1024, 132, 1), (2, 256, 128, 1024, 3)):
    x = torch.randn(B, Cin, T, device="cuda", generator=g) * 3
    w = torch.randn(Cout, Cin, kw, device="cuda", generator=g) / (Cin * kw) ** 0.5
    b = torch.randn(Cout, device="cuda", generator=g)
    ref = F.relu(F.conv1d(x.double(), w.double(), b.double(), padding=kw // 2))
    out = {}
    for impl in ("ffma", "tc5"):
        Fn.conv_select(impl)
        y = Fn.conv1d(x, w, b, relu=True)
        out[impl] = float((y.double() - ref).abs().max() / ref.abs().max())
    Fn.conv_select("auto")
    print(f"B={B} Cin={Cin} Cout={Cout} T={T} kw={kw}: maxabs-rel err ffma {out['ffma']:.2e}  tc5 {out['tc5']:.2e}", flush=True)
for (B, Cin, Cout, T, kw) in ((2048, 256, 256, 1024, 3), (2048, 256, 128, 1024, 3), (2048, 128, 1024, 1024, 1)):
    x = torch.randn(B, Cin, T, device="cuda", generator=g)
    w = torch.randn(Cout, Cin, kw, device="cuda", generator=g) / (Cin * kw) ** 0.5
    b = torch.randn(Cout, device="cuda", generator=g)
    res = {}
    for impl in ("ffma", "tc5"):
        Fn.conv_select(impl)
        ms = timed(lambda: Fn.conv1d(x, w, b, relu=True))
        res[impl] = ms
    Fn.conv_select("auto")
    fl = 2.0 * B * T * Cin * Cout * kw
    print(f"B={B} Cin={Cin} Cout={Cout} T={T} kw={kw}: ffma {res['ffma']:.2f} ms ({fl/res['ffma']/1e9:.1f} TFLOP/s)  tc5 {res['tc5']:.2f} ms ({fl/res['tc5']/1e9:.1f} TFLOP/s)", flush=True)
    del x
