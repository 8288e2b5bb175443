import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from vhmm_b200 import functional as Fn
g = torch.Generator(device="cuda").manual_seed(0)
for (B, Cin, Cout, T, kw) in ((2048, 128, 1024, 1024, 1), (2048, 128, 256, 1024, 3)):
    x = torch.randn(B, Cin, T, device="cuda", generator=g)
    w = torch.randn(Cout, Cin, kw, device="cuda", generator=g) / (Cin * kw) ** 0.5
    b = torch.randn(Cout, device="cuda", generator=g)
    Fn.conv_select("tc5")
    for _ in range(2):
        Fn.conv1d(x, w, b, relu=True)
    torch.cuda.synchronize()
