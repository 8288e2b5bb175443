import sys, torch
sys.path.insert(0, '/root/repo')
import vhmm_b200 as V
torch.manual_seed(0)
m = V.VAE_HMM(5, 64, 3, 32, u_dim=4).cuda()
B, T = 8192, 512
x = torch.randn(B, 5, T, device='cuda'); u = torch.randn(B, 4, T, device='cuda')
L = torch.full((B,), T, dtype=torch.int64)
for _ in range(2):
    m.zero_grad(); loss = m.compute_loss(x, u, L, 1.0); loss.backward()
torch.cuda.synchronize()
