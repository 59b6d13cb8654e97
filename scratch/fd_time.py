"""throughput of the device FD producer at 512^3 (field + 6 derivatives)"""
import time, numpy as np, torch, epde_b200
from epde_b200 import producers
n = 512
dev = torch.device('cuda', 0)
t = torch.linspace(0, 1, n, dtype=torch.float64, device=dev)
u = (torch.exp(-t)[:, None, None] * torch.sin(3 * t)[None, :, None] * torch.cos(2 * t)[None, None, :]).contiguous()
coords = [np.linspace(0, 1, n)] * 3
eng = epde_b200.Engine((n, n, n), device=0)
producers.fd_derivatives(eng, u, coords, (2, 2, 2), 0, first_token_id=1, field_token_id=0)
torch.cuda.synchronize()
t0 = time.perf_counter()
for _ in range(3):
    producers.fd_derivatives(eng, u, coords, (2, 2, 2), 0, first_token_id=1, field_token_id=0)
torch.cuda.synchronize()
dt = (time.perf_counter() - t0) / 3
N = n ** 3
alg = 8 * N * (6 * 2 + 7 * 2)          # 6 gradients (read + write) + 7 crops (read + write)
print(f'fd_derivatives 512^3 order (2,2,2): {dt * 1e3:.2f} ms, algorithmic {alg / 1e9:.1f} GB -> {alg / dt / 1e9:.0f} GB/s')
