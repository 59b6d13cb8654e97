"""one pass of both device token producers at 512^3 (for an ncu capture and for event timing):
fd_derivatives orders (2,2,2) and poly_derivatives orders (2,2,2), each followed by the boundary crops"""
import time, numpy as np, torch, epde_b200
from epde_b200 import producers
n = 512
dev = torch.device('cuda', 0)
t = torch.linspace(0, 1, n, dtype=torch.float64, device=dev)
u = (torch.exp(-t)[:, None, None] * torch.sin(3 * t)[None, :, None] * torch.cos(2 * t)[None, None, :]).contiguous()
coords = [np.linspace(0, 1, n)] * 3
eng = epde_b200.Engine((n, n, n), device=0)
N = n ** 3
for name, fn in (('fd_derivatives', producers.fd_derivatives), ('poly_derivatives', producers.poly_derivatives)):
    fn(eng, u, coords, (2, 2, 2), 0, first_token_id=1, field_token_id=0)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    fn(eng, u, coords, (2, 2, 2), 0, first_token_id=1, field_token_id=0)
    torch.cuda.synchronize()
    dt = time.perf_counter() - t0
    alg = 8 * N * (6 * 2 + 7 * 2)          # 6 derivatives (read + write) + 7 crops (read + write)
    print(f'{name} 512^3 orders (2,2,2): {dt * 1e3:.2f} ms wall (host weight tables included), algorithmic {alg / 1e9:.1f} GB -> {alg / dt / 1e9:.0f} GB/s')
