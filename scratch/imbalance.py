"""which equations make a population-sharded rank slow?  kernel time of every part of the 8-way partition of heat256
on one GPU, then a bisection of the slowest part."""
import sys
import numpy as np, torch
sys.path.insert(0, '.')
import bench, epde_b200
from epde_b200 import sharding, synthetic as syn

shape, kind, eqs, evals = bench.build_workload('heat256', None, 5)
n_ev = [0] * len(eqs)
for e, _ in evals:
    n_ev[e] += 1
parts = sharding.population_partition([len(q) for q in eqs], n_ev, 8)
dev = torch.device('cuda', 0)
eng = epde_b200.Engine(shape, device=0, stream=torch.cuda.current_stream().cuda_stream)
for i in range(len(syn.token_labels(kind))):
    tk = syn.make_tokens(shape, kind, xp=torch, device=dev, seed=5, only=[i])[i]
    eng.upload_device(i, tk.data_ptr(), tk.numel())
    torch.cuda.synchronize()
    del tk


def cost(idx):
    sub = [eqs[i] for i in idx]
    enc = epde_b200.encode(sub, syn.make_evals(sub))
    for _ in range(2):
        eng.evaluate(enc, want_sw=False, want_l2=False)
    eng.set_timing(True)
    for _ in range(3):
        eng.evaluate(enc, want_sw=False, want_l2=False)
    kt = eng.kernel_times()
    eng.set_timing(False)
    return {k: round(v[0] / 3, 2) for k, v in kt.items()}


times = []
for r, p in enumerate(parts):
    c = cost(p)
    times.append(c['gram'] + c['residual'])
    print('part', r, c, flush=True)
slow = parts[int(np.argmax(times))]
fast = parts[int(np.argmin(times))]
print('slowest part', int(np.argmax(times)), 'fastest', int(np.argmin(times)))
# leave-one-out on the slow part: which equation removes the excess?
base = cost(slow)
for i in slow:
    c = cost([j for j in slow if j != i])
    d = (base['gram'] - c['gram'], base['residual'] - c['residual'])
    terms = [[(f[0], f[1]) for f in t] for t in eqs[i]]
    print(f'eq {i}: T {len(eqs[i])} evals {n_ev[i]} saves gram {d[0]:.2f} resid {d[1]:.2f}  slots {[sum(int(f[1]) for f in t) for t in eqs[i]]}', flush=True)
