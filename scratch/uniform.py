"""experiment: pass A / pass B time when every equation runs the SAME code variant (no instruction-cache pressure)"""
import sys, numpy as np, torch, epde_b200
from epde_b200 import synthetic as syn
shape, pop, kind = syn.CONFIGS['heat512']
mode = sys.argv[1] if len(sys.argv) > 1 else 'uniform'
rng = np.random.default_rng(1)
if mode == 'uniform':
    eqs = []
    for k in range(128):
        toks = rng.permutation(6)[:5] + 1
        eq = [[(int(t), 1.0, 0, 0.0)] for t in toks[:4]] + [[(0, 1.0, 0, 0.0)]]
        eqs.append(eq)
else:
    eqs = syn.make_population(7, 128, seed=5)
evals = syn.make_evals(eqs)
dev = torch.device('cuda', 0)
eng = epde_b200.Engine(shape, device=0)
for i in range(7):
    tk = syn.make_tokens(shape, kind, xp=torch, device=dev, seed=5, only=[i])[i]
    eng.upload_device(i, tk.data_ptr(), tk.numel()); torch.cuda.synchronize(); del tk
enc = epde_b200.encode(eqs, evals)
eng.evaluate(enc, want_sw=False)
eng.set_timing(True)
for _ in range(2):
    res = eng.evaluate(enc, want_sw=False)
kt = eng.kernel_times()
print(mode, 'evals', len(evals), {k: round(v[0] / 2, 2) for k, v in kt.items()}, 'n_active', np.bincount(res.n_active).tolist())
