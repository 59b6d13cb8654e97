import numpy as np, torch, epde_b200
from epde_b200 import synthetic as syn
import collections
shape, pop, kind = syn.CONFIGS['heat256']
eqs = syn.make_population(7, 128, seed=5); evals = syn.make_evals(eqs)
eng = epde_b200.Engine(shape, device=0)
dev = torch.device('cuda', 0)
for i in range(7):
    tk = syn.make_tokens(shape, kind, xp=torch, device=dev, seed=5, only=[i])[i]
    eng.upload_device(i, tk.data_ptr(), tk.numel()); torch.cuda.synchronize()
res = eng.evaluate(epde_b200.encode(eqs, evals), want_sw=False)
print('n_active hist', collections.Counter(res.n_active.tolist()))
nu=[]; nv=[]
for e, eq in enumerate(eqs):
    vs=[v for v,(ee,t) in enumerate(evals) if ee==e]
    used=set()
    for v in vs:
        t=evals[v][1]; used.add(t)
        feat=[i for i in range(len(eq)) if i!=t]
        for j,cf in zip(feat,res.coef[v]):
            if cf!=0: used.add(j)
    nu.append(len(used)); nv.append(len(vs))
print('T', collections.Counter(len(e) for e in eqs)); print('NU', collections.Counter(nu)); print('NV', collections.Counter(nv))
print('sum NU*NV', sum(a*b for a,b in zip(nu,nv)), 'sum (nact+1)', int(res.n_active.sum()+len(evals)), 'sumNU', sum(nu))
