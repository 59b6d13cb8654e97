"""host-side cost of one evaluate() of the bench population (128 equations, 463 evals) on a SMALL 3-D grid (64^3: the
kernels take ~1 ms, W = 99 windows as at 512^3), phase by phase: what e2e pays on top of the device step"""
import time, numpy as np, epde_b200
from epde_b200 import synthetic as syn
shape, kind = (64, 64, 64), 'heat'
tokens = syn.make_tokens(shape, kind, seed=5)
eqs = syn.make_population(len(tokens), 128, seed=5)
evals = syn.make_evals(eqs)
enc = epde_b200.encode(eqs, evals)
with epde_b200.Engine(shape, device=0) as eng:
    for i, t in enumerate(tokens):
        eng.upload(i, t)
    for _ in range(3):
        eng.evaluate(enc)
    acc = {'create': 0.0, 'launch': 0.0, 'sync': 0.0, 'fetch': 0.0, 'close': 0.0, 'total': 0.0}
    n = 20
    for _ in range(n):
        t0 = time.perf_counter()
        b = eng.batch(enc); t1 = time.perf_counter()
        b.set_norms(want_l2=False); b.gram(); b.solve(); b.residual(); t2 = time.perf_counter()
        eng.synchronize(); t3 = time.perf_counter()
        res = b.fetch(want_sw=True); t4 = time.perf_counter()
        b.close(); t5 = time.perf_counter()
        for k, v in zip(acc, (t1 - t0, t2 - t1, t3 - t2, t4 - t3, t5 - t4, t5 - t0)):
            acc[k] += v
    print({k: round(v / n * 1e3, 3) for k, v in acc.items()}, 'ms per call;', 'sw bytes', res.sw_weights.nbytes)
    t0 = time.perf_counter()
    for _ in range(n):
        e2 = epde_b200.encode(eqs, evals)
    print('encode (python):', round((time.perf_counter() - t0) / n * 1e3, 3), 'ms')
