"""latency of one small Engine.evaluate call (what the per-equation operator mirrors pay)"""
import time, numpy as np, epde_b200
from epde_b200 import synthetic as syn
for cfg, neq in [('wave61', 1), ('wave61', 8), ('burgers156', 1), ('burgers156', 16)]:
    shape, _, kind = syn.CONFIGS[cfg]
    tokens = syn.make_tokens(shape, kind, seed=1)
    eqs = syn.make_population(len(tokens), neq, seed=1, t_fresh=(6, 6), pruned_fraction=0.0)
    evals = syn.make_evals(eqs)
    enc = epde_b200.encode(eqs, evals)
    with epde_b200.Engine(shape, device=0) as eng:
        for i, t in enumerate(tokens):
            eng.upload(i, t)
        for _ in range(5):
            eng.evaluate(enc)
        t0 = time.perf_counter()
        for _ in range(50):
            eng.evaluate(enc)
        dt = (time.perf_counter() - t0) / 50
        eng.set_timing(True)
        res = eng.evaluate(enc)
        kt = eng.kernel_times()
    print(f'{cfg} {neq} eq / {len(evals)} evals: sweeps {np.asarray(res.iters).tolist()[:8]} active {np.asarray(res.n_active).tolist()[:8]}')
    print(f'{cfg} {neq} eq / {len(evals)} evals: {dt * 1e6:.0f} us per call; kernels {dict((k, round(v[0] * 1e3)) for k, v in kt.items())} us')
