import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import epde_b200
from epde_b200 import synthetic as syn
shape = tuple(int(x) for x in sys.argv[1].split('x'))
kind = 'wave' if len(shape) == 2 else 'heat'
tokens = syn.make_tokens(shape, kind, seed=9, noise=1e-3)
eqs = syn.make_population(len(tokens), int(sys.argv[2]), seed=3)
evals = syn.make_evals(eqs)
with epde_b200.Engine(shape) as eng:
    for i, t in enumerate(tokens):
        eng.upload(i, t)
    enc = epde_b200.encode(eqs, evals)
    for stage in ('gram', 'solve', 'residual'):
        with eng.batch(enc) as b:
            try:
                b.gram(); eng.synchronize(); print('gram ok')
                if stage == 'gram': continue
                b.solve(); eng.synchronize(); print('solve ok')
                if stage == 'solve': continue
                b.residual(); eng.synchronize(); print('residual ok')
            except Exception as e:
                print('FAILED in', stage, e); break
