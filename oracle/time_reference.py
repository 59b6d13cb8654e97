"""TEST / MEASUREMENT INFRASTRUCTURE ONLY (build container; needs /root/reference): wall-clock of the REFERENCE's own
evaluator -- ``EqRightPartSelector.apply`` + ``L2LRFitness.apply`` (right_part_selection.py:46-94, fitness.py:102-183),
unmodified, loaded from /root/reference -- on the grids of BASELINE.json that fit host RAM (BASELINE.md section 3
item 1: configs 1-3 and the 64^3 / 128^3 stand-ins of configs 4-5).  The numbers cannot be produced on the GPU box
(the reference does not travel), so they are generated here once and committed: profiles/reference_cpu_timing.json.

Unit as in the bench: one fitness eval = one (equation, target) LASSO + fitness; the selector runs D of them out of
place and one more in place.  usage: python oracle/time_reference.py [config ...]
"""
import json
import os
import platform
import signal
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import ref_loader as rl            # noqa: E402
from oracle import make_golden as mg           # noqa: E402
from epde_b200 import synthetic as syn         # noqa: E402

# name -> (inner shape, kind, equations to time, BASELINE config it stands for)
CASES = {
    'wave61': ((61, 61), 'wave', 8, 'config 1: wave 81x81, boundary 10 -> 61x61'),
    'burgers156': ((156, 156), 'wave', 8, 'config 2: Burgers 256x256, boundary 50 -> 156x156'),
    'heat64': ((64, 64, 64), 'heat', 4, 'stand-in of config 4 (256^3 does not fit: ~100 GB of window copies)'),
    'wave1024': ((1024, 1024), 'wave', 2, 'config 3: 1024x1024 (13.4 GB of window copies per LASSO)'),
    'heat128': ((128, 128, 128), 'heat', 2, 'stand-in of config 5 (512^3 is impossible for the reference)'),
}


class _Timeout(Exception):
    pass


def _alarm(signum, frame):
    raise _Timeout()


def time_case(name):
    shape, kind, n_eq, what = CASES[name]
    # the selector re-randomises terms in unbounded loops on degenerate equations (right_part_selection.py:55-57,
    # 76-78,142-144): bound every equation with an alarm, sized from the grid
    limit_s = 60 if int(np.prod(shape)) < 100000 else 3000
    rl.load_reference()
    coords = np.meshgrid(*[np.linspace(0.0, 1.0, n) for n in shape], indexing='ij')
    u = syn.make_tokens(shape, kind, seed=5)[0].reshape(shape)
    search = rl.make_search(coords, u, 0, (2,) * len(shape))
    l2lr, _, rps = rl.make_operators()
    labels = [k[0] for k in rl.cache_tokens().keys()]
    deriv_labels = [l for l in labels if l != 'u']
    rng = np.random.default_rng(17)
    eqs = mg.random_equations(rng, labels, deriv_labels, n_eq, t_range=(4, 6))
    rows, t_all, ev_all = [], 0.0, 0
    signal.signal(signal.SIGALRM, _alarm)
    for qi, terms in enumerate(eqs):
        T = len(terms)
        D = sum(1 for t in terms if sum(l in deriv_labels for l, _ in t) == 1)
        np.random.seed(1000 + qi)
        _, eq = rl.translate(mg.eq_text(terms, T - 1), search)
        signal.alarm(limit_s)
        t0 = time.perf_counter()
        try:
            rps.apply(eq, {})
            l2lr.apply(eq, {})
            ok = True
        except _Timeout:
            ok = False
        except Exception as exc:                      # noqa: BLE001  (degenerate random equations: skip, say so)
            print(f'[{name}] eq {qi}: {type(exc).__name__}: {exc}')
            ok = False
        finally:
            signal.alarm(0)
        dt = time.perf_counter() - t0
        if ok:
            rows.append(dict(terms=T, evals=D + 1, seconds=round(dt, 4)))
            t_all += dt
            ev_all += D + 1
        print(f'[{name}] eq {qi}: T={T} D={D} ok={ok} {dt:.2f} s', flush=True)
    return dict(case=name, stands_for=what, inner_shape=list(shape), equations=rows, evals=ev_all,
                seconds=round(t_all, 3), evals_per_s=(ev_all / t_all if t_all > 0 else None))


def main():
    names = sys.argv[1:] or ['wave61', 'burgers156', 'heat64']
    out_path = os.path.join(ROOT, 'profiles', 'reference_cpu_timing.json')
    try:
        with open(out_path) as f:
            doc = json.load(f)
    except Exception:                                 # noqa: BLE001
        doc = {}
    doc['_what'] = ("the UNMODIFIED reference (EqRightPartSelector.apply + L2LRFitness.apply from /root/reference) timed in the "
                    "build container; unit = fitness evals (D out-of-place + 1 in-place per equation); random 4-6 term equations "
                    "over u and its FD derivatives up to order 2, synthetic fields of epde_b200.synthetic")
    doc['_host'] = dict(cores=len(os.sched_getaffinity(0)), cpu=platform.processor() or platform.machine(),
                        numpy=np.__version__, blas_threads='default (unset)')
    for n in names:
        doc[n] = time_case(n)
        with open(out_path, 'w') as f:
            json.dump(doc, f, indent=1)
    print(json.dumps({k: v.get('evals_per_s') for k, v in doc.items() if not k.startswith('_')}))


if __name__ == '__main__':
    main()
