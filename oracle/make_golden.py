"""TEST INFRASTRUCTURE ONLY -- generate tests/golden/*.npz by running the REAL reference.

Run in the build container (needs /root/reference):   python oracle/make_golden.py
The fixtures it writes are committed; the GPU box only ever reads the fixtures.

Each fixture holds the inputs (masked token vectors exactly as the reference's tensor cache
stores them, inner grid shape, equation encodings) and the outputs of the reference's own
operators on them:
  * per (equation, target): LASSOSparsity + L2LRFitness(force_out_of_place=True)
    (operators/common/sparsity.py:178-211, fitness.py:105-150): weights_internal, intercept,
    cached sliding-window weights, L1 fitness (NaN where the reference returns None);
    plus the stability objective of those cached weights (fitness.py:160-169) and
    L2Fitness' value for the same weights (fitness.py:70-91);
  * per equation: EqRightPartSelector.apply + in-place L2LRFitness.apply
    (right_part_selection.py:46-94): chosen target / kept terms / fitness / stability;
  * raw micro-fixtures for calculate_weights and PhysicsInformedLasso.fit on plain arrays,
    including ill-conditioned inputs.
"""
import os
import signal
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
sys.path.insert(0, ROOT)

from oracle import ref_loader as rl  # noqa: E402

OUT = os.path.join(ROOT, 'tests', 'golden')
KMAX = 8      # padded width of the per-eval arrays (T <= 7 in the fixtures)


def random_equations(rng, labels, deriv_labels, n_eq, t_range=(3, 6)):
    """Distinct-term random equations: list of list of [(label, power), ...]."""
    eqs = []
    while len(eqs) < n_eq:
        T = int(rng.integers(t_range[0], t_range[1] + 1))
        terms, seen = [], set()
        tries = 0
        while len(terms) < T and tries < 200:
            tries += 1
            nf = 1 if rng.random() < 0.7 else 2
            chosen = rng.choice(len(labels), size=nf, replace=False)
            term = []
            for c in chosen:
                lab = labels[int(c)]
                pmax = 3 if lab == 'u' else 2
                p = int(rng.integers(1, pmax + 1)) if rng.random() < 0.35 else 1
                term.append((lab, p))
            key = frozenset(term)
            if key in seen:
                continue
            seen.add(key)
            terms.append(term)
        n_adm = sum(1 for t in terms if sum(l in deriv_labels for l, _ in t) == 1)
        if len(terms) == T and n_adm >= 1:
            eqs.append(terms)
    return eqs


def eq_text(terms, target):
    def tstr(term):
        return ' * '.join('%s{power: %d}' % (l, p) for l, p in term)
    feats = [t for i, t in enumerate(terms) if i != target]
    return ' + '.join('1.0 * ' + tstr(t) for t in feats) + ' + 0. = ' + tstr(terms[target])


def _alarm(signum, frame):
    raise TimeoutError('reference call exceeded its time box')


def term_key(term):
    return frozenset((f.label, float(f.params[0])) for f in term.structure)


def run_case(name, coords, data, boundary, max_deriv_order, n_random, seed, extra_eqs=(),
             derivs=None, preproc='FD', preproc_kwargs=None, extra_arrays=None, t_range=(3, 6)):
    rl.load_reference()
    import epde.globals as gv
    search = rl.make_search(coords, data, boundary, max_deriv_order, derivs=derivs, preproc=preproc,
                            preproc_kwargs=preproc_kwargs)
    l2lr, l2, rps = rl.make_operators()
    tokens = rl.cache_tokens()
    labels = [k[0] for k in tokens.keys()]
    deriv_labels = [l for l in labels if l != 'u']
    tok_id = {l: i for i, l in enumerate(labels)}
    shape = np.array(gv.grid_cache.inner_shape, dtype=np.int64)
    ndim = len(shape)
    g = gv.grid_cache.g_func[gv.grid_cache.g_func_mask]
    assert np.all(g == 1.0)

    rng = np.random.default_rng(seed)
    eqs = list(extra_eqs) + random_equations(rng, labels, deriv_labels, n_random, t_range=t_range)

    term_off, fac_off, fac_tok, fac_pow = [0], [0], [], []
    ev_eq, ev_tgt, ev_coef, ev_b, ev_fit, ev_fit_l2, ev_stab, ev_sw, ev_nact = [], [], [], [], [], [], [], [], []
    W = None
    rps_ok, rps_target, rps_keep, rps_fit, rps_stab, rps_wfinal = [], [], [], [], [], []
    for qi, terms in enumerate(eqs):
        for term in terms:
            for lab, p in term:
                fac_tok.append(tok_id[lab])
                fac_pow.append(float(p))
            fac_off.append(len(fac_tok))
        term_off.append(len(fac_off) - 1)
        T = len(terms)
        admissible = [i for i, t in enumerate(terms) if sum(l in deriv_labels for l, _ in t) == 1]
        # ---- per-target out-of-place evaluation (the loop body of right_part_selection.py:60-74)
        for tgt in admissible:
            _, eq = rl.translate(eq_text(terms, tgt), search)
            order = [i for i in range(T) if i != tgt] + [tgt]
            # the translated equation lists features in `order` with the target last
            eq.reset_state(True)
            eq.target_idx = T - 1
            fit = l2lr.apply(eq, {}, force_out_of_place=True)
            wi = np.asarray(eq.weights_internal, dtype=np.float64)
            sw = np.asarray(eq._cached_sw_weights, dtype=np.float64)
            b = float(sw.mean(axis=0)[-1])
            W = sw.shape[0]
            coef = np.full(KMAX, np.nan)
            coef[:T - 1] = wi                      # feature order == term order minus target
            swp = np.full((W, KMAX), np.nan)
            swp[:, :sw.shape[1]] = sw
            # stability of the cached weights as the in-place call computes it (fitness.py:160-169)
            with np.errstate(divide='ignore', invalid='ignore'):
                std = sw.std(axis=0, ddof=1)
                mu = sw.mean(axis=0)
                cv = (std ** 2) / (mu ** 2)
                cv[mu == 0] = 0.0
            stab = float(sum(cv[:-1]) / ndim)
            # L2Fitness on the same weights (fitness.py:70-91)
            if fit is not None:
                eq.weights_internal_evald = True
                eq.weights_final_evald = True
                eq.weights_final = np.append(wi, b)       # L2Fitness indexes it as an ndarray
                eq._eval_cache = {}
                try:
                    l2.apply(eq, {}, force_out_of_place=False)
                    fit_l2 = float(eq.fitness_value)
                except ValueError:
                    # fitness.py:81-82 multiplies an (N,) residual by the FULL-grid g_func_flat:
                    # it only broadcasts when boundary == 0 (SURVEY.md section 7 'bug-compatibility')
                    fit_l2 = np.nan
            else:
                fit_l2 = np.nan
            ev_eq.append(qi); ev_tgt.append(tgt); ev_coef.append(coef); ev_b.append(b)
            ev_fit.append(np.nan if fit is None else float(fit)); ev_fit_l2.append(fit_l2)
            ev_stab.append(stab); ev_sw.append(swp); ev_nact.append(sw.shape[1] - 1)
        # ---- full selector + in-place fitness
        np.random.seed(seed * 1000 + qi)
        _, eq = rl.translate(eq_text(terms, T - 1), search)
        order = list(range(T))                      # target last in text == original order
        keys0 = [term_key(t) for t in eq.structure]
        ok = True
        # the selector re-randomises terms in unbounded loops on degenerate equations
        # (right_part_selection.py:55-57,76-78,142-144): bound it with an alarm
        signal.signal(signal.SIGALRM, _alarm)
        signal.alarm(30)
        try:
            rps.apply(eq, {})
            l2lr.apply(eq, {})
        except Exception as exc:                    # noqa: BLE001
            print('   rps failed:', type(exc).__name__, exc)
            ok = False
        finally:
            signal.alarm(0)
        keep = np.zeros(KMAX, dtype=np.int8)
        tgt_orig, fitv, stabv = -1, np.nan, np.nan
        wf = np.full(KMAX + 1, np.nan)
        if ok:
            keys1 = [term_key(t) for t in eq.structure]
            if all(k in keys0 for k in keys1):
                idx = [keys0.index(k) for k in keys1]
                keep[idx] = 1
                tgt_orig = idx[eq.target_idx]
                fitv, stabv = float(eq.fitness_value), float(eq.coefficients_stability)
                wfl = list(eq.weights_final)
                wf[:len(wfl)] = wfl
            else:
                ok = False                           # simplification rewrote terms (RNG-dependent)
        rps_ok.append(ok); rps_target.append(tgt_orig); rps_keep.append(keep)
        rps_fit.append(fitv); rps_stab.append(stabv); rps_wfinal.append(wf)
        print(f'[{name}] eq {qi}: T={T} admissible={admissible} rps_ok={ok} target={tgt_orig} '
              f'fit={fitv:.6g} stab={stabv:.6g}')

    np.savez_compressed(
        os.path.join(OUT, name + '.npz'),
        tokens=np.stack([tokens[(l, (1.0,))] for l in labels]), labels=np.array(labels),
        shape=shape, term_off=np.array(term_off, dtype=np.int32), fac_off=np.array(fac_off, dtype=np.int32),
        fac_tok=np.array(fac_tok, dtype=np.int32), fac_pow=np.array(fac_pow, dtype=np.float64),
        ev_eq=np.array(ev_eq, dtype=np.int32), ev_tgt=np.array(ev_tgt, dtype=np.int32),
        ev_coef=np.array(ev_coef), ev_b=np.array(ev_b), ev_fit=np.array(ev_fit),
        ev_fit_l2=np.array(ev_fit_l2), ev_stab=np.array(ev_stab), ev_sw=np.array(ev_sw),
        ev_nact=np.array(ev_nact, dtype=np.int32),
        rps_ok=np.array(rps_ok), rps_target=np.array(rps_target, dtype=np.int32),
        rps_keep=np.array(rps_keep), rps_fit=np.array(rps_fit), rps_stab=np.array(rps_stab),
        rps_wfinal=np.array(rps_wfinal), **(extra_arrays or {}))


def micro_fixtures():
    """calculate_weights / PhysicsInformedLasso.fit on plain arrays."""
    rl.load_reference()
    from epde.supplementary import calculate_weights
    from epde.operators.common.sparsity import PhysicsInformedLasso
    rng = np.random.default_rng(7)
    out = {}
    cases = [((61, 61), 4, 'plain'), ((40, 52), 5, 'plain'), ((20, 24, 18), 5, 'plain'),
             ((33, 17, 29), 6, 'collinear'), ((64, 48), 4, 'duplicate'), ((30, 34), 3, 'nearconst'),
             ((50, 50), 5, 'zerointercept'), ((9, 8), 2, 'tiny'), ((156, 156), 6, 'plain')]
    for ci, (shape, F, kind) in enumerate(cases):
        axes = [np.linspace(0, 1, n) for n in shape]
        gr = np.meshgrid(*axes, indexing='ij')
        cols = [np.sin((k + 1) * np.pi * gr[k % len(shape)]) * np.cos((k + 2) * gr[(k + 1) % len(shape)])
                + 0.3 * rng.standard_normal(shape) for k in range(F)]
        X = np.stack([c.reshape(-1) for c in cols], axis=1)
        w = rng.standard_normal(F) * (rng.random(F) < 0.6)
        y = X @ w + 0.5 + 0.05 * rng.standard_normal(X.shape[0])
        if kind == 'collinear':
            X[:, 1] = X[:, 0] * 1.0000001 + 1e-7 * rng.standard_normal(X.shape[0])
        if kind == 'duplicate':
            X[:, 2] = X[:, 0]
        if kind == 'nearconst':
            X[:, 1] = 3.0 + 1e-6 * rng.standard_normal(X.shape[0])
        if kind == 'zerointercept':
            y = X @ w
        g = np.ones(X.shape[0])
        cw = calculate_weights(X, y, g, np.array(shape))
        est = PhysicsInformedLasso(grid_shape=np.array(shape))
        est.fit(X, y, g)
        out[f'c{ci}_shape'] = np.array(shape)
        out[f'c{ci}_X'] = X
        out[f'c{ci}_y'] = y
        out[f'c{ci}_cw'] = cw
        out[f'c{ci}_coef'] = est.coef_
        out[f'c{ci}_b'] = np.float64(est.intercept_)
        out[f'c{ci}_cached'] = est.cached_weights_
        out[f'c{ci}_kind'] = np.array(kind)
        print(f'[micro] case {ci} {shape} F={F} {kind}: selected {(est.coef_ != 0).astype(int)}')
    out['n_cases'] = np.int64(len(cases))
    np.savez_compressed(os.path.join(OUT, 'micro.npz'), **out)


def micro2_fixtures():
    """round 2 additions, same entry points as micro_fixtures(): (a) WIDE systems -- F = 11, 15, 29 features, i.e.
    equations of T = 12, 16, 30 terms (north_star 'n_terms ~ 5-30'; SURVEY 8d stress runs); (b) NON-TRIVIAL sample
    weights g (a user `function_form`, interface.py:511-550): the window regressions are weighted
    (supplementary.py:455) while the LASSO sums stay unweighted (sparsity.py:54-55,73-76)."""
    rl.load_reference()
    from epde.supplementary import calculate_weights
    from epde.operators.common.sparsity import PhysicsInformedLasso
    rng = np.random.default_rng(11)
    out = {}
    cases = [((40, 44), 11, 'wide', False), ((36, 40), 15, 'wide', False), ((48, 52), 29, 'wide', False),
             ((20, 18, 16), 15, 'wide', False),
             ((61, 61), 4, 'weighted', True), ((40, 52), 5, 'weighted', True), ((20, 24, 18), 5, 'weighted', True),
             ((36, 40), 11, 'weighted', True)]
    for ci, (shape, F, kind, weighted) in enumerate(cases):
        axes = [np.linspace(0, 1, n) for n in shape]
        gr = np.meshgrid(*axes, indexing='ij')
        nd = len(shape)
        cols = [np.sin((k % 5 + 1) * np.pi * gr[k % nd] + 0.37 * k) * np.cos((k % 7 + 2) * gr[(k + 1) % nd] + 0.11 * k)
                + 0.3 * rng.standard_normal(shape) for k in range(F)]
        X = np.stack([c.reshape(-1) for c in cols], axis=1)
        w = rng.standard_normal(F) * (rng.random(F) < 0.5)
        y = X @ w + 0.5 + 0.05 * rng.standard_normal(X.shape[0])
        if weighted:
            # a smooth positive bump, like the inverse-polynomial test function the reference keeps commented out
            bump = np.ones(shape)
            for a in gr:
                bump = bump * (1.0 - 0.9 * (2 * a - 1) ** 2)
            g = (0.05 + bump).reshape(-1)
        else:
            g = np.ones(X.shape[0])
        cw = calculate_weights(X, y, g, np.array(shape))
        est = PhysicsInformedLasso(grid_shape=np.array(shape))
        est.fit(X, y, g)
        out[f'c{ci}_shape'] = np.array(shape)
        out[f'c{ci}_X'] = X
        out[f'c{ci}_y'] = y
        out[f'c{ci}_g'] = g
        out[f'c{ci}_cw'] = cw
        out[f'c{ci}_coef'] = est.coef_
        out[f'c{ci}_b'] = np.float64(est.intercept_)
        out[f'c{ci}_cached'] = est.cached_weights_
        out[f'c{ci}_kind'] = np.array(kind)
        print(f'[micro2] case {ci} {shape} F={F} {kind}: selected {int((est.coef_ != 0).sum())} of {F}')
    out['n_cases'] = np.int64(len(cases))
    np.savez_compressed(os.path.join(OUT, 'micro2.npz'), **out)


def procedural_case(ref):
    """(round 2) procedural token families evaluated by CustomEvaluator in the reference (evaluators.py:47-89 with
    trig_eval_fun_np :141 and grid_eval_fun_np :150): sin / cos of a coordinate and powers of a coordinate, mixed with
    the data tokens of the wave case.  Stores the per-target LASSOSparsity + L2LRFitness(out of place) results of the
    unmodified reference for hand-written equations; the engine evaluates the same factors in-kernel (kinds 1, 2 and
    0 on the coordinate token)."""
    rl.load_reference()
    import epde.globals as gv
    from epde.interface.prepared_tokens import TrigonometricTokens, GridTokens
    d = np.loadtxt(os.path.join(ref, 'examples/data/wave_sln_80.csv'), delimiter=',').T
    t = np.linspace(0, 1, d.shape[0]); x = np.linspace(0, 1, d.shape[1])
    coords = np.meshgrid(t, x, indexing='ij')
    trig = TrigonometricTokens(freq=(0.5, 6.5), dimensionality=1)
    grid = GridTokens(['t', 'x'], max_power=2, dimensionality=1)
    search = rl.make_search(coords, d, 8, (2, 2), additional_tokens=[trig, grid])
    l2lr, _, _ = rl.make_operators()
    tokens = rl.cache_tokens()
    labels = [k[0] for k in tokens.keys()]
    mask = gv.grid_cache.g_func != 0
    grids = [np.asarray(gv.grid_cache.get(str(dim)))[mask].reshape(-1) for dim in range(2)]
    tok_arrays = [np.asarray(v, dtype=np.float64).reshape(-1) for v in tokens.values()] + grids
    tok_id = {l: i for i, l in enumerate(labels)}
    gid = {0: len(labels), 1: len(labels) + 1}
    KIND = {'sin': 1, 'cos': 2}
    # factor = (label, power) for data tokens, (label, power, freq, dim) for trig, (label, power, dim) for grids
    eqs = [
        [[('sin', 1, 3.0, 1), ('du/dx1', 1)], [('d^2u/dx1^2', 1)], [('u', 1)], [('d^2u/dx0^2', 1)]],
        [[('cos', 1, 2.0, 0)], [('u', 1), ('sin', 1, 1.5, 1)], [('d^2u/dx1^2', 1)], [('du/dx0', 1)]],
        [[('x', 1, 1), ('du/dx1', 1)], [('t', 2, 0)], [('d^2u/dx1^2', 1)], [('u', 2)], [('d^2u/dx0^2', 1)]],
        [[('cos', 1, 6.0, 1), ('u', 1)], [('x', 2, 1)], [('du/dx1', 1)], [('d^2u/dx1^2', 1)], [('du/dx0', 1)]],
        [[('sin', 1, 0.7, 0), ('cos', 1, 4.2, 1)], [('u', 1)], [('d^2u/dx1^2', 1)], [('d^2u/dx0^2', 1)]],
    ]

    def ftext(f):
        if f[0] in KIND:
            return '%s{power: %d, freq: %r, dim: %d}' % (f[0], f[1], float(f[2]), f[3])
        if f[0] in ('t', 'x'):
            return '%s{power: %d, dim: %d}' % (f[0], f[1], f[2])
        return '%s{power: %d}' % (f[0], f[1])

    def text(terms, target):
        ts = [' * '.join(ftext(f) for f in term) for term in terms]
        feats = [q for i, q in enumerate(ts) if i != target]
        return ' + '.join('1.0 * ' + q for q in feats) + ' + 0. = ' + ts[target]

    term_off, fac_off, fac_tok, fac_pow, fac_kind, fac_par = [0], [0], [], [], [], []
    ev_eq, ev_tgt, ev_coef, ev_b, ev_fit, ev_sw, ev_nact = [], [], [], [], [], [], []
    for qi, terms in enumerate(eqs):
        for term in terms:
            for f in term:
                if f[0] in KIND:
                    fac_tok.append(gid[f[3]]); fac_pow.append(float(f[1])); fac_kind.append(KIND[f[0]]); fac_par.append(float(f[2]))
                elif f[0] in ('t', 'x'):
                    fac_tok.append(gid[f[2]]); fac_pow.append(float(f[1])); fac_kind.append(0); fac_par.append(0.0)
                else:
                    fac_tok.append(tok_id[f[0]]); fac_pow.append(float(f[1])); fac_kind.append(0); fac_par.append(0.0)
            fac_off.append(len(fac_tok))
        term_off.append(len(fac_off) - 1)
        T = len(terms)
        for tgt in range(T):
            if not any(f[0].startswith('d') for f in terms[tgt]):
                continue
            _, eq = rl.translate(text(terms, tgt), search)
            eq.reset_state(True)
            eq.target_idx = T - 1
            fit = l2lr.apply(eq, {}, force_out_of_place=True)
            wi = np.asarray(eq.weights_internal, dtype=np.float64)
            sw = np.asarray(eq._cached_sw_weights, dtype=np.float64)
            coef = np.full(KMAX, np.nan); coef[:T - 1] = wi
            swp = np.full((sw.shape[0], KMAX), np.nan); swp[:, :sw.shape[1]] = sw
            ev_eq.append(qi); ev_tgt.append(tgt); ev_coef.append(coef); ev_b.append(float(sw.mean(axis=0)[-1]))
            ev_fit.append(np.nan if fit is None else float(fit)); ev_sw.append(swp); ev_nact.append(sw.shape[1] - 1)
            print(f'[procedural] eq {qi} target {tgt}: fit={fit} weights={wi}')
    np.savez_compressed(
        os.path.join(OUT, 'procedural.npz'), shape=np.array(gv.grid_cache.inner_shape, dtype=np.int64),
        labels=np.array(labels + ['grid0', 'grid1']), tokens=np.stack(tok_arrays),
        term_off=np.array(term_off, dtype=np.int32), fac_off=np.array(fac_off, dtype=np.int32),
        fac_tok=np.array(fac_tok, dtype=np.int32), fac_pow=np.array(fac_pow), fac_kind=np.array(fac_kind, dtype=np.int32),
        fac_par=np.array(fac_par), ev_eq=np.array(ev_eq, dtype=np.int32), ev_tgt=np.array(ev_tgt, dtype=np.int32),
        ev_coef=np.array(ev_coef), ev_b=np.array(ev_b), ev_fit=np.array(ev_fit), ev_sw=np.array(ev_sw),
        ev_nact=np.array(ev_nact, dtype=np.int32))


def main():
    os.makedirs(OUT, exist_ok=True)
    ref = rl.REFERENCE_ROOT
    which = sys.argv[1:] or ['micro', 'wave', 'burgers', 'heat']
    if 'micro' in which:
        micro_fixtures()
    if 'micro2' in which:
        micro2_fixtures()
    if 'procedural' in which:
        procedural_case(ref)
    if 'wave' in which:
        wave_cases(ref)
    if 'burgers' in which:
        burgers_case(ref)
    if 'burgers_poly' in which:
        burgers_poly_case(ref)
    if 'heat' in which:
        heat_case()
    if 'heat2' in which:
        heat_case2()


def wave_cases(ref):

    # (1) wave, examples/data/wave_sln_80.csv, the two equations of projects/pic/data/wave/wave.py:102-103
    d = np.loadtxt(os.path.join(ref, 'examples/data/wave_sln_80.csv'), delimiter=',').T
    t = np.linspace(0, 1, d.shape[0]); x = np.linspace(0, 1, d.shape[1])
    coords = np.meshgrid(t, x, indexing='ij')
    wave_eqs = [[[('d^2u/dx1^2', 1)], [('d^2u/dx0^2', 1)]],
                [[('d^2u/dx1^2', 1), ('du/dx0', 1)], [('d^2u/dx0^2', 1), ('du/dx0', 1)]]]
    run_case('wave_b5', coords, d, 5, (2, 2), n_random=14, seed=1, extra_eqs=wave_eqs)
    run_case('wave_b10', coords, d, 10, (2, 2), n_random=10, seed=2, extra_eqs=wave_eqs)
    # round 2: a third wave fixture so that the wave data carries >= 50 reference-evaluated equations
    run_case('wave_b8', coords, d, 8, (2, 2), n_random=26, seed=7, extra_eqs=wave_eqs)


def burgers_case(ref):
    # (2) Burgers 256x256, boundary 50 (projects/burgers/burgers_eq.py:96)
    import scipy.io as sio
    m = sio.loadmat(os.path.join(ref, 'projects/burgers/burgers.mat'))
    u = np.ascontiguousarray(m['u'].T if m['u'].shape[0] != m['t'].size else m['u'])
    tb = m['t'].reshape(-1); xb = m['x'].reshape(-1)
    coords = np.meshgrid(tb, xb, indexing='ij')
    burgers_eqs = [[[('u', 1), ('du/dx1', 1)], [('d^2u/dx1^2', 1)], [('du/dx0', 1)]]]
    run_case('burgers_b50', coords, u, 50, (1, 2), n_random=8, seed=3, extra_eqs=burgers_eqs)


def burgers_poly_case(ref):
    # (2b, round 2) BASELINE.json config 2 as written: Burgers 256x256, boundary 50, NOISY data (noise_data of
    # projects/pic/data/burgers/burgers.py:38-40, 1 %) differentiated by the reference's 'poly' preprocessor
    # (PolynomialDeriv: per-point Chebyshev fits over 9-point windows, deriv_calculators.py:111-134, cheb.py:14-85).
    # The full-grid noisy field and the coordinate vectors are stored too: the device Chebyshev producer is checked
    # against the derivative tokens of this fixture.
    import scipy.io as sio
    m = sio.loadmat(os.path.join(ref, 'projects/burgers/burgers.mat'))
    u = np.ascontiguousarray(m['u'].T if m['u'].shape[0] != m['t'].size else m['u'])
    tb = m['t'].reshape(-1); xb = m['x'].reshape(-1)
    np.random.seed(12)
    noisy = 1.0 * 0.01 * np.std(u) * np.random.normal(size=u.shape) + u
    coords = np.meshgrid(tb, xb, indexing='ij')
    burgers_eqs = [[[('u', 1), ('du/dx1', 1)], [('d^2u/dx1^2', 1)], [('du/dx0', 1)]]]
    run_case('burgers_poly_b50', coords, noisy, 50, (1, 2), n_random=8, seed=6, extra_eqs=burgers_eqs, preproc='poly',
             preproc_kwargs={'use_smoothing': False, 'mp_poolsize': 4, 'polynomial_window': 9, 'poly_order': None},
             extra_arrays={'data_full': noisy, 'axis0': tb, 'axis1': xb, 'boundary': np.int64(50),
                           'polynomial_window': np.int64(9)})


def heat_case2():
    # (round 2) second 3-D fixture: 36^3, boundary 3 -> 30^3, equations of 4..7 terms
    n = 36
    tt = np.linspace(0, 1, n); xx = np.linspace(0, np.pi, n); yy = np.linspace(0, np.pi, n)
    T3, X3, Y3 = np.meshgrid(tt, xx, yy, indexing='ij')
    u3 = (np.exp(-T3) * np.sin(X3) * np.sin(Y3) + 0.4 * np.exp(-2.5 * T3) * np.sin(2 * X3) * np.sin(Y3)
          + 0.2 * np.exp(-4 * T3) * np.sin(X3) * np.sin(3 * Y3))
    u3 = u3 + 1e-3 * np.random.default_rng(8).standard_normal(u3.shape)
    run_case('heat3d_b3', (T3, X3, Y3), u3, 3, (1, 2, 2), n_random=8, seed=10, t_range=(4, 7))


def heat_case():
    # (3) synthetic 3-D heat, 32^3, boundary 4 (SURVEY.md section 6 probe recipe)
    n = 32
    tt = np.linspace(0, 1, n); xx = np.linspace(0, np.pi, n); yy = np.linspace(0, np.pi, n)
    T3, X3, Y3 = np.meshgrid(tt, xx, yy, indexing='ij')
    k = 0.5
    u3 = (np.exp(-2 * k * T3) * np.sin(X3) * np.sin(Y3) + 0.3 * np.exp(-5 * k * T3) * np.sin(2 * X3) * np.sin(Y3))
    u3 = u3 + 1e-3 * np.random.default_rng(5).standard_normal(u3.shape)
    heat_eqs = [[[('d^2u/dx1^2', 1)], [('d^2u/dx2^2', 1)], [('u', 1)], [('du/dx1', 1)], [('du/dx0', 1)]]]
    run_case('heat3d_b4', (T3, X3, Y3), u3, 4, (1, 2, 2), n_random=7, seed=4, extra_eqs=heat_eqs)


if __name__ == '__main__':
    main()
