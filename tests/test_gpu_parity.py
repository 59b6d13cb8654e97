"""Parity of the CUDA path (through the C ABI) with the reference's golden vectors and with the
numpy oracle on the same seeded inputs.  Bar (BASELINE.json): identical selected term sets;
coefficients / fitness / stability within 1e-6 relative (fp64)."""
import numpy as np
import pytest

from conftest import EQ_FIXTURES, fixture_equations, load_fixture
from helpers import assert_rel, assert_sw, natural_scales

import epde_b200
from epde_b200 import synthetic as syn
from oracle import epde_oracle as orc

pytestmark = pytest.mark.gpu
COND_MAX = 1e9      # oracle.window_condition: beyond this the reference itself is rounding noise


def make_engine(shape, tokens, **kw):
    eng = epde_b200.Engine(shape, **kw)
    for i, t in enumerate(tokens):
        eng.upload(i, t)
    return eng


class Scales(np.ndarray):
    """natural_scales() array carrying ||target||_2 for the L2 fitness floor."""
    t_norm = 0.0


def eval_scales(tokens, eq, tgt):
    cols = orc.term_columns(tokens, eq)
    sc = natural_scales(np.delete(cols, tgt, axis=1), cols[:, tgt]).view(Scales)
    sc.t_norm = float(np.linalg.norm(cols[:, tgt]))
    return sc


def compare_eval(res, v, ref, T, ndim, what, check_sw=True):
    """res: BatchResult row v; ref: oracle EvalResult-like dict."""
    F = T - 1
    coef = res.coef[v, :F]
    assert np.array_equal(coef != 0, ref['coef'] != 0), f'{what}: selected set differs {coef} vs {ref["coef"]}'
    assert np.all(res.coef[v, F:] == 0)
    sc = ref['scales']                      # natural scale per feature, then the intercept
    assert_rel(coef, ref['coef'], f'{what} coef', atol=1e-9 * sc[:F])
    assert_rel(res.intercept[v], ref['intercept'], f'{what} intercept', atol=1e-9 * sc[-1])
    nact = int(np.count_nonzero(ref['coef']))
    assert int(res.n_active[v]) == nact
    assert bool(res.all_zero[v]) == (nact == 0)
    if check_sw and ref.get('sw') is not None:
        act = np.append(ref['coef'] != 0, True)
        assert_sw(res.cached_weights(v), ref['sw'], f'{what} sw', col_scales=sc[act])
    if ref.get('fitness') is not None and not np.isnan(ref['fitness']):
        assert_rel(res.fitness_l1[v], ref['fitness'], f'{what} fitness', atol=1e-10)
    if ref.get('fitness_l2') is not None and not np.isnan(ref['fitness_l2']):
        assert_rel(res.fitness_l2[v], ref['fitness_l2'], f'{what} fitness_l2', atol=1e-9 * getattr(sc, 't_norm', 0.0))
    if ref.get('stability') is not None and not np.isnan(ref['stability']):
        assert_rel(res.stability[v], ref['stability'], f'{what} stability', rtol=1e-5, atol=1e-14)


@pytest.mark.parametrize('name', EQ_FIXTURES)
def test_golden_equations(name):
    """every (equation, target) eval of the reference-generated fixtures."""
    fx = load_fixture(name)
    eqs = fixture_equations(fx)
    shape = tuple(int(s) for s in fx['shape'])
    evals = list(zip(fx['ev_eq'].tolist(), fx['ev_tgt'].tolist()))
    with make_engine(shape, fx['tokens']) as eng:
        assert eng.n_windows == fx['ev_sw'].shape[1]
        res = eng.evaluate(epde_b200.encode(eqs, evals))
    # conditioning of each eval (oracle, CPU) decides whether strict parity is meaningful
    n_ill = 0
    for e, eq in enumerate(eqs):
        sel = [v for v, (ee, _) in enumerate(evals) if ee == e]
        if not sel:
            continue
        ores = orc.evaluate_equation(fx['tokens'], shape, eq, targets=[evals[v][1] for v in sel])
        for v, o in zip(sel, ores):
            if o.cond > COND_MAX:
                n_ill += 1
                assert np.all(np.isfinite(res.coef[v]))
                continue
            T = len(eq)
            nact = int(fx['ev_nact'][v])
            ref = dict(coef=fx['ev_coef'][v, :T - 1], intercept=fx['ev_b'][v],
                       sw=fx['ev_sw'][v, :, :nact + 1], fitness=fx['ev_fit'][v],
                       fitness_l2=fx['ev_fit_l2'][v], stability=fx['ev_stab'][v],
                       scales=eval_scales(fx['tokens'], eq, evals[v][1]))
            compare_eval(res, v, ref, T, len(shape), f'{name} eval {v}')
            # the reference returns None when everything vanished (fitness.py:123)
            assert bool(res.all_zero[v]) == bool(np.isnan(fx['ev_fit'][v]))
    assert n_ill <= 0.25 * len(evals)


@pytest.mark.parametrize('ci', range(9))
def test_golden_micro(ci):
    """calculate_weights / PhysicsInformedLasso.fit on plain arrays: columns uploaded as tokens."""
    z = load_fixture('micro')
    X, y = z[f'c{ci}_X'], z[f'c{ci}_y']
    shape = tuple(int(s) for s in z[f'c{ci}_shape'])
    kind = str(z[f'c{ci}_kind'])
    F = X.shape[1]
    toks = [X[:, j] for j in range(F)] + [y]
    eq = [[(j, 1.0, 0, 0.0)] for j in range(F + 1)]
    enc = epde_b200.encode([eq], [(0, F)])
    with make_engine(shape, toks) as eng:
        res = eng.evaluate(enc)
        # mode 1: window solves on ALL features = calculate_weights(X, y)
        with eng.batch(enc) as b:
            b.set_given([1], np.zeros((1, epde_b200.MAX_TERMS)), [0.0])
            b.gram(); b.solve(); b.residual()
            full = b.fetch()
    ref_coef, ref_b = z[f'c{ci}_coef'], float(z[f'c{ci}_b'])
    assert np.array_equal(res.coef[0, :F] != 0, ref_coef != 0), f'{kind}: selected set differs'
    Xa = np.column_stack([X, np.ones(X.shape[0])])
    if kind in ('collinear', 'duplicate', 'nearconst'):
        ref_pred = Xa @ np.append(ref_coef, ref_b)
        pred = Xa @ np.append(res.coef[0, :F], res.intercept[0])
        assert np.abs(pred - ref_pred).max() <= 1e-5 * np.abs(ref_pred).max()
        ref_pw = Xa @ z[f'c{ci}_cw'].T
        assert np.abs(Xa @ full.sw_weights[0, :, :F + 1].T - ref_pw).max() <= 1e-5 * np.abs(ref_pw).max()
        return
    sc = natural_scales(X, y)
    assert_rel(res.coef[0, :F], ref_coef, 'coef', atol=1e-9 * sc[:F])
    assert_rel(res.intercept[0], ref_b, 'intercept', atol=1e-9 * sc[-1])
    assert_sw(res.cached_weights(0), z[f'c{ci}_cached'], 'cached weights', col_scales=sc[np.append(ref_coef != 0, True)])
    assert_sw(full.sw_weights[0, :, :F + 1], z[f'c{ci}_cw'], 'calculate_weights', col_scales=sc)


@pytest.mark.parametrize('ci', range(8))
def test_golden_micro2_wide_and_weighted(ci):
    """reference-generated fixtures of round 2: equations of T = 12, 16, 30 terms (one warp per window system, four
    column blocks of the dictionary) and NON-TRIVIAL sample weights (weighted window Grams, unweighted LASSO sums:
    supplementary.py:455 vs sparsity.py:54-55,73-76)."""
    z = load_fixture('micro2')
    X, y, g = z[f'c{ci}_X'], z[f'c{ci}_y'], z[f'c{ci}_g']
    shape = tuple(int(s) for s in z[f'c{ci}_shape'])
    kind = str(z[f'c{ci}_kind'])
    F = X.shape[1]
    toks = [X[:, j] for j in range(F)] + [y]
    eq = [[(j, 1.0, 0, 0.0)] for j in range(F + 1)]
    enc = epde_b200.encode([eq], [(0, F)])
    with make_engine(shape, toks) as eng:
        if kind == 'weighted':
            eng.set_sample_weights(g)
        res = eng.evaluate(enc)
        with eng.batch(enc) as b:
            b.set_given([1], np.zeros((1, epde_b200.MAX_TERMS)), [0.0])
            b.gram(); b.solve(); b.residual()
            full = b.fetch()
    ref_coef, ref_b = z[f'c{ci}_coef'], float(z[f'c{ci}_b'])
    assert np.array_equal(res.coef[0, :F] != 0, ref_coef != 0), f'{kind} F={F}: selected set differs'
    sc = natural_scales(X, y)
    assert_rel(res.coef[0, :F], ref_coef, 'coef', atol=1e-9 * sc[:F])
    assert_rel(res.intercept[0], ref_b, 'intercept', atol=1e-9 * sc[-1])
    assert_sw(res.cached_weights(0), z[f'c{ci}_cached'], 'cached weights', col_scales=sc[np.append(ref_coef != 0, True)])
    assert_sw(full.sw_weights[0, :, :F + 1], z[f'c{ci}_cw'], 'calculate_weights', col_scales=sc)
    # the L1 fitness is unweighted whatever g is (fitness.py:142)
    pred = X @ ref_coef + ref_b
    fit = np.abs(y - pred).sum() / np.abs(y).sum() if np.any(ref_coef != 0) else np.abs(y - y.mean()).sum() / np.abs(y).sum()
    assert_rel(res.fitness_l1[0], fit, 'L1 fitness', rtol=1e-6)
    # L2Fitness multiplies its residual by g (fitness.py:81-82)
    assert_rel(res.fitness_l2[0], np.linalg.norm((pred - y) * g), 'L2 fitness (weighted residual)', rtol=1e-6)


def oracle_refs(tokens, shape, eqs, evals):
    out = {}
    for e, eq in enumerate(eqs):
        sel = [v for v, (ee, _) in enumerate(evals) if ee == e]
        if not sel:
            continue
        for v, o in zip(sel, orc.evaluate_equation(tokens, shape, eq, targets=[evals[v][1] for v in sel])):
            out[v] = o
    return out


def check_against_oracle(res, tokens, shape, eqs, evals, what, min_checked=1):
    refs = oracle_refs(tokens, shape, eqs, evals)
    checked = 0
    for v, o in refs.items():
        if not np.isfinite(o.cond) or o.cond > COND_MAX:
            continue
        eq = eqs[evals[v][0]]
        ref = dict(coef=o.coef, intercept=o.intercept, sw=o.sw_weights, fitness=o.fitness,
                   fitness_l2=o.fitness_l2, stability=o.stability,
                   scales=eval_scales(tokens, eq, evals[v][1]))
        compare_eval(res, v, ref, len(eq), len(shape), f'{what} eval {v}')
        checked += 1
    assert checked >= min_checked, f'{what}: only {checked} well-posed evals'
    return checked


@pytest.mark.parametrize('cfg,pop,seed', [('wave61', 24, 1), ('burgers156', 16, 2), ('heat64', 12, 3)])
def test_synthetic_vs_oracle(cfg, pop, seed):
    shape, _, kind = syn.CONFIGS[cfg]
    tokens = syn.make_tokens(shape, kind, seed=seed)
    eqs = syn.make_population(len(tokens), pop, seed=seed)
    evals = syn.make_evals(eqs)
    with make_engine(shape, tokens) as eng:
        res = eng.evaluate(epde_b200.encode(eqs, evals))
    check_against_oracle(res, tokens, shape, eqs, evals, cfg, min_checked=len(evals) // 2)


def test_golden_procedural_tokens():
    """reference-generated fixture with sin / cos / coordinate-power factors (CustomEvaluator families,
    evaluators.py:47-89,141,150) evaluated in-kernel from the coordinate tokens (factor kinds 1, 2, 0)."""
    from conftest import fixture_equations_procedural
    fx = load_fixture('procedural')
    shape = tuple(int(s) for s in fx['shape'])
    eqs = fixture_equations_procedural(fx)
    evals = [(int(e), int(t)) for e, t in zip(fx['ev_eq'], fx['ev_tgt'])]
    with make_engine(shape, list(fx['tokens'])) as eng:
        res = eng.evaluate(epde_b200.encode(eqs, evals))
    checked = 0
    for v, (e, t) in enumerate(evals):
        o = orc.evaluate_equation(fx['tokens'], shape, eqs[e], targets=[t])[0]
        if o.cond > COND_MAX:
            continue
        F = len(eqs[e]) - 1
        ref = fx['ev_coef'][v, :F]
        assert np.array_equal(res.coef[v, :F] != 0, ref != 0), f'eval {v}: selected set'
        assert_rel(res.coef[v, :F], ref, f'eval {v} coef', rtol=1e-6, atol=1e-9 * max(1.0, np.abs(ref).max()))
        assert_rel(res.intercept[v], fx['ev_b'][v], f'eval {v} intercept', rtol=1e-6, atol=1e-9)
        assert bool(res.all_zero[v]) == bool(np.isnan(fx['ev_fit'][v]))
        if not res.all_zero[v]:
            assert_rel(res.fitness_l1[v], fx['ev_fit'][v], f'eval {v} fitness', rtol=1e-6)
        checked += 1
    assert checked >= 8


@pytest.mark.parametrize('kind,pop', [('distinct', 48), ('wide16', 6), ('wide30', 4)])
def test_stress_populations_vs_oracle(kind, pop):
    """populations that stress the column dictionary: 'distinct' = no term shared between equations (several launch
    groups, the grouping must not change any result); 'wide16' / 'wide30' = 16 / 30-term equations (SURVEY 8d stress
    runs; window systems of up to 31 unknowns, pass-B evals with many surviving terms)."""
    shape = (24, 20, 22) if kind == 'distinct' else (40, 36)
    knd = 'heat' if len(shape) == 3 else 'wave'
    tokens = syn.make_tokens(shape, knd, seed=4, noise=5e-2)
    if kind == 'distinct':
        eqs = syn.make_population_distinct(len(tokens), pop, seed=2)
    else:
        eqs = syn.make_population_wide(len(tokens), pop, int(kind[4:]), seed=2, max_power=2)
    evals = syn.make_evals(eqs)
    with make_engine(shape, tokens) as eng:
        enc = epde_b200.encode(eqs, evals)
        with eng.batch(enc) as b:
            info = b.info()
            b.gram(); b.solve(); b.residual()
            res = b.fetch()
        # the same evals one equation at a time: grouping / batch composition must not matter beyond rounding
        solo = [eng.evaluate(epde_b200.encode([eq], [(0, t) for (e, t) in evals if e == i])) for i, eq in enumerate(eqs[:6])]
    if kind == 'distinct':
        assert info['groups'] >= 2, info
    v0 = 0
    for i, r in enumerate(solo):
        n = r.coef.shape[0]
        F = len(eqs[i]) - 1
        assert np.array_equal(res.coef[v0:v0 + n, :F] != 0, r.coef[:, :F] != 0)
        assert_rel(res.fitness_l1[v0:v0 + n], r.fitness_l1, f'{kind} eq {i} fitness batch vs solo', rtol=1e-7)
        v0 += n
    check_against_oracle(res, tokens, shape, eqs, evals, kind, min_checked=max(1, len(evals) // 4))


@pytest.mark.parametrize('shape', [(9, 8), (2, 2), (5, 300), (300, 5), (7, 6, 5), (33, 17, 29), (31, 260), (4, 3, 513)])
def test_ragged_shapes_vs_oracle(shape):
    """odd / tiny / long-thin grids: window plan edge cases (supplementary.py:411-425)."""
    rng = np.random.default_rng(11)
    N = int(np.prod(shape))
    axes = np.meshgrid(*[np.linspace(0, 1, n) for n in shape], indexing='ij')
    toks = [np.sin(3 * axes[0] + k) + np.cos((k + 2) * axes[-1]) + 0.2 * rng.standard_normal(shape) for k in range(4)]
    toks = [t.reshape(-1) for t in toks]
    eqs = [[[(0, 1)], [(1, 1)], [(2, 1)], [(3, 1)]],
           [[(0, 2)], [(1, 1), (2, 1)], [(3, 1)]],
           [[(0, 1)], [(3, 3)]]]
    eqs = [[[(t, float(p), 0, 0.0) for t, p in term] for term in eq] for eq in eqs]
    evals = [(0, 0), (0, 3), (1, 0), (1, 2), (2, 1), (2, 0)]
    with make_engine(shape, toks) as eng:
        res = eng.evaluate(epde_b200.encode(eqs, evals))
    if min(shape) < 4:
        # a window of a single slab: ddof=1 statistics of <=2 windows; only finiteness of coefs
        assert res.coef.shape[0] == len(evals)
        return
    check_against_oracle(res, toks, shape, eqs, evals, f'shape {shape}')


def test_factor_kinds_vs_oracle():
    """sin / cos / inverse procedural factors (evaluators.py:141-151) and non-integer powers."""
    shape = (40, 36)
    rng = np.random.default_rng(5)
    axes = np.meshgrid(np.linspace(0.1, 2.0, shape[0]), np.linspace(0.5, 3.0, shape[1]), indexing='ij')
    toks = [axes[0].reshape(-1), axes[1].reshape(-1),
            (np.sin(2 * axes[0]) * axes[1] + 0.05 * rng.standard_normal(shape)).reshape(-1),
            (np.cos(axes[1]) + axes[0] ** 2 + 0.05 * rng.standard_normal(shape)).reshape(-1)]
    K = epde_b200
    eqs = [[[(2, 1.0, 0, 0.0)], [(0, 1.0, K.KIND_SIN, 2.0)], [(1, 1.0, K.KIND_COS, 1.0)], [(3, 1.0, 0, 0.0)]],
           [[(3, 1.0, 0, 0.0)], [(0, 2.0, K.KIND_INVERSE, 0.0)], [(1, 2.0, K.KIND_SIN, 1.3), (0, 1.0, 0, 0.0)], [(0, 2.0, 0, 0.0)]],
           [[(2, 1.0, 0, 0.0)], [(1, 1.5, 0, 0.0)], [(0, 3.0, K.KIND_COS, 0.7)]]]
    evals = [(0, 0), (0, 3), (1, 0), (2, 0)]
    with make_engine(shape, toks) as eng:
        res = eng.evaluate(K.encode(eqs, evals))
    check_against_oracle(res, toks, shape, eqs, evals, 'kinds', min_checked=3)


def test_high_integer_powers_vs_oracle():
    """integer powers above 3 (multiplications by squaring in the kernel, numpy's pow in the oracle), alone and inside
    products, negative bases included; the distinct stress population of the bench uses powers up to 5."""
    shape = (36, 30, 12)
    rng = np.random.default_rng(9)
    axes = np.meshgrid(*[np.linspace(0.0, 1.0 + d, n) for d, n in enumerate(shape)], indexing='ij')
    toks = [(np.sin(2 * axes[0] + k) * np.cos(axes[1] - k) + 0.4 * axes[2] - 0.3 + 0.05 * rng.standard_normal(shape)).reshape(-1)
            for k in range(4)]
    eqs = [[[(0, 1)], [(1, 4)], [(2, 5)], [(3, 1)]],
           [[(1, 1)], [(0, 4), (2, 1)], [(3, 7)], [(2, 6), (0, 2)]],
           [[(2, 1)], [(3, 5), (1, 4), (0, 3)], [(0, 8)]]]
    eqs = [[[(t, float(p), 0, 0.0) for t, p in term] for term in eq] for eq in eqs]
    evals = [(0, 0), (0, 3), (1, 0), (2, 0)]
    with make_engine(shape, toks) as eng:
        res = eng.evaluate(epde_b200.encode(eqs, evals))
    check_against_oracle(res, toks, shape, eqs, evals, 'high powers', min_checked=3)


def test_staged_equals_oneshot_and_is_deterministic():
    shape, _, kind = syn.CONFIGS['wave61']
    tokens = syn.make_tokens(shape, kind, seed=4)
    eqs = syn.make_population(len(tokens), 10, seed=4)
    enc = epde_b200.encode(eqs, syn.make_evals(eqs))
    with make_engine(shape, tokens) as eng:
        a = eng.evaluate(enc)
        with eng.batch(enc) as b:
            b.gram(); b.solve(); b.residual()
            c = b.fetch()
        d = eng.evaluate(enc)
        assert eng.launch_count > 0
    for k in ('coef', 'intercept', 'fitness_l1', 'fitness_l2', 'stability', 'sw_weights', 'n_active', 'iters'):
        assert np.array_equal(getattr(a, k), getattr(c, k)), k
        assert np.array_equal(getattr(a, k), getattr(d, k)), k


def test_l1_only_mode_equals_default():
    """epde_batch_set_norms(0): pass B skips the sum of squares; every other output is bit-identical to the
    default run and fitness_l2 comes back as NaN."""
    for cfg, pop, seed in [('wave61', 12, 6), ('heat64', 8, 7)]:
        shape, _, kind = syn.CONFIGS[cfg]
        tokens = syn.make_tokens(shape, kind, seed=seed)
        eqs = syn.make_population(len(tokens), pop, seed=seed)
        enc = epde_b200.encode(eqs, syn.make_evals(eqs))
        with make_engine(shape, tokens) as eng:
            a = eng.evaluate(enc)
            b = eng.evaluate(enc, want_l2=False)
        for k in ('coef', 'intercept', 'fitness_l1', 'stability', 'sw_weights', 'n_active', 'all_zero', 'iters'):
            assert np.array_equal(getattr(a, k), getattr(b, k)), (cfg, k)
        assert np.all(np.isnan(b.fitness_l2)) and np.all(np.isfinite(a.fitness_l2))


def test_given_weights_modes():
    """in-place fitness with known weights (fitness.py:70-96,130-174)."""
    shape, _, kind = syn.CONFIGS['wave61']
    tokens = syn.make_tokens(shape, kind, seed=6)
    eq = [[(2, 1.0, 0, 0.0)], [(4, 1.0, 0, 0.0)], [(0, 1.0, 0, 0.0)]]
    enc = epde_b200.encode([eq], [(0, 0)])
    w = np.zeros((1, epde_b200.MAX_TERMS)); w[0, 1] = 0.04; w[0, 2] = -0.01
    cols = orc.term_columns(tokens, eq)
    X = cols[:, 1:]; y = cols[:, 0]
    pred = X @ np.array([0.04, -0.01]) + 0.003
    l1 = np.sum(np.abs(y - pred)) / np.sum(np.abs(y))
    l2 = np.linalg.norm(pred - y)
    cw = orc.calculate_weights_direct(X, y, np.ones(len(y)), shape)
    with make_engine(shape, tokens) as eng:
        for mode in (1, 2):
            with eng.batch(enc) as b:
                b.set_given([mode], w, [0.003])
                b.gram(); b.solve(); b.residual()
                r = b.fetch()
            assert_rel(r.fitness_l1[0], l1, 'given l1')
            assert_rel(r.fitness_l2[0], l2, 'given l2')
            assert_rel(r.coef[0, :2], [0.04, -0.01], 'given coef', rtol=0)
            if mode == 1:
                assert_sw(r.sw_weights[0, :, :3], cw, 'window weights', col_scales=natural_scales(X, y))
                assert_rel(r.stability[0], orc.stability_from_weights(cw, 2), 'stability', rtol=1e-6)
            else:
                assert np.isnan(r.stability[0])


def test_cache_api():
    shape = (12, 10)
    rng = np.random.default_rng(0)
    a = rng.standard_normal(120)
    with epde_b200.Engine(shape) as eng:
        assert 3 not in eng
        eng.upload(3, a)
        assert 3 in eng and eng.consumed_memory == 120 * 8
        assert np.array_equal(eng.download(3), a)
        eng.upload(3, 2 * a)                       # overwrite
        assert np.array_equal(eng.download(3), 2 * a)
        with pytest.raises(epde_b200.EpdeB200Error, match='expects'):
            eng.upload(4, a[:50])
        with pytest.raises(epde_b200.EpdeB200Error, match='not in the cache'):
            eng.evaluate(epde_b200.encode([[[(9, 1.0)], [(3, 1.0)]]]))
        with pytest.raises(epde_b200.EpdeB200Error):
            eng.evaluate(epde_b200.encode([[[(3, 1.0)]] * (epde_b200.MAX_TERMS + 1)]))          # too many terms
        eng.set_sample_weights(np.ones(120))
        with pytest.raises(epde_b200.EpdeB200Error, match='size mismatch'):
            eng.set_sample_weights(np.full(50, 0.5))
        eng.delete(3)
        assert 3 not in eng and eng.consumed_memory == 0


def two_shards_vs_whole(shape, cuts, enc, whole_engine, shard_engines):
    """two shard contexts on one device + a manual sum of their segment Grams / solution blocks / residual sums
    (exactly what evaluate_grid_sharded does with NCCL all-reduces) against the unsharded evaluation."""
    import torch
    full = whole_engine.evaluate(enc)
    batches, segs, sums = [], [], []
    for e in shard_engines:
        b = e.batch(enc)
        batches.append(b)
        s = torch.zeros(b.gram_size, dtype=torch.float64, device='cuda')
        b.gram(s.data_ptr())
        segs.append(s)
    torch.cuda.synchronize()
    tot = segs[0] + segs[1]
    # every shard solves half of the evals (epde_batch_solve_part); the solution blocks are merged by a sum
    sols = []
    for r in range(2):
        segs[r].copy_(tot)
        sol = torch.zeros(batches[r].solution_size(True), dtype=torch.float64, device='cuda')
        batches[r].solve_part(segs[r].data_ptr(), r, 2, sol.data_ptr(), with_sw=True)
        sols.append(sol)
    torch.cuda.synchronize()
    assert float((sols[0][:enc.n_eval * epde_b200.MAX_TERMS].abs() > 0).sum()) > 0
    merged = sols[0] + sols[1]
    for r in range(2):
        batches[r].set_solution(merged.data_ptr(), with_sw=True)
        s = torch.zeros(batches[r].residual_size, dtype=torch.float64, device='cuda')
        batches[r].residual(s.data_ptr())
        sums.append(s)
    torch.cuda.synchronize()
    tots = sums[0] + sums[1]
    outs = [batches[r].fetch(tots.data_ptr()) for r in range(2)]
    for b in batches:
        b.close()
    for k in ('coef', 'n_active', 'iters', 'sw_weights'):
        assert np.array_equal(getattr(outs[0], k), getattr(outs[1], k)), k
    assert outs[0].n_active.max() > 0 and outs[0].iters.max() > 0
    assert np.array_equal(outs[0].coef != 0, full.coef != 0)
    # the slabs sum their cells in another order than the whole grid: the contract tolerance (1e-6 relative)
    # applies, random equations of these populations reach ~1e-8 through their conditioning
    assert_rel(outs[0].coef, full.coef, 'sharded coef', rtol=1e-6, atol=1e-14)
    assert_rel(outs[0].fitness_l1, full.fitness_l1, 'sharded fitness', rtol=1e-6, atol=1e-10)
    assert_rel(outs[0].stability, full.stability, 'sharded stability', rtol=1e-6, atol=1e-14)
    return full


def test_grid_sharding_single_gpu():
    """uneven cut of a small grid; the NCCL variant only replaces the sums."""
    shape = (40, 24, 28)
    tokens = syn.make_tokens(shape, 'heat', seed=8)
    eqs = syn.make_population(len(tokens), 8, seed=8)
    enc = epde_b200.encode(eqs, syn.make_evals(eqs))
    cuts = [0, 13, 40]
    row = shape[1] * shape[2]
    shards = []
    for r in range(2):
        lo, hi = cuts[r], cuts[r + 1]
        e = epde_b200.Engine((hi - lo,) + shape[1:], shard=(lo, shape[0]))
        for i, t in enumerate(tokens):
            e.upload(i, t[lo * row:hi * row])
        shards.append(e)
    with make_engine(shape, tokens) as eng:
        two_shards_vs_whole(shape, cuts, enc, eng, shards)
    for e in shards:
        e.close()


def device_engine(shape, kind, seed, axis0_range=None, global_rows=None):
    """engine whose tokens are generated on the device (counter-based noise: a slab holds the rows of the whole grid)."""
    import torch
    lo, hi = axis0_range if axis0_range else (0, shape[0])
    kw = dict(shard=(lo, global_rows)) if axis0_range else {}
    eng = epde_b200.Engine((hi - lo,) + tuple(shape[1:]), device=0, **kw)
    for i in range(len(syn.token_labels(kind))):
        tk = syn.make_tokens(shape, kind, xp=torch, device=torch.device('cuda', 0), seed=seed, axis0_range=(lo, hi), only=[i])[i]
        torch.cuda.synchronize()
        eng.upload_device(i, tk.data_ptr(), tk.numel())
        eng.synchronize()
        del tk
    return eng


def test_slab_sharding_at_size_512x256x256():
    """a 2-shard slab run of a 512 x 256 x 256 grid (population 16, every admissible target) against
    the unsharded run of the same grid: identical selected sets, coefficients / fitness / stability at 1e-6."""
    shape, kind = (512, 256, 256), 'heat'
    eqs = syn.make_population(len(syn.token_labels(kind)), 16, seed=12)
    enc = epde_b200.encode(eqs, syn.make_evals(eqs))
    cuts = [0, 256, 512]
    whole = device_engine(shape, kind, 12)
    shards = [device_engine(shape, kind, 12, axis0_range=(cuts[r], cuts[r + 1]), global_rows=shape[0]) for r in range(2)]
    try:
        full = two_shards_vs_whole(shape, cuts, enc, whole, shards)
        assert full.coef.shape[0] == enc.n_eval >= 16
    finally:
        for e in [whole] + shards:
            e.close()


def test_heat256_population_vs_streamed_oracle():
    """parity at a size the bench quotes -- heat 256^3, population 16 evaluated as one batch; every
    eval of the first 4 equations against the streamed oracle (oracle.evaluate_equation_streamed)."""
    shape, kind = (256, 256, 256), 'heat'
    tokens = syn.make_tokens(shape, kind, seed=6)
    eqs = syn.make_population(len(tokens), 16, seed=6)
    evals = syn.make_evals(eqs)
    with make_engine(shape, tokens) as eng:
        res = eng.evaluate(epde_b200.encode(eqs, evals))
    checked = 0
    for e in range(4):
        sel = [v for v, (ee, _) in enumerate(evals) if ee == e]
        outs = orc.evaluate_equation_streamed(tokens, shape, eqs[e], targets=[evals[v][1] for v in sel])
        for v, o in zip(sel, outs):
            if not np.isfinite(o.cond) or o.cond > COND_MAX:
                continue
            ref = dict(coef=o.coef, intercept=o.intercept, sw=o.sw_weights, fitness=o.fitness, fitness_l2=o.fitness_l2,
                       stability=o.stability, scales=eval_scales(tokens, eqs[e], evals[v][1]))
            compare_eval(res, v, ref, len(eqs[e]), 3, f'heat256 eq {e} eval {v}')
            checked += 1
    assert checked >= 4, checked


def test_large_grid_properties():
    """config 3 size (1024^2): properties that need no oracle at full size --
    (i) duplicated equations inside a batch give bit-identical rows; (ii) permuting the terms of
    an equation permutes the coefficients; (iii) a planted exact linear relation is recovered with
    ~zero residual; (iv) spot parity with the oracle on 2 equations."""
    shape, _, kind = syn.CONFIGS['wave1024']
    tokens = syn.make_tokens(shape, kind, seed=9, noise=1e-3)
    planted = 0.5 * tokens[4] - 2.0 * tokens[1] + 0.25
    toks = tokens + [planted]
    base = [[(5, 1.0, 0, 0.0)], [(4, 1.0, 0, 0.0)], [(1, 1.0, 0, 0.0)]]
    perm = [base[0], base[2], base[1]]
    wave = [[(2, 1.0, 0, 0.0)], [(4, 1.0, 0, 0.0)], [(0, 1.0, 0, 0.0)], [(3, 1.0, 0, 0.0), (0, 2.0, 0, 0.0)]]
    eqs = [base, perm, base, wave]
    evals = [(0, 0), (1, 0), (2, 0), (3, 0), (3, 1)]
    with make_engine(shape, toks) as eng:
        res = eng.evaluate(epde_b200.encode(eqs, evals))
    assert np.array_equal(res.coef[0], res.coef[2]) and res.fitness_l1[0] == res.fitness_l1[2]
    assert np.array_equal(res.sw_weights[0], res.sw_weights[2])
    assert_rel(res.coef[0, :2], res.coef[1, [1, 0]], 'term permutation', rtol=1e-9)
    assert_rel(res.coef[0, :2], [0.5, -2.0], 'planted coefficients', rtol=1e-6)
    assert_rel(res.intercept[0], 0.25, 'planted intercept', rtol=1e-6)
    assert res.fitness_l1[0] < 1e-9
    check_against_oracle(res, toks, shape, eqs, [(0, 0), (1, 0), (2, 0), (3, 0), (3, 1)], 'wave1024', min_checked=4)
