"""The numpy oracle (oracle/epde_oracle.py) replayed against vectors produced by the REAL
reference (tests/golden/*.npz, written by oracle/make_golden.py).  CPU only."""
import numpy as np
import pytest

from conftest import EQ_FIXTURES, fixture_equations, load_fixture
from helpers import assert_rel, assert_sw, natural_scales
from oracle import epde_oracle as orc


COND_MAX = 1e9      # see oracle.window_condition: beyond this the reference itself is rounding noise
ILL = ('collinear', 'duplicate', 'nearconst')


def _micro_cases():
    z = load_fixture('micro')
    return z, int(z['n_cases'])


@pytest.mark.parametrize('ci', range(9))
def test_calculate_weights_micro(ci):
    z, n = _micro_cases()
    assert ci < n
    X, y, shape = z[f'c{ci}_X'], z[f'c{ci}_y'], z[f'c{ci}_shape']
    kind = str(z[f'c{ci}_kind'])
    g = np.ones(X.shape[0])
    direct = orc.calculate_weights_direct(X, y, g, shape)
    plan = orc.WindowPlan(shape)
    cols = np.column_stack([X, y])
    win = orc.window_grams(orc.segment_grams(cols, plan), plan)
    F = X.shape[1]
    gram = orc.calc_weights_gram(win, list(range(F)) + [F + 1], F)
    if kind in ILL:
        # singular up to the 1e-6 ridge: individual coefficients are not determined to 1e-6 by
        # any summation order (LAPACK's included); the fitted values are.
        Xa = np.column_stack([X, np.ones(X.shape[0])])
        ref_pred = Xa @ z[f'c{ci}_cw'].T
        scale = np.abs(ref_pred).max()
        for w in (direct, gram):
            assert np.abs(Xa @ w.T - ref_pred).max() <= 1e-5 * scale, kind
        return
    assert_rel(direct, z[f'c{ci}_cw'], 'direct calculate_weights', rtol=1e-8, atol=1e-12)
    assert_rel(gram, z[f'c{ci}_cw'], 'gram calculate_weights', rtol=1e-6, atol=1e-10)


@pytest.mark.parametrize('ci', range(9))
def test_lasso_micro(ci):
    z, _ = _micro_cases()
    X, y, shape = z[f'c{ci}_X'], z[f'c{ci}_y'], z[f'c{ci}_shape']
    kind = str(z[f'c{ci}_kind'])
    g = np.ones(X.shape[0])
    ref_coef, ref_b, ref_cached = z[f'c{ci}_coef'], float(z[f'c{ci}_b']), z[f'c{ci}_cached']
    plan = orc.WindowPlan(shape)
    cols = np.column_stack([X, y])
    seg = orc.segment_grams(cols, plan)
    F = X.shape[1]
    coef, b, cached = orc.lasso_fit_gram(orc.window_grams(seg, plan), seg[0].sum(axis=0), list(range(F)), F)
    assert np.array_equal(coef != 0, ref_coef != 0), f'{kind}: selected set differs'
    if kind in ILL:
        Xa = np.column_stack([X, np.ones(X.shape[0])])
        ref_pred = Xa @ np.append(ref_coef, ref_b)
        assert np.abs(Xa @ np.append(coef, b) - ref_pred).max() <= 1e-5 * np.abs(ref_pred).max(), kind
        return
    assert_rel(coef, ref_coef, 'coef', rtol=1e-6, atol=1e-12)
    assert_rel(b, ref_b, 'intercept', rtol=1e-6, atol=1e-10)
    assert_rel(cached, ref_cached, 'cached weights', rtol=1e-6, atol=1e-10)
    if X.shape[0] <= 4000:
        coef_d, b_d, _ = orc.lasso_fit_direct(X, y, g, shape)
        assert np.array_equal(coef_d != 0, ref_coef != 0)
        assert_rel(coef_d, ref_coef, 'direct coef', rtol=1e-8, atol=1e-13)


@pytest.mark.parametrize('name', EQ_FIXTURES)
def test_equation_evals(name):
    fx = load_fixture(name)
    eqs = fixture_equations(fx)
    tokens = fx['tokens']
    shape = fx['shape']
    ndim = len(shape)
    n_ill = 0
    for e, eq in enumerate(eqs):
        sel = np.nonzero(fx['ev_eq'] == e)[0]
        if sel.size == 0:
            continue
        res = orc.evaluate_equation(tokens, shape, eq, targets=[int(fx['ev_tgt'][v]) for v in sel])
        for v, r in zip(sel, res):
            T = len(eq)
            ref_coef = fx['ev_coef'][v, :T - 1]
            if r.cond > COND_MAX:
                n_ill += 1
                continue
            assert np.array_equal(r.coef != 0, ref_coef != 0), f'{name} eval {v}: selected set differs'
            assert_rel(r.coef, ref_coef, f'{name} eval {v} coef', atol=1e-14)
            assert_rel(r.intercept, fx['ev_b'][v], f'{name} eval {v} intercept', atol=1e-12)
            nact = int(fx['ev_nact'][v])
            assert r.sw_weights.shape[1] == nact + 1
            cols = orc.term_columns(tokens, eq)
            sc = natural_scales(np.delete(cols, r.target, axis=1), cols[:, r.target])
            assert_sw(r.sw_weights, fx['ev_sw'][v, :, :nact + 1], f'{name} eval {v} sw',
                      col_scales=sc[np.append(ref_coef != 0, True)])
            if np.isnan(fx['ev_fit'][v]):
                assert r.all_zero
            else:
                assert not r.all_zero
                assert_rel(r.fitness, fx['ev_fit'][v], f'{name} eval {v} fitness')
            assert_rel(r.stability, fx['ev_stab'][v], f'{name} eval {v} stability', rtol=1e-5, atol=1e-14)
            if not np.isnan(fx['ev_fit_l2'][v]):
                assert_rel(r.fitness_l2, fx['ev_fit_l2'][v], f'{name} eval {v} l2')
    assert n_ill <= 0.25 * len(fx['ev_eq']), f'{n_ill} ill-posed evals'


@pytest.mark.parametrize('name', EQ_FIXTURES)
def test_right_part_selection(name):
    """argmin-fitness target of the reference's EqRightPartSelector on equations where the
    selector finished without re-randomising terms (rps_ok)."""
    fx = load_fixture(name)
    eqs = fixture_equations(fx)
    checked = 0
    for e, eq in enumerate(eqs):
        if not fx['rps_ok'][e]:
            continue
        sel = np.nonzero(fx['ev_eq'] == e)[0]
        res = orc.evaluate_equation(fx['tokens'], fx['shape'], eq, targets=[int(fx['ev_tgt'][v]) for v in sel])
        best = orc.select_right_part(res)
        if best is None or any(r.cond > COND_MAX for r in res):
            continue
        # the reference may loop (simplification) and re-select; only first-pass equivalents are
        # comparable: kept terms == target + nonzero features of the chosen eval
        r = res[best]
        keep = np.zeros(8, dtype=np.int8)
        keep[r.target] = 1
        feat = [i for i in range(len(eq)) if i != r.target]
        for j, c in zip(feat, r.coef):
            if c != 0:
                keep[j] = 1
        if not np.array_equal(keep, fx['rps_keep'][e]):
            continue
        assert r.target == int(fx['rps_target'][e])
        assert_rel(r.fitness, fx['rps_fit'][e], f'{name} eq {e} rps fitness')
        checked += 1
    assert checked >= 1


# ------------------------------------------------------------------------------------------
# round 2 fixtures (oracle/make_golden.py micro2_fixtures): wide systems (T = 12, 16, 30) and non-trivial
# sample weights, generated by the unmodified reference
# ------------------------------------------------------------------------------------------
N_MICRO2 = 8


@pytest.mark.parametrize('ci', range(N_MICRO2))
def test_micro2_wide_and_weighted(ci):
    z = load_fixture('micro2')
    assert int(z['n_cases']) == N_MICRO2
    X, y, g, shape = z[f'c{ci}_X'], z[f'c{ci}_y'], z[f'c{ci}_g'], z[f'c{ci}_shape']
    kind = str(z[f'c{ci}_kind'])
    F = X.shape[1]
    plan = orc.WindowPlan(shape)
    cols = np.column_stack([X, y])
    weighted = kind == 'weighted'
    seg_w = orc.segment_grams(cols, plan, g if weighted else None)
    win = orc.window_grams(seg_w, plan)
    total = orc.segment_grams(cols, plan, None)[0].sum(axis=0) if weighted else seg_w[0].sum(axis=0)
    gram = orc.calc_weights_gram(win, list(range(F)) + [F + 1], F)
    assert_rel(gram, z[f'c{ci}_cw'], f'{kind} calculate_weights (gram form)', rtol=1e-6, atol=1e-9)
    coef, b, cached = orc.lasso_fit_gram(win, total, list(range(F)), F)
    ref_coef = z[f'c{ci}_coef']
    assert np.array_equal(coef != 0, ref_coef != 0), f'{kind} F={F}: selected set differs'
    assert_rel(coef, ref_coef, 'coef', rtol=1e-6, atol=1e-11)
    assert_rel(b, float(z[f'c{ci}_b']), 'intercept', rtol=1e-6, atol=1e-9)
    assert_rel(cached, z[f'c{ci}_cached'], 'cached weights', rtol=1e-6, atol=1e-9)


def test_procedural_tokens_fixture():
    """sin / cos / coordinate-power tokens (CustomEvaluator families, evaluators.py:141,150): oracle == reference."""
    from conftest import fixture_equations_procedural
    fx = load_fixture('procedural')
    shape = tuple(int(s) for s in fx['shape'])
    eqs = fixture_equations_procedural(fx)
    assert any(f[2] in (1, 2) for eq in eqs for term in eq for f in term)
    checked = 0
    for v in range(len(fx['ev_eq'])):
        eq = eqs[int(fx['ev_eq'][v])]
        o = orc.evaluate_equation(fx['tokens'], shape, eq, targets=[int(fx['ev_tgt'][v])])[0]
        if o.cond > COND_MAX:
            continue
        F = len(eq) - 1
        ref = fx['ev_coef'][v, :F]
        assert np.array_equal(o.coef != 0, ref != 0), f'eval {v}: selected set'
        assert_rel(o.coef, ref, f'eval {v} coef', rtol=1e-6, atol=1e-9 * max(1.0, np.abs(ref).max()))
        if np.isnan(fx['ev_fit'][v]):
            assert o.all_zero
        else:
            assert_rel(o.fitness, fx['ev_fit'][v], f'eval {v} fitness', rtol=1e-6)
        checked += 1
    assert checked >= 8


def test_poly_derivative_restatement_matches_reference_fixture():
    """oracle.poly_derivative_points (restated cheb.py / fin_der.py) == the derivative tokens the reference's 'poly'
    preprocessor produced for the noisy Burgers field (a strip of the grid: the per-point fits are slow on the host),
    and the linear-functional form the device producer uses (producers.chebyshev_window_weights) == both."""
    from epde_b200.producers import chebyshev_window_weights
    fx = load_fixture('burgers_poly_b50')
    u, t, x, b = fx['data_full'], fx['axis0'], fx['axis1'], int(fx['boundary'])
    labels = [str(l) for l in fx['labels']]
    inner = (u.shape[0] - 2 * b, u.shape[1] - 2 * b)
    cols = slice(b, b + 6)                                   # 6 columns of the interior
    for lab, axis, m, nder in (('du/dx0', 0, 1, 1), ('d^2u/dx1^2', 1, 2, 2)):
        ref = fx['tokens'][labels.index(lab)].reshape(inner)
        if axis == 0:
            got = orc.poly_derivative_points(u[:, cols], t, 0, m, nder)[b:-b, :]
            want = ref[:, :6]
        else:
            got = orc.poly_derivative_points(u[cols, :], x, 1, m, nder)[:, b:-b]
            want = ref[:6, :]
        assert_rel(got, want, f'{lab} restatement', rtol=1e-10, atol=1e-10 * np.abs(want).max())
        W = chebyshev_window_weights([t, x][axis], m, 9, n_der=nder)
        i = b + 3
        win = u[i - 4:i + 5, b] if axis == 0 else u[b, i - 4:i + 5]
        assert_rel(float(W[i] @ win), ref[3, 0] if axis == 0 else ref[0, 3], f'{lab} functional', rtol=1e-9,
                   atol=1e-10 * np.abs(want).max())
