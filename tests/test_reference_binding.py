"""CPU tests that run only where the reference tree exists (the build container): the operator mirrors
are checked against the reference's REAL classes and objects -- same keys / tags / params contract,
and `encode_equation` walks genuine `Equation` / `Term` / `Factor` instances into an encoding whose
columns (evaluated by the oracle) equal `Equation.evaluate`'s.  Skipped on the GPU box."""
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

from oracle import ref_loader as rl  # noqa: E402

pytestmark = pytest.mark.skipif(not rl.reference_available(), reason='reference tree not present')


@pytest.fixture(scope='module')
def wave_search():
    rl.load_reference()
    d = np.loadtxt(os.path.join(rl.REFERENCE_ROOT, 'examples/data/wave_sln_80.csv'), delimiter=',').T
    t = np.linspace(0, 1, d.shape[0]); x = np.linspace(0, 1, d.shape[1])
    coords = np.meshgrid(t, x, indexing='ij')
    return rl.make_search(coords, d, 10, (2, 2))


def test_operator_contract_equals_reference():
    rl.load_reference()
    from epde.operators.common import fitness as rf, sparsity as rs, coeff_calculation as rc, right_part_selection as rr
    from epde_b200 import operators as ops
    pairs = [(ops.LASSOSparsity, rs.LASSOSparsity), (ops.LinRegBasedCoeffsEquation, rc.LinRegBasedCoeffsEquation),
             (ops.L2Fitness, rf.L2Fitness), (ops.L2LRFitness, rf.L2LRFitness),
             (ops.EqRightPartSelector, rr.EqRightPartSelector)]
    for mine, ref in pairs:
        assert mine.key == ref.key
        assert mine().operator_tags == ref().operator_tags, mine.__name__
        # params contract: keys must equal param_keys (template.py:49-55)
        a, b = mine(['penalty_coeff']), ref(['penalty_coeff'])
        a.params = {'penalty_coeff': 0.2}; b.params = {'penalty_coeff': 0.2}
        for op in (a, b):
            with pytest.raises(KeyError):
                op.params = {'other': 1}
    # a reference operator is accepted as suboperator of a mirror and vice versa (duck-typed containers)
    m = ops.L2LRFitness(['penalty_coeff'])
    m.set_suboperators({'sparsity': ops.LASSOSparsity(), 'coeff_calc': rc.LinRegBasedCoeffsEquation()})
    r = rf.L2LRFitness(['penalty_coeff'])
    r.set_suboperators({'sparsity': rs.LASSOSparsity(), 'coeff_calc': rc.LinRegBasedCoeffsEquation()})
    assert set(m.suboperators.keys()) == {'sparsity', 'coeff_calc'}


def test_encode_real_equations_matches_equation_evaluate(wave_search):
    """encode_equation on the reference's own Equation objects; oracle columns == Equation.evaluate."""
    from epde_b200 import operators as ops
    from oracle import epde_oracle as orc
    tokens = rl.cache_tokens()
    b = ops.Binding(engine=None)
    vecs = []
    for (label, params), vec in tokens.items():
        if params == (1.0,):
            assert b.token_id(label, create=True) == len(vecs)
            vecs.append(np.asarray(vec, dtype=np.float64))
    texts = ['1.0 * d^2u/dx1^2{power: 1} + 0. = d^2u/dx0^2{power: 1}',
             '1.0 * u{power: 2} * du/dx1{power: 1} + 1.0 * d^2u/dx1^2{power: 1} + 1.0 * u{power: 3} + 0. = du/dx0{power: 1}',
             '1.0 * du/dx0{power: 2} + 1.0 * u{power: 1} * d^2u/dx0^2{power: 1} + 0. = d^2u/dx1^2{power: 1}']
    for text in texts:
        _, eq = rl.translate(text, wave_search)
        eq.target_idx = len(eq.structure) - 1
        enc = ops.encode_equation(eq, b)
        assert len(enc) == len(eq.structure)
        for term, eterm in zip(eq.structure, enc):
            assert len(eterm) == len(term.structure)
            for f, ef in zip(term.structure, eterm):
                assert b.token_ids[f.label] == ef[0] and ef[2] == 0
        cols = orc.term_columns(vecs, enc)
        _, target, features = eq.evaluate(normalize=True, return_val=False)
        np.testing.assert_allclose(cols[:, eq.target_idx], target, rtol=1e-12, atol=0)
        feat_idx = [i for i in range(len(enc)) if i != eq.target_idx]
        np.testing.assert_allclose(cols[:, feat_idx], features, rtol=1e-12, atol=0)


def test_selector_structural_steps_on_real_equation(wave_search):
    """simplify_equation of the mirror == the reference's on a real equation whose kept terms share a factor."""
    rl.load_reference()
    from epde.operators.common.right_part_selection import EqRightPartSelector as RefSel
    from epde_b200.operators import EqRightPartSelector as MySel
    text = '1.0 * u{power: 2} * du/dx1{power: 1} + 1.0 * u{power: 1} * d^2u/dx1^2{power: 1} + 0. = u{power: 3} * du/dx0{power: 1}'
    outs = []
    for cls in (RefSel, MySel):
        _, eq = rl.translate(text, wave_search)
        eq.target_idx = len(eq.structure) - 1
        eq.weights_internal = np.array([0.5, -2.0])
        eq.weights_internal_evald = True
        changed = cls().simplify_equation(eq)
        outs.append((changed, [sorted((f.label, float(f.params[0])) for f in t.structure) for t in eq.structure]))
    assert outs[0] == outs[1]
    assert outs[0][0] is True       # the shared factor u was divided out


# ------------------------------------------------------------------------------------------
# the drop-in proper: mirrors under EPDE's own mappers / directors / EpdeSearch.fit
# (engine = tests/oracle_engine.OracleEngine, the CPU stand-in; the GPU suite runs the same operator code on
# the CUDA engine through tests/standins.py)
# ------------------------------------------------------------------------------------------
@pytest.fixture()
def integration():
    rl.load_reference()
    from epde_b200 import epde_integration as ei, operators as ops
    from oracle_engine import OracleEngine
    made = []

    def factory(shape):
        made.append(OracleEngine(shape))
        return made[-1]
    ei.set_engine_factory(factory)
    ops.unbind()
    yield ei, ops, made
    ei.set_engine_factory(None)
    ops.unbind()


def _wire(fit_cls, sp_cls, cc_cls, penalty=0.2):
    op = fit_cls(['penalty_coeff'])
    op.params = {'penalty_coeff': penalty}
    op.set_suboperators({'sparsity': sp_cls(), 'coeff_calc': cc_cls()})
    return op


def test_mirrors_pass_the_reference_type_checks(integration):
    """template.py:72-78 (issubclass check in set_suboperators) and operator_mappers.py:37-80."""
    ei, ops, _ = integration
    from epde.operators.utils.template import CompoundOperator as RefBase
    from epde.operators.utils.operator_mappers import map_operator_between_levels
    from epde.operators.common.right_part_selection import EqRightPartSelector as RefRPS
    from epde.operators.common.fitness import L2LRFitness as RefL2LR
    for cls in (ops.LASSOSparsity, ops.LinRegBasedCoeffsEquation, ops.L2Fitness, ops.L2LRFitness,
                ops.EqRightPartSelector, ops.RandomRHPSelector):
        assert issubclass(cls, RefBase), cls
    fit = _wire(ops.L2LRFitness, ops.LASSOSparsity, ops.LinRegBasedCoeffsEquation)
    mapped = map_operator_between_levels(fit, 'gene level', 'chromosome level')
    assert 'chromosome level' in mapped.operator_tags and 'gene level' not in mapped.operator_tags
    ref_rps = RefRPS()
    ref_rps.set_suboperators({'fitness_calculation': fit})          # a mirror inside a reference operator
    my_rps = ops.EqRightPartSelector()
    my_rps.set_suboperators({'fitness_calculation': _wire(RefL2LR, *_ref_sparsity_pair())})   # and vice versa
    map_operator_between_levels(my_rps, 'gene level', 'chromosome level')
    map_operator_between_levels(ops.LASSOSparsity(), 'gene level', 'chromosome level')


def _ref_sparsity_pair():
    from epde.operators.common.sparsity import LASSOSparsity
    from epde.operators.common.coeff_calculation import LinRegBasedCoeffsEquation
    return LASSOSparsity, LinRegBasedCoeffsEquation


def _eq_state(eq):
    return dict(target=int(eq.target_idx), n_terms=len(eq.structure),
                wi=np.array(eq.weights_internal, dtype=float), wf=np.array(eq.weights_final, dtype=float),
                fit=eq.fitness_value, stab=eq.coefficients_stability)


def _assert_same_state(a, b, tag=''):
    assert a['target'] == b['target'] and a['n_terms'] == b['n_terms'], (tag, a, b)
    assert np.array_equal(a['wi'] != 0, b['wi'] != 0), (tag, a['wi'], b['wi'])
    np.testing.assert_allclose(a['wi'], b['wi'], rtol=1e-6, atol=1e-12, err_msg=tag)
    np.testing.assert_allclose(a['wf'], b['wf'], rtol=1e-6, atol=1e-12, err_msg=tag)
    np.testing.assert_allclose(a['fit'], b['fit'], rtol=1e-6, err_msg=tag)
    np.testing.assert_allclose(a['stab'], b['stab'], rtol=1e-5, atol=1e-12, err_msg=tag)


def test_selector_and_fitness_equal_reference_on_real_equations(wave_search, integration):
    """mirror selector + in-place fitness == reference selector + fitness on deep copies of the same real
    Equation objects, incl. a copy made BETWEEN sparsity and fitness (the stash outside __slots__ is lost:
    ADVICE r1) and the selector scoring through a reference fitness suboperator (sequential path)."""
    import copy
    ei, ops, made = integration
    ref_l2lr, _, ref_rps = rl.make_operators()
    my_fit = _wire(ops.L2LRFitness, ops.LASSOSparsity, ops.LinRegBasedCoeffsEquation)
    my_rps = ops.EqRightPartSelector()
    my_rps.set_suboperators({'fitness_calculation': my_fit})
    texts = ['1.0 * d^2u/dx1^2{power: 1} + 1.0 * u{power: 1} + 0. = d^2u/dx0^2{power: 1}',
             '1.0 * u{power: 2} * du/dx1{power: 1} + 1.0 * d^2u/dx1^2{power: 1} + 1.0 * u{power: 3} + 0. = du/dx0{power: 1}',
             '1.0 * du/dx0{power: 1} + 1.0 * u{power: 1} * d^2u/dx0^2{power: 1} + 1.0 * du/dx1{power: 1} + 0. = d^2u/dx1^2{power: 1}']
    for text in texts:
        _, eq0 = rl.translate(text, wave_search)
        eq0.target_idx = len(eq0.structure) - 1
        a, b, c = copy.deepcopy(eq0), copy.deepcopy(eq0), copy.deepcopy(eq0)
        np.random.seed(3); ref_rps.apply(a, {}); ref_l2lr.apply(a, {})
        np.random.seed(3); my_rps.apply(b, {})
        b2 = copy.deepcopy(b)                    # drops _b200_stability, keeps _cached_sw_weights
        assert not hasattr(b2, '_b200_stability') and b2._cached_sw_weights is not None
        my_fit.apply(b, {}); my_fit.apply(b2, {})
        _assert_same_state(_eq_state(a), _eq_state(b), text)
        _assert_same_state(_eq_state(a), _eq_state(b2), text + ' (deep-copied)')
        # selector wired with the REFERENCE fitness: applied target by target like right_part_selection.py:65
        mixed = ops.EqRightPartSelector()
        mixed.set_suboperators({'fitness_calculation': ref_l2lr})
        np.random.seed(3); mixed.apply(c, {}); ref_l2lr.apply(c, {})
        _assert_same_state(_eq_state(a), _eq_state(c), text + ' (reference suboperator)')
    assert made and made[0].n_batches > 0


def test_selector_honours_an_l2_fitness_suboperator(wave_search, integration):
    """a director that wires L2Fitness (never None, penalty division) must get L2-ranked targets
    (right_part_selection.py:65 calls whatever suboperator is set).  The reference's own L2Fitness cannot run
    after its LASSOSparsity in this snapshot (fitness.py:78 indexes the weights_final LIST with a mask ->
    TypeError), so the expectation is built from the oracle values of every candidate target."""
    import copy
    ei, ops, made = integration
    from oracle import epde_oracle as orc
    my_l2 = _wire(ops.L2Fitness, ops.LASSOSparsity, ops.LinRegBasedCoeffsEquation)
    my_rps = ops.EqRightPartSelector()
    my_rps.set_suboperators({'fitness_calculation': my_l2})
    text = '1.0 * d^2u/dx1^2{power: 1} + 1.0 * u{power: 1} + 1.0 * du/dx0{power: 1} + 0. = d^2u/dx0^2{power: 1}'
    _, eq0 = rl.translate(text, wave_search)
    eq0.target_idx = len(eq0.structure) - 1
    b = copy.deepcopy(eq0)
    enc = ops.encode_equation(b)
    cand = [i for i, term in enumerate(b.structure) if term.contains_deriv('u')]
    eng = made[-1]
    rs = orc.evaluate_equation(eng._token_list(), eng.inner_shape, enc, targets=cand)
    l2 = [r.fitness_l2 / (0.2 if (np.sum(r.coef) + r.intercept) == 0 else 1.0) for r in rs]
    l1 = [np.inf if r.all_zero else r.fitness for r in rs]
    np.random.seed(5); my_rps.apply(b, {})
    chosen = eq0.structure[cand[int(np.argmin(l2))]]
    assert frozenset(f.label for f in b.structure[b.target_idx].structure) == frozenset(f.label for f in chosen.structure)
    assert len(l1) == len(l2)


def test_director_runs_epde_search_fit(integration):
    """EpdeSearch(..., use_default_strategy=False, director=B200MOEADDDirector).fit on the wave data, then every
    equation of the final population is re-evaluated with the REFERENCE operators: same fitness / stability."""
    import copy
    ei, ops, made = integration
    from epde.interface.interface import EpdeSearch
    d = np.loadtxt(os.path.join(rl.REFERENCE_ROOT, 'examples/data/wave_sln_80.csv'), delimiter=',').T
    t = np.linspace(0, 1, d.shape[0]); x = np.linspace(0, 1, d.shape[1])
    grids = np.meshgrid(t, x, indexing='ij')
    np.random.seed(11)
    director = ei.make_moeadd_director()
    s = EpdeSearch(multiobjective_mode=True, use_default_strategy=False, director=director, use_solver=False,
                   use_pic=True, boundary=10, coordinate_tensors=tuple(grids), device='cpu',
                   verbose_params={'show_iter_idx': False})
    s.set_preprocessor(default_preprocessor_type='FD', preprocessor_kwargs={})
    s.set_moeadd_params(population_size=4, training_epochs=1)
    s.fit(data=d, variable_names=['u'], max_deriv_order=(2, 2), equation_terms_max_number=4, data_fun_pow=2,
          additional_tokens=[], equation_factors_max_number={'factors_num': [1, 2], 'probas': [0.7, 0.3]},
          eq_sparsity_interval=(1e-6, 1e-4))
    assert made and made[-1].n_batches > 0 and made[-1].n_evals > 0          # the engine did the numeric work
    ref_l2lr, _, _ = rl.make_operators()
    checked = 0
    for level in s.optimizer.pareto_levels.levels:
        for soeq in level:
            for eq in soeq:
                assert eq.fitness_calculated and eq.right_part_selected
                ref = copy.deepcopy(eq)
                ref.fitness_calculated = False
                ref._cached_sw_weights = None            # reference recomputes the windows (fitness.py:158-159)
                ref_l2lr.apply(ref, {})
                np.testing.assert_allclose(eq.fitness_value, ref.fitness_value, rtol=1e-6)
                np.testing.assert_allclose(eq.coefficients_stability, ref.coefficients_stability, rtol=1e-5, atol=1e-12)
                checked += 1
    assert checked >= 4


def test_device_tensor_cache_as_global_cache_and_persistence(integration):
    """with install_device_tensor_cache() EPDE's own `global_var.tensor_cache` keeps its
    base tensors in the engine only (no numpy copies), every consumer of the Cache API still works (create_pool,
    Factor.evaluate through Equation.evaluate, the operator mirrors, a short MOEADD fit), and the cache survives a
    pickle round trip of its host form (what EPDELoader stores) with a re-upload."""
    import copy
    import pickle
    ei, ops, made = integration
    import epde.globals as gv
    from epde.interface.interface import EpdeSearch
    undo = ei.install_device_tensor_cache()
    try:
        d = np.loadtxt(os.path.join(rl.REFERENCE_ROOT, 'examples/data/wave_sln_80.csv'), delimiter=',').T
        t = np.linspace(0, 1, d.shape[0]); x = np.linspace(0, 1, d.shape[1])
        grids = np.meshgrid(t, x, indexing='ij')
        np.random.seed(3)
        s = EpdeSearch(multiobjective_mode=True, use_default_strategy=False, director=ei.make_moeadd_director(),
                       use_solver=False, use_pic=True, boundary=10, coordinate_tensors=tuple(grids), device='cpu',
                       verbose_params={'show_iter_idx': False})
        s.set_preprocessor(default_preprocessor_type='FD', preprocessor_kwargs={})
        s.create_pool(data=d, variable_names=['u'], max_deriv_order=(2, 2), additional_tokens=[])
        tc = gv.tensor_cache
        assert isinstance(tc, ei.DeviceTensorCache)
        store = tc.memory_default['numpy']
        n_pts = int(np.prod(gv.grid_cache.inner_shape))
        assert len(store._ids) >= 5 and not any(np.asarray(v).size == n_pts for v in store._host.values())
        eng = made[-1]
        assert tc.consumed_memory == 8 * n_pts * len(store._ids) == eng.consumed_memory
        # reference-side consumers read through the same object
        key = ('du/dx0', (1.0,))
        assert (key, False) in tc and key in tc
        ref_vec = np.gradient(d, t, axis=0)[10:-10, 10:-10].reshape(-1)
        np.testing.assert_allclose(tc.get(key), ref_vec, rtol=1e-12, atol=1e-14)
        _, eq = rl.translate('1.0 * u{power: 2} * du/dx1{power: 1} + 1.0 * d^2u/dx1^2{power: 1} + 0. = d^2u/dx0^2{power: 1}', s)
        eq.target_idx = 2
        _, target, feats = eq.evaluate(normalize=True, return_val=False)
        u_in = d[10:-10, 10:-10].reshape(-1)
        np.testing.assert_allclose(feats[:, 0], u_in ** 2 * np.gradient(d, x, axis=1)[10:-10, 10:-10].reshape(-1), rtol=1e-12)
        # the mirrors use the tokens that are already resident (no second upload of the same label)
        before = len(eng.tokens)
        fit = _wire(ops.L2LRFitness, ops.LASSOSparsity, ops.LinRegBasedCoeffsEquation)
        val = fit.apply(eq, {}, force_out_of_place=True)
        ref_l2lr, _, _ = rl.make_operators()
        eq2 = copy.deepcopy(eq)
        ref_val = ref_l2lr.apply(eq2, {}, force_out_of_place=True)
        assert (val is None) == (ref_val is None)
        if val is not None:
            np.testing.assert_allclose(val, ref_val, rtol=1e-6)
        assert len([k for k in eng.tokens]) <= before + 0
        # persistence: host form -> pickle -> restore -> re-upload
        blob = pickle.dumps(tc.to_host_cache())
        host = pickle.loads(blob)
        assert type(host).__name__ == 'Cache' and isinstance(host.memory_default['numpy'], dict)
        restored = ei.DeviceTensorCache.from_host_cache(host)
        np.testing.assert_array_equal(restored.get(key), tc.get(key))
        assert set(restored.memory_default['numpy'].keys()) == set(store.keys())
        # and a short search on top of the device cache
        s.set_moeadd_params(population_size=3, training_epochs=1)
        s.fit(data=d, variable_names=['u'], max_deriv_order=(2, 2), equation_terms_max_number=3, data_fun_pow=2,
              additional_tokens=[], equation_factors_max_number={'factors_num': [1, 2], 'probas': [0.8, 0.2]},
              eq_sparsity_interval=(1e-6, 1e-4))
        assert all(eq.fitness_calculated for lvl in s.optimizer.pareto_levels.levels for soeq in lvl for eq in soeq)
    finally:
        undo()
