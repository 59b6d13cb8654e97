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
