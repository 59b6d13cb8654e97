"""TEST INFRASTRUCTURE ONLY -- loader for the real (Python) EPDE reference.

Only ``oracle/make_golden.py`` and CPU tests that are skipped when
``/root/reference`` is absent may import this module.  It never travels to the
GPU box (the reference tree does not exist there) and no product code under
``epde_b200/`` imports it.

What it does (recipe of SURVEY.md Appendix A):
  * registers stub modules for matplotlib / seaborn (not installed; the reference
    imports them at module top level: fitness.py:13-14, sparsity.py:16,
    supplementary.py:18);
  * puts ``/root/reference`` on ``sys.path`` and imports ``epde``;
  * offers helpers that build a token pool from explicit arrays and wire the
    reference's own operators (LASSOSparsity / L2LRFitness / EqRightPartSelector)
    the way ``projects/pic/data/wave/wave.py:73-86`` does.
"""
import os
import sys
import warnings
from unittest import mock

REFERENCE_ROOT = os.environ.get('EPDE_REFERENCE_ROOT', '/root/reference')


def reference_available() -> bool:
    return os.path.isdir(os.path.join(REFERENCE_ROOT, 'epde'))


def load_reference():
    """Import the unmodified reference package and return the module."""
    if not reference_available():
        raise RuntimeError(f'EPDE reference tree not found at {REFERENCE_ROOT}')
    for name in ('matplotlib', 'matplotlib.pyplot', 'matplotlib.cm', 'matplotlib.colors',
                 'matplotlib.ticker', 'matplotlib.animation', 'mpl_toolkits',
                 'mpl_toolkits.mplot3d', 'seaborn'):
        if name not in sys.modules:
            sys.modules[name] = mock.MagicMock()
    if REFERENCE_ROOT not in sys.path:
        sys.path.insert(0, REFERENCE_ROOT)
    warnings.filterwarnings('ignore', category=SyntaxWarning)
    warnings.filterwarnings('ignore', category=DeprecationWarning)
    import epde  # noqa: F401
    return epde


def make_search(coordinate_tensors, data, boundary, max_deriv_order, derivs=None,
                additional_tokens=None, preproc='FD', preproc_kwargs=None):
    """EpdeSearch + pool from arrays; FD derivatives (or the ``preproc`` preprocessor, e.g. 'poly' = per-point
    Chebyshev fits, preprocessing/deriv_calculators.py:111-134) unless ``derivs`` is given."""
    load_reference()
    from epde.interface.interface import EpdeSearch
    s = EpdeSearch(use_solver=False, use_pic=True, boundary=boundary,
                   coordinate_tensors=tuple(coordinate_tensors), device='cpu',
                   verbose_params={'show_iter_idx': False})
    s.set_preprocessor(default_preprocessor_type=preproc, preprocessor_kwargs=preproc_kwargs or {})
    kwargs = dict(data=data, variable_names=['u'], max_deriv_order=tuple(max_deriv_order),
                  additional_tokens=additional_tokens or [])
    if derivs is not None:
        kwargs['derivs'] = [derivs]
    s.create_pool(**kwargs)
    return s


def make_operators(penalty_coeff=0.2):
    """(fitness L2LR, fitness L2, right-part selector) wired like the reference's directors
    (optimizers/moeadd/strategy.py:46-63)."""
    load_reference()
    from epde.operators.common.fitness import L2LRFitness, L2Fitness
    from epde.operators.common.sparsity import LASSOSparsity
    from epde.operators.common.coeff_calculation import LinRegBasedCoeffsEquation
    from epde.operators.common.right_part_selection import EqRightPartSelector

    def wire(cls):
        op = cls(['penalty_coeff'])
        op.params = {'penalty_coeff': penalty_coeff}
        op.set_suboperators({'sparsity': LASSOSparsity(), 'coeff_calc': LinRegBasedCoeffsEquation()})
        return op

    l2lr = wire(L2LRFitness)
    l2 = wire(L2Fitness)
    rps = EqRightPartSelector()
    rps.set_suboperators({'fitness_calculation': l2lr})
    return l2lr, l2, rps


def translate(text, search):
    """text -> Equation of variable 'u' (interface/equation_translator.py:25)."""
    from epde.interface.equation_translator import translate_equation
    soeq = translate_equation(text, search.pool, all_vars=['u'])
    eq = soeq.vals['u']
    eq.main_var_to_explain = 'u'
    return soeq, eq


def cache_tokens():
    """{label: (N,) float64} for every power-1 token in the reference's tensor cache."""
    import epde.globals as gv
    out = {}
    for key, val in gv.tensor_cache.memory_default['numpy'].items():
        if isinstance(key, tuple) and len(key) == 2 and isinstance(key[0], str):
            out[key] = val
    return out
