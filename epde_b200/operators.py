"""Drop-in mirrors of the reference's evaluator operators and tensor cache, backed by the B200 engine.

Same class names, ``key``s, constructor / ``params`` / ``set_suboperators`` / ``apply`` signatures,
tags, return conventions and ``Equation`` mutations as the reference:

    CompoundOperator            epde/operators/utils/template.py:26-170
    LASSOSparsity               epde/operators/common/sparsity.py:153-215      key 'LASSOBasedSparsity'
    LinRegBasedCoeffsEquation   epde/operators/common/coeff_calculation.py:16-47  key 'LinRegCoeffCalc'
    L2Fitness                   epde/operators/common/fitness.py:27-99         key 'DiscrepancyBasedFitness'
    L2LRFitness                 epde/operators/common/fitness.py:102-183       key 'DiscrepancyBasedFitnessWithCV'
    EqRightPartSelector         epde/operators/common/right_part_selection.py:19-152
                                                                         key 'FitnessCheckingRightPartSelector'
    RandomRHPSelector           epde/operators/common/right_part_selection.py:154-207
                                                                         key 'RandomRightPartSelector'
    Cache                       epde/cache/cache.py:141-524 (the subset the evaluator path touches)

**Base class.**  When the ``epde`` package is importable the mirrors DERIVE FROM EPDE's own
``CompoundOperator`` (template.py:26), so ``set_suboperators`` (template.py:72-78, an ``issubclass``
check), ``map_operator_between_levels`` / ``OperatorMapper`` (operator_mappers.py:37-80) and the
directors (moeadd/strategy.py:31-96) accept them unchanged.  Where EPDE is not installed (the GPU test
box) a standalone base with the same contract is used.  :func:`install_operator_classes` rebuilds the
classes against a given base (used when EPDE becomes importable after this module was imported).

The objects handed to ``apply`` are the reference's own ``Equation`` / ``Term`` / ``Factor`` instances
(duck-typed: only the attributes named in :func:`encode_equation` and in the operators are read).  All
arithmetic runs in libepde_b200.so; this module packs encodings and writes results back.  No CPU
fallback: without a bound engine every ``apply`` raises.

B200-first difference: the reference evaluates the D admissible targets of an equation one after the
other (right_part_selection.py:60-74); here they are ONE batch (one Gram pass serves all D targets),
and :meth:`EqRightPartSelector.apply_population` batches a whole population the same way
(SURVEY.md section 8f item 1).
"""
from __future__ import annotations

import numpy as np

from . import engine as _engine

# ------------------------------------------------------------------------------------------
# binding: which engine / token numbering the operators use (the reference keeps the equivalent in
# module globals: epde/globals.py:26-42 tensor_cache / grid_cache)
# ------------------------------------------------------------------------------------------
_TRIG_KIND = {'sin': _engine.KIND_SIN, 'cos': _engine.KIND_COS}


class Binding:
    """Engine + label -> token id table.  ``grid_label(dim)`` names the coordinate token of axis
    ``dim`` (the reference's grid cache keys them '0', '1', ..., cache.py:66-84).

    ``source``: optional callable ``label -> (N,) float64 vector or None``; a label that is not in
    the table yet is fetched through it and uploaded to HBM on first use (this is how the operators
    find the tensors EPDE's preprocessing put into ``global_var.tensor_cache``)."""

    def __init__(self, engine, source=None):
        self.engine = engine
        self.token_ids = {}
        self.families = {}        # factor label -> kind of a procedural family
        self.source = source

    # -- tokens
    def token_id(self, label, create=False):
        if label not in self.token_ids:
            if not create:
                vec = self.source(label) if self.source is not None else None
                if vec is None:
                    raise KeyError(f'token {label!r} is not in the B200 cache')
                self.token_ids[label] = len(self.token_ids)
                self.engine.upload(self.token_ids[label], np.asarray(vec, dtype=np.float64).reshape(-1))
                return self.token_ids[label]
            self.token_ids[label] = len(self.token_ids)
        return self.token_ids[label]

    @staticmethod
    def grid_label(dim):
        return ('grid', int(dim))

    def register_family(self, label, kind):
        """procedural factor family evaluated in-kernel on the coordinate token of its ``dim``
        parameter: kind KIND_SIN / KIND_COS (evaluators.py:141), KIND_POWER for grid tokens
        (evaluators.py:150), KIND_INVERSE (evaluators.py:147)."""
        self.families[label] = int(kind)


_binding: Binding | None = None
_auto_binder = None          # callable () -> Binding | None, installed by epde_b200.epde_integration


def bind(engine, source=None) -> Binding:
    """Install the engine the operators use (call where the reference calls globals.init_caches)."""
    global _binding
    _binding = Binding(engine, source)
    for lab, kind in _TRIG_KIND.items():
        _binding.register_family(lab, kind)
    return _binding


def binding() -> Binding:
    global _binding
    if _binding is None and _auto_binder is not None:
        _auto_binder()              # binds from EPDE's module globals (epde_integration.bind_from_epde_globals)
    if _binding is None:
        raise RuntimeError('epde_b200.operators: no engine bound (call epde_b200.operators.bind(engine)); '
                           'this path has no CPU fallback')
    return _binding


def unbind():
    global _binding
    _binding = None


def set_auto_binder(fn):
    global _auto_binder
    _auto_binder = fn


# ------------------------------------------------------------------------------------------
# Cache mirror (cache.py:141-524): same method names / return conventions for the calls the
# evaluator path makes; tensors live in HBM, nothing is kept on the host.
# (epde_b200.epde_integration.DeviceTensorCache is the variant installable as global_var.tensor_cache.)
# ------------------------------------------------------------------------------------------
class Cache:
    def __init__(self, bound: Binding | None = None):
        self._b = bound or binding()

    @property
    def inner_shape(self):                       # cache.py:250-270
        return self._b.engine.inner_shape

    @property
    def consumed_memory(self):                   # cache.py:305-330
        return self._b.engine.consumed_memory

    def add(self, label, tensor, normalized=False, structural=False, deriv_code=None) -> bool:
        """Cache.add (cache.py:333-395).  Returns True when stored.  The reference refuses (False)
        when its tensor-count budget is full; here the only limit is HBM and exhausting it raises."""
        if normalized or structural:
            raise NotImplementedError('only the base (non-normalised, non-structural) tensors are on the '
                                      'evaluator path (main_structures.py:593,606 pass structural=False)')
        vec = np.asarray(tensor, dtype=np.float64).reshape(-1)
        if vec.size != self._b.engine.n_points:
            raise ValueError(f'tensor of {vec.size} values does not match inner shape {self.inner_shape}')
        self._b.engine.upload(self._b.token_id(_base_label(label), create=True), vec)
        return True

    def get(self, label, normalized=False, structural=False, saved_as=None, torch_mode=False, deriv_code=None):
        """Cache.get (cache.py:397-452): host copy of the stored vector; a power != 1 key
        (supplementary.py:216-220) is served as base ** power like simple_function_evaluator
        (evaluators.py:128-135)."""
        if torch_mode:
            raise NotImplementedError('torch_mode tensors are not part of the evaluator path')
        key, power = _split_label(label)
        if key not in self._b.token_ids:
            raise KeyError(f'Can not get tensor from cache with label {label}')
        vec = self._b.engine.download(self._b.token_ids[key])
        return vec if power == 1.0 else vec ** power

    def __contains__(self, obj):                 # cache.py:454-470: (label, structural) or label
        label = obj[0] if (isinstance(obj, tuple) and len(obj) == 2 and isinstance(obj[1], bool)) else obj
        key, _ = _split_label(label)
        return key in self._b.token_ids and self._b.token_ids[key] in self._b.engine

    def delete_entry(self, entry_label):         # cache.py:472-490
        key = _base_label(entry_label)
        if key not in self._b.token_ids:
            raise ValueError('deleted element already not in memory')
        self._b.engine.delete(self._b.token_ids[key])

    def get_all(self, normalized=False, structural=False, mode='numpy'):
        keys = list(self._b.token_ids)
        return keys, [self._b.engine.download(self._b.token_ids[k]) for k in keys]

    def upload_grid(self, dim, values):
        """coordinate tensor of axis ``dim`` restricted to the interior (upload_grids, cache.py:66-84,
        then the g_func mask of evaluators.py:87)."""
        return self.add(Binding.grid_label(dim), values)


def _split_label(label):
    """reference cache keys are (label, (power, ...)) (supplementary.py:216-220); the HBM table holds
    the power-1 vectors and applies powers in-kernel -> (base key, power)."""
    if isinstance(label, tuple) and len(label) == 2 and isinstance(label[1], tuple) and \
            not (label[0] == 'grid'):
        power = float(label[1][0]) if len(label[1]) >= 1 else 1.0
        return label[0], power
    return label, 1.0


def _base_label(label):
    key, power = _split_label(label)
    if power != 1.0:
        raise KeyError(f'{label}: only power-1 tensors are stored; powers are applied in-kernel')
    return key


# ------------------------------------------------------------------------------------------
# Equation -> encoding (structure/factor.py:218-269, evaluators.py:47-151)
# ------------------------------------------------------------------------------------------
def _param(factor, name, default=None):
    for idx, descr in factor.params_description.items():
        if descr['name'] == name:
            return factor.params[idx]
    if default is None:
        raise KeyError(f'factor {factor.label!r} has no parameter {name!r}')
    return default


def encode_factor(factor, bound: Binding):
    """(token id, power, kind, param) of a reference ``Factor`` (reads .label, .params,
    .params_description)."""
    power = float(_param(factor, 'power', 1.0))
    fam = bound.families.get(factor.label)
    if fam is None and getattr(factor, 'ftype', None) == 'grids':
        fam = _engine.KIND_POWER         # GridTokens (prepared_tokens.py:312-356): coordinate of axis `dim` ** power
    if fam is None:
        return (bound.token_id(factor.label), power, _engine.KIND_POWER, 0.0)
    dim = int(_param(factor, 'dim'))
    tok = bound.token_id(Binding.grid_label(dim))
    if fam in (_engine.KIND_SIN, _engine.KIND_COS):
        return (tok, power, fam, float(_param(factor, 'freq')))
    return (tok, power, fam, 0.0)


def encode_equation(objective, bound: Binding | None = None):
    """nested encoding [[(token, power, kind, param), ...] per term] of ``objective.structure``
    (Equation.structure -> Term.structure -> Factor; main_structures.py:207-235)."""
    bound = bound or binding()
    return [[encode_factor(f, bound) for f in term.structure] for term in objective.structure]


def _describe(objective):
    try:
        return ' + '.join(' * '.join(f'{f.label}^{_param(f, "power", 1.0):g}' for f in t.structure)
                          for t in objective.structure)
    except Exception:       # noqa: BLE001
        return repr(objective)


def _check_limits(objective, eq):
    """clear error (naming the equation) before the engine rejects it deep inside a batch."""
    if len(eq) > _engine.MAX_TERMS:
        raise ValueError(f'equation with {len(eq)} terms exceeds the engine limit EPDE_MAX_TERMS = '
                         f'{_engine.MAX_TERMS}: {_describe(objective)}')
    for term in eq:
        if len(term) > _engine.MAX_FACTORS:
            raise ValueError(f'term with {len(term)} factors exceeds the engine limit EPDE_MAX_FACTORS = '
                             f'{_engine.MAX_FACTORS}: {_describe(objective)}')


def _evaluate(objectives, targets_per_eq, given=None, bound=None):
    """one engine batch over several equations; targets_per_eq[i] = list of target term indices."""
    bound = bound or binding()
    eqs = [encode_equation(o, bound) for o in objectives]
    for o, eq in zip(objectives, eqs):
        _check_limits(o, eq)
    evals = [(e, int(t)) for e, ts in enumerate(targets_per_eq) for t in ts]
    enc = _engine.encode(eqs, evals)
    if given is None:
        return bound.engine.evaluate(enc), evals
    mode, wterm, b = given
    with bound.engine.batch(enc) as bt:
        bt.set_given(mode, wterm, b)
        bt.gram(); bt.solve(); bt.residual()
        return bt.fetch(), evals


def _write_sparsity(objective, res, v):
    """the Equation mutations of LASSOSparsity.apply (sparsity.py:205-211)."""
    F = len(objective.structure) - 1
    coef = np.array(res.coef[v, :F], dtype=np.float64)
    objective.weights_internal = coef
    objective.weights_internal_evald = True
    full = np.append(coef, res.intercept[v])
    objective.weights_final = [w for w in full if w != 0]
    objective.weights_final_evald = True
    objective._cached_sw_weights = np.array(res.cached_weights(v), dtype=np.float64)
    objective._eval_cache = {}
    # not a reference attribute: the stability of exactly these cached weights (fitness.py:160-169),
    # computed by the same solve kernel, so the in-place call need not re-solve the windows.  It lives in
    # the instance __dict__ and is lost by Equation.__deepcopy__ (main_structures.py:673 copies __slots__
    # only); _inplace_l2lr then recomputes it from _cached_sw_weights on the host.
    try:
        objective._b200_stability = (objective._cached_sw_weights, float(res.stability[v]))
    except AttributeError:      # a slots-only object: nothing to stash
        pass


def stability_from_cached(sw, ndim):
    """fitness.py:160-169 on an already cached (W, K) window-weight array: cv = var(ddof=1) / mean**2 per
    column, cv := 0 where mean == 0, sum over the feature columns (not the intercept) / ndim."""
    w = np.asarray(sw, dtype=np.float64)
    std = w.std(axis=0, ddof=1)
    mu = w.mean(axis=0)
    with np.errstate(divide='ignore', invalid='ignore'):
        cv = (std ** 2) / (mu ** 2)
        cv[mu == 0] = 0.0
    return float(sum(cv[:-1]) / ndim)


# ------------------------------------------------------------------------------------------
# operator base: EPDE's own CompoundOperator when importable, otherwise a standalone twin
# ------------------------------------------------------------------------------------------
class StandaloneCompoundOperator:
    """template.py:26-170 -- the contract the reference's directors / mappers rely on; used only where
    the ``epde`` package is absent (the mirrors then cannot be mixed with reference operators)."""
    key = None

    def __init__(self, param_keys: list = []):
        self.param_keys = param_keys
        self._params = {}
        self._suboperators = {}
        self.use_default_tags()

    @property
    def params(self):
        return self._params

    @params.setter
    def params(self, param_dict: dict):
        if set(self.param_keys) != set(param_dict.keys()):
            raise KeyError('Wrong keys of param dict')
        self._params = param_dict

    @property
    def suboperators(self):
        return self._suboperators

    def set_suboperators(self, operators: dict, probas: dict = {}):
        for k, v in operators.items():
            if not (isinstance(k, str) and (isinstance(v, (list, tuple, dict)) or hasattr(v, 'apply'))):
                raise TypeError('The suboperators of an evolutionary operator must be declared in format '
                                'key : value, where key is str and value - CompoundOperator, list, tuple or dict')
        self._suboperators = dict(operators)

    @property
    def arguments(self):
        return set()

    def get_suboperator_args(self, personal=False):
        args = set(self.arguments)
        if not personal:
            for op in self._suboperators.values():
                if hasattr(op, 'get_suboperator_args'):
                    args |= op.get_suboperator_args()
        return args - {'arguments', 'objective', 'obj', 'self'}

    def parse_suboperator_args(self, arguments: dict):
        arguments = arguments or {}
        sub = {k: {a: arguments[a] for a in op.get_suboperator_args()} if hasattr(op, 'get_suboperator_args') else {}
               for k, op in self._suboperators.items()}
        return {a: arguments[a] for a in self.get_suboperator_args(True)}, sub

    def apply(self, objective, arguments: dict):
        raise NotImplementedError('Trying to apply abstract superclass of the operator.')

    def use_default_tags(self):
        self._tags = set()

    @property
    def operator_tags(self):
        return self._tags


def reference_operator_base():
    """EPDE's CompoundOperator (template.py:26) if the package can be imported, else None."""
    try:
        from epde.operators.utils.template import CompoundOperator as ref_base
        return ref_base
    except Exception:       # noqa: BLE001  (ImportError, or a missing plotting dependency of epde)
        return None


def _check_penalty(params):
    if not (params['penalty_coeff'] > 0. and params['penalty_coeff'] < 1.):
        raise ValueError('Incorrect penalty coefficient set, value shall be in (0, 1).')


def _given_from_weights(objective, dense: bool):
    """(wterm row, intercept) for the in-place calls: the equation already carries weights.
    dense=True : L2LRFitness, features = ALL non-target terms and weights_final[:-1] must have one
                 entry per feature (fitness.py:127,138-140 -- numpy raises otherwise);
    dense=False: L2Fitness, features = nonzero-weight terms, weights = weights_final[:-1][wi != 0]
                 (fitness.py:70-74)."""
    T = len(objective.structure)
    tgt = int(objective.target_idx)
    feat_terms = [t for t in range(T) if t != tgt]
    wf = objective.weights_final
    row = np.zeros(_engine.MAX_TERMS)
    if dense:
        coefs = np.asarray(wf[:-1], dtype=np.float64)
        if coefs.shape[0] != len(feat_terms):
            raise ValueError(f'shapes ({len(feat_terms)},) and ({coefs.shape[0]},) not aligned: weights_final '
                             f'must hold one coefficient per non-target term')
        row[feat_terms] = coefs
    else:
        wi = np.asarray(objective.weights_internal)
        coefs = np.asarray(wf, dtype=np.float64)[:-1][wi != 0]
        row[np.asarray(feat_terms)[wi != 0]] = coefs
    return row, float(wf[-1])


def _inplace_l2lr(objectives):
    """[(fitness, stability)] of the in-place L2LRFitness call for several equations in ONE batch
    (fitness.py:127-174).  Window weights cached on the equation are reused (fitness.py:155-156): their
    stability comes from the stash the sparsity mirror left, or -- when the equation was deep-copied,
    unpickled or filled by someone else (the stash lives outside Equation.__slots__) -- is computed on the
    host from ``_cached_sw_weights`` exactly as the reference does (std**2 / mu**2).  Without cached
    weights the windows are solved on all features on the device (mode 1)."""
    ndim = len(binding().engine.inner_shape)
    modes, rows, bs, cached = [], [], [], []
    for o in objectives:
        row, b = _given_from_weights(o, dense=True)
        sw = getattr(o, '_cached_sw_weights', None)
        stash = getattr(o, '_b200_stability', None)
        if sw is None:
            cached.append(None)
        elif stash is not None and stash[0] is sw:
            cached.append(stash[1])
        else:
            cached.append(stability_from_cached(sw, ndim))
        modes.append(1 if sw is None else 2)
        rows.append(row); bs.append(b)
    res, _ = _evaluate(objectives, [[o.target_idx] for o in objectives],
                       given=(np.asarray(modes, np.int32), np.stack(rows), np.asarray(bs)))
    return [(float(res.fitness_l1[v]), float(cached[v] if cached[v] is not None else res.stability[v]))
            for v in range(len(objectives))]


def _write_inplace_fitness(objective, fitness, stability):
    """fitness.py:152-174."""
    objective.aic = None
    objective.aic_calculated = True
    objective.fitness_calculated = True
    objective.fitness_value = fitness
    objective.stability_calculated = True
    objective.coefficients_stability = stability


# ------------------------------------------------------------------------------------------
# operators (built against a base class by make_operator_classes)
# ------------------------------------------------------------------------------------------
def make_operator_classes(Base):
    """The mirror classes derived from ``Base`` (EPDE's CompoundOperator or the standalone twin)."""

    class LASSOSparsity(Base):
        """sparsity.py:153-215: fit the CV-thresholded LASSO of the equation's current target and store
        weights_internal / weights_final / _cached_sw_weights on it."""
        key = 'LASSOBasedSparsity'
        b200_mirror = True

        def apply(self, objective, arguments: dict):
            res, _ = _evaluate([objective], [[objective.target_idx]])
            _write_sparsity(objective, res, 0)
            objective._b200_last = (objective.target_idx, res)

        def use_default_tags(self):
            self._tags = {'sparsity', 'gene level', 'no suboperators', 'inplace'}

    class LinRegBasedCoeffsEquation(Base):
        """coeff_calculation.py:16-47: in this snapshot of the reference an assertion only (the sklearn
        refit is commented out, :48-94)."""
        key = 'LinRegCoeffCalc'
        b200_mirror = True

        def apply(self, objective, arguments: dict = None):
            assert objective.weights_internal_evald, \
                'Trying to calculate final weights before evaluating intermeidate ones (no sparsity).'

        def use_default_tags(self):
            self._tags = {'coefficient calculation', 'gene level', 'no suboperators', 'inplace'}

    class _FitnessMirror(Base):
        """shared plumbing of the two discrepancy fitness operators."""
        b200_mirror = True

        def batchable(self):
            """True when the out-of-place call can be scored from one engine batch: the sparsity and
            coefficient suboperators are the B200 mirrors (anything else is applied one by one)."""
            try:
                sp, cc = self.suboperators['sparsity'], self.suboperators['coeff_calc']
            except KeyError:
                return False
            return getattr(sp, 'b200_mirror', False) and getattr(cc, 'b200_mirror', False) and \
                sp.key == 'LASSOBasedSparsity' and cc.key == 'LinRegCoeffCalc'

        def _last_result(self, objective):
            last = getattr(objective, '_b200_last', None)
            if last is not None and last[0] == objective.target_idx:
                return last[1]
            return None

        def use_default_tags(self):
            self._tags = {'fitness evaluation', 'gene level', 'contains suboperators', 'inplace'}

    class L2Fitness(_FitnessMirror):
        """fitness.py:27-99: || (F w + b - t) * g ||_2, divided by penalty_coeff when sum(w) == 0."""
        key = 'DiscrepancyBasedFitness'

        def score(self, objective, res, v):
            """value of apply(force_out_of_place=True) from row ``v`` of a batch that already ran the LASSO
            of (objective, its current target); the equation carries that row's weights."""
            _check_penalty(self.params)
            value = float(res.fitness_l2[v])
            if np.sum(objective.weights_final) == 0:
                value /= self.params['penalty_coeff']
            return value

        def apply(self, objective, arguments: dict, force_out_of_place: bool = False):
            _, subop_args = self.parse_suboperator_args(arguments=arguments)
            if force_out_of_place:
                self.suboperators['sparsity'].apply(objective, subop_args.get('sparsity', {}))
            self.suboperators['coeff_calc'].apply(objective, subop_args.get('coeff_calc', {}))
            _check_penalty(self.params)
            last = self._last_result(objective) if force_out_of_place else None
            if last is not None:
                value = float(last.fitness_l2[0])
            else:
                row, b = _given_from_weights(objective, dense=False)
                res, _ = _evaluate([objective], [[objective.target_idx]],
                                   given=(np.array([2], np.int32), row[None, :], np.array([b])))
                value = float(res.fitness_l2[0])
            if np.sum(objective.weights_final) == 0:
                value /= self.params['penalty_coeff']
            if force_out_of_place:
                return value
            objective.fitness_calculated = True
            objective.fitness_value = value

    class L2LRFitness(_FitnessMirror):
        """fitness.py:102-183: L1 discrepancy ratio; in place also the coefficient-stability objective."""
        key = 'DiscrepancyBasedFitnessWithCV'

        def score(self, objective, res, v):
            if res.all_zero[v]:                      # fitness.py:122-123
                return None
            _check_penalty(self.params)
            return float(res.fitness_l1[v])

        def apply(self, objective, arguments: dict, force_out_of_place: bool = False):
            _, subop_args = self.parse_suboperator_args(arguments=arguments)
            if force_out_of_place:
                self.suboperators['sparsity'].apply(objective, subop_args.get('sparsity', {}))
                if all(objective.weights_internal == 0):
                    return None
            self.suboperators['coeff_calc'].apply(objective, subop_args.get('coeff_calc', {}))
            _check_penalty(self.params)
            if force_out_of_place:
                last = self._last_result(objective)
                if last is not None:
                    return float(last.fitness_l1[0])
                row, b = _given_from_weights(objective, dense=False)
                res, _ = _evaluate([objective], [[objective.target_idx]],
                                   given=(np.array([2], np.int32), row[None, :], np.array([b])))
                return float(res.fitness_l1[0])
            fitness, stability = _inplace_l2lr([objective])[0]
            _write_inplace_fitness(objective, fitness, stability)

        def apply_population(self, objectives, arguments: dict = None):
            """in-place call on many equations as ONE engine batch (what InitialParetoLevelSorting /
            OffspringUpdater trigger equation by equation, moeadd_specific.py:409-465,507-540)."""
            objectives = list(objectives)
            if not objectives:
                return
            for o in objectives:
                self.suboperators['coeff_calc'].apply(o, {})
            _check_penalty(self.params)
            for o, (fit, stab) in zip(objectives, _inplace_l2lr(objectives)):
                _write_inplace_fitness(o, fit, stab)

    class EqRightPartSelector(Base):
        """right_part_selection.py:19-152.  The numeric loop over candidate targets (:60-74) is one engine
        batch scored through the wired ``fitness_calculation`` suboperator (its ``score``), when that
        suboperator is a B200 fitness mirror over the B200 sparsity mirror; any other suboperator is
        applied target by target exactly as the reference does (:65).  The structural steps call the
        Equation's own methods where the reference does (reset_state, restore_property, randomize,
        remove_zero_terms)."""
        key = 'FitnessCheckingRightPartSelector'
        b200_mirror = True

        def apply(self, objective, arguments: dict):
            self.apply_population([objective], arguments)

        def _fitness_op(self):
            try:
                return self.suboperators['fitness_calculation']
            except KeyError:
                raise KeyError("EqRightPartSelector: suboperator 'fitness_calculation' is not set "
                               '(right_part_selection.py:65)') from None

        def apply_population(self, objectives, arguments: dict = None):
            """Run the selector on many equations, batching every round's candidate targets of all
            unfinished equations into one engine call."""
            arguments = arguments or {}
            _, subop_args = self.parse_suboperator_args(arguments=arguments)
            fit_args = subop_args.get('fitness_calculation', {})
            fit_op = self._fitness_op()
            batched = getattr(fit_op, 'b200_mirror', False) and hasattr(fit_op, 'score') and fit_op.batchable()
            for o in objectives:
                assert len(o.structure) == len(o.terms_labels)
            active = [o for o in objectives if not (o.simplified and o.is_correct_right_part)]
            done_directly = [o for o in objectives if (o.simplified and o.is_correct_right_part)]
            while active:
                cand = []
                for o in active:
                    o.reset_state(True)
                    var = o.main_var_to_explain
                    while not any(term.contains_deriv(var) for term in o.structure):
                        o.restore_property(mandatory_family=False, deriv=True)
                    cand.append([i for i, term in enumerate(o.structure) if term.contains_deriv(var)])
                if batched:
                    res, evals = _evaluate(active, cand)
                    rows = {}
                    for v, (e, t) in enumerate(evals):
                        rows.setdefault(e, []).append((v, t))
                nxt = []
                for e, o in enumerate(active):
                    best, best_fit = None, np.inf
                    if batched:
                        for v, t in rows.get(e, []):
                            o.target_idx = t
                            _write_sparsity(o, res, v)           # what the suboperator's sparsity leaves (:65)
                            fitness = fit_op.score(o, res, v)
                            if fitness is not None and fitness < best_fit:
                                best_fit = fitness
                                best = (t, o.weights_internal, list(o.weights_final), o._cached_sw_weights,
                                        getattr(o, '_b200_stability', None))
                            o.weights_internal_evald = False
                            o.weights_final_evald = False
                    else:
                        for t in cand[e]:
                            o.target_idx = t
                            fitness = fit_op.apply(o, arguments=fit_args, force_out_of_place=True)
                            if fitness is not None and fitness < best_fit:
                                best_fit = fitness
                                best = (t, o.weights_internal, [w for w in o.weights_final if w != 0],
                                        o._cached_sw_weights, getattr(o, '_b200_stability', None))
                            o.weights_internal_evald = False
                            o.weights_final_evald = False
                    if best is None:                              # every target gave None / inf (:76-78)
                        o.randomize()
                        nxt.append(o)
                        continue
                    o.target_idx = best[0]
                    o.weights_internal = best[1]
                    o.weights_final = best[2]
                    o._cached_sw_weights = best[3]
                    if best[4] is not None:
                        try:
                            o._b200_stability = best[4]
                        except AttributeError:
                            pass
                    o.weights_internal_evald = True
                    o.weights_final_evald = True
                    if not self.simplify_equation(o):
                        o.simplified = True
                    if o.structure[o.target_idx].contains_deriv(o.main_var_to_explain):
                        o.is_correct_right_part = True
                    if o.simplified and o.is_correct_right_part:
                        done_directly.append(o)
                    else:
                        nxt.append(o)
                active = nxt
            for o in done_directly:
                o.right_part_selected = True
                o.remove_zero_terms()

        @staticmethod
        def _factor_family_label(factor):
            """label used to intersect terms: token label, plus the dimension for families that carry
            one (right_part_selection.py:113-116)."""
            lab = factor.cache_label
            return lab[0] if len(factor.params) == 1 else (lab[0], lab[1][-1])

        def simplify_equation(self, objective) -> bool:
            """right_part_selection.py:96-148: when every surviving term (nonzero features + target)
            shares a factor, lower that factor's power in all of them by the smallest power present;
            factors reaching power 0 are dropped, emptied terms are re-randomised.  True = changed."""
            tgt = objective.target_idx
            feats = [t for i, t in enumerate(objective.structure) if i != tgt]
            kept = [t for t, w in zip(feats, objective.weights_internal) if w != 0]
            kept.append(objective.structure[tgt])
            if len(kept) < 2:
                return False
            common = list(frozenset.intersection(*[t.term_label_without_power for t in kept]))
            for shared in common:
                lowest, dims = np.inf, []
                for term in kept:
                    for f in term.structure:
                        if self._factor_family_label(f) == shared:
                            if len(f.params) > 1:
                                dims.append(f.params[-1])
                            lowest = min(lowest, f.cache_label[1][0])
                if len(set(dims)) >= 2:
                    continue
                for term in kept:
                    dropped = []
                    for f in term.structure:
                        if self._factor_family_label(f) != shared:
                            continue
                        for i in range(len(f.params_description)):
                            if f.params_description[i]['name'] == 'power':
                                f.params[i] -= lowest
                                if f.params[i] == 0:
                                    dropped.append(f)
                    term.structure = [f for f in term.structure if f not in dropped]
                    term.reset_saved_state()
                    while (len(term.structure) == 0 or not term.contains_meaningful()
                           or len(objective.terms_labels) != len(objective.structure)):
                        term.randomize()
                return True
            return False

        def use_default_tags(self):
            self._tags = {'equation right part selection', 'gene level', 'contains suboperators', 'inplace'}

    class RandomRHPSelector(Base):
        """right_part_selection.py:154-207: the single-objective selector (pairs with L2Fitness): a random
        admissible term becomes the target; when none holds a derivative of the main variable a random
        term is replaced by a freshly created derivative term (:188-197).  Purely structural -- no numeric
        work, so nothing runs on the device here; mirrored so that a director can take every operator of
        the path from this module."""
        key = 'RandomRightPartSelector'
        b200_mirror = True

        def apply(self, objective, arguments: dict):
            self.parse_suboperator_args(arguments=arguments or {})
            if objective.right_part_selected:
                return
            var = objective.main_var_to_explain
            admissible = [i for i, term in enumerate(objective.structure) if term.contains_deriv(variable=var)]
            if len(admissible) == 0:
                idx = np.random.choice([i for i, _ in enumerate(objective.structure)])
                prev = objective.structure[idx]
                while True:
                    candidate = type(prev)(pool=prev.pool, mandatory_family=var,
                                           max_factors_in_term=len(prev.structure), create_derivs=True)
                    if candidate.contains_deriv(variable=var):
                        break
                objective.structure[idx] = candidate
            else:
                idx = np.random.choice(admissible)
            objective.target_idx = idx
            objective.reset_explaining_term(idx)
            objective.right_part_selected = True

        def use_default_tags(self):
            self._tags = {'equation right part selection', 'gene level', 'contains suboperators', 'inplace'}

    return dict(LASSOSparsity=LASSOSparsity, LinRegBasedCoeffsEquation=LinRegBasedCoeffsEquation,
                L2Fitness=L2Fitness, L2LRFitness=L2LRFitness, EqRightPartSelector=EqRightPartSelector,
                RandomRHPSelector=RandomRHPSelector, CompoundOperator=Base)


def install_operator_classes(base=None):
    """(Re)build the mirror classes against ``base`` (default: EPDE's CompoundOperator when importable,
    else the standalone twin) and publish them as module attributes.  Returns the base used."""
    if base is None:
        base = reference_operator_base() or StandaloneCompoundOperator
    globals().update(make_operator_classes(base))
    globals()['OPERATOR_BASE'] = base
    return base


install_operator_classes()


def evaluate_population(objectives, bound: Binding | None = None):
    """In-place fitness + stability objectives of a whole population in one batch (what
    InitialParetoLevelSorting / OffspringUpdater trigger equation by equation through
    SoEq.obj_fun, moeadd_specific.py:409-465,507-540; eq_mo_objectives.py:19-41,106-115)."""
    objectives = list(objectives)
    out = _inplace_l2lr(objectives)
    for o, (fit, stab) in zip(objectives, out):
        _write_inplace_fitness(o, fit, stab)
    return out
