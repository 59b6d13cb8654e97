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
    Cache                       epde/cache/cache.py:141-524 (the subset the evaluator path touches)

The objects handed to ``apply`` are the reference's own ``Equation`` / ``Term`` / ``Factor`` instances
(duck-typed: only the attributes named in :func:`encode_equation` and in the operators are read), so
an EPDE director can construct these classes instead of its own (moeadd/strategy.py:46-63,
single_criterion/strategy.py:41-45) without touching any other EPDE file.  All arithmetic runs in
libepde_b200.so; this module packs encodings and writes results back.  No CPU fallback: without a
bound engine every ``apply`` raises.

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
    ``dim`` (uploaded by the caller; the reference's grid cache keys them '0', '1', ...,
    cache.py:66-84)."""

    def __init__(self, engine: _engine.Engine):
        self.engine = engine
        self.token_ids = {}
        self.families = {}        # factor label -> (kind, 'grid')  procedural families

    # -- tokens
    def token_id(self, label, create=False):
        if label not in self.token_ids:
            if not create:
                raise KeyError(f'token {label!r} is not in the B200 cache')
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


def bind(engine: _engine.Engine) -> Binding:
    """Install the engine the operators use (call where the reference calls globals.init_caches)."""
    global _binding
    _binding = Binding(engine)
    for lab, kind in _TRIG_KIND.items():
        _binding.register_family(lab, kind)
    return _binding


def binding() -> Binding:
    if _binding is None:
        raise RuntimeError('epde_b200.operators: no engine bound (call epde_b200.operators.bind(engine)); '
                           'this path has no CPU fallback')
    return _binding


def unbind():
    global _binding
    _binding = None


# ------------------------------------------------------------------------------------------
# Cache mirror (cache.py:141-524): same method names / return conventions for the calls the
# evaluator path makes; tensors live in HBM, nothing is kept on the host.
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
        """Cache.get (cache.py:397-452): host copy of the stored vector."""
        if torch_mode:
            raise NotImplementedError('torch_mode tensors are not part of the evaluator path')
        key = _base_label(label)
        if key not in self._b.token_ids:
            raise KeyError(f'Can not get tensor from cache with label {label}')
        return self._b.engine.download(self._b.token_ids[key])

    def __contains__(self, obj):                 # cache.py:454-470: (label, structural) or label
        label = obj[0] if (isinstance(obj, tuple) and len(obj) == 2 and isinstance(obj[1], bool)) else obj
        key = _base_label(label)
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


def _base_label(label):
    """reference cache keys are (label, (power, ...)) (supplementary.py:216-220); the HBM table holds
    the power-1 vectors and applies powers in-kernel, so key on the label part."""
    if isinstance(label, tuple) and len(label) == 2 and isinstance(label[1], tuple) and \
            not (label[0] == 'grid'):
        if len(label[1]) >= 1 and float(label[1][0]) != 1.0:
            raise KeyError(f'{label}: only power-1 tensors are stored; powers are applied in-kernel')
        return label[0]
    return label


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


def _evaluate(objectives, targets_per_eq, given=None, bound=None):
    """one engine batch over several equations; targets_per_eq[i] = list of target term indices."""
    bound = bound or binding()
    eqs = [encode_equation(o, bound) for o in objectives]
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
    # computed by the same solve kernel, so the in-place call need not re-solve the windows
    objective._b200_stability = (objective._cached_sw_weights, float(res.stability[v]))


# ------------------------------------------------------------------------------------------
# operators
# ------------------------------------------------------------------------------------------
class CompoundOperator:
    """template.py:26-170 -- the contract the reference's directors / mappers rely on."""
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


class LASSOSparsity(CompoundOperator):
    """sparsity.py:153-215: fit the CV-thresholded LASSO of the equation's current target and store
    weights_internal / weights_final / _cached_sw_weights on it."""
    key = 'LASSOBasedSparsity'

    def apply(self, objective, arguments: dict):
        res, _ = _evaluate([objective], [[objective.target_idx]])
        _write_sparsity(objective, res, 0)
        objective._b200_last = (objective.target_idx, res)

    def use_default_tags(self):
        self._tags = {'sparsity', 'gene level', 'no suboperators', 'inplace'}


class LinRegBasedCoeffsEquation(CompoundOperator):
    """coeff_calculation.py:16-47: in this snapshot of the reference an assertion only (the sklearn
    refit is commented out, :48-94)."""
    key = 'LinRegCoeffCalc'

    def apply(self, objective, arguments: dict = None):
        assert objective.weights_internal_evald, \
            'Trying to calculate final weights before evaluating intermeidate ones (no sparsity).'

    def use_default_tags(self):
        self._tags = {'coefficient calculation', 'gene level', 'no suboperators', 'inplace'}


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


class L2Fitness(CompoundOperator):
    """fitness.py:27-99: || (F w + b - t) * g ||_2, divided by penalty_coeff when sum(w) == 0."""
    key = 'DiscrepancyBasedFitness'

    def apply(self, objective, arguments: dict, force_out_of_place: bool = False):
        _, subop_args = self.parse_suboperator_args(arguments=arguments)
        if force_out_of_place:
            self.suboperators['sparsity'].apply(objective, subop_args.get('sparsity', {}))
        self.suboperators['coeff_calc'].apply(objective, subop_args.get('coeff_calc', {}))
        _check_penalty(self.params)
        last = getattr(objective, '_b200_last', None)
        if force_out_of_place and last is not None and last[0] == objective.target_idx:
            value = float(last[1].fitness_l2[0])
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

    def use_default_tags(self):
        self._tags = {'fitness evaluation', 'gene level', 'contains suboperators', 'inplace'}


class L2LRFitness(CompoundOperator):
    """fitness.py:102-183: L1 discrepancy ratio; in place also the coefficient-stability objective."""
    key = 'DiscrepancyBasedFitnessWithCV'

    def apply(self, objective, arguments: dict, force_out_of_place: bool = False):
        _, subop_args = self.parse_suboperator_args(arguments=arguments)
        if force_out_of_place:
            self.suboperators['sparsity'].apply(objective, subop_args.get('sparsity', {}))
            if all(objective.weights_internal == 0):
                return None
        self.suboperators['coeff_calc'].apply(objective, subop_args.get('coeff_calc', {}))
        _check_penalty(self.params)
        if force_out_of_place:
            last = getattr(objective, '_b200_last', None)
            if last is not None and last[0] == objective.target_idx:
                return float(last[1].fitness_l1[0])
            row, b = _given_from_weights(objective, dense=False)
            res, _ = _evaluate([objective], [[objective.target_idx]],
                               given=(np.array([2], np.int32), row[None, :], np.array([b])))
            return float(res.fitness_l1[0])
        fitness, stability = _inplace_l2lr([objective])[0]
        objective.aic = None
        objective.aic_calculated = True
        objective.fitness_calculated = True
        objective.fitness_value = fitness
        objective.stability_calculated = True
        objective.coefficients_stability = stability

    def use_default_tags(self):
        self._tags = {'fitness evaluation', 'gene level', 'contains suboperators', 'inplace'}


def _inplace_l2lr(objectives):
    """[(fitness, stability)] of the in-place L2LRFitness call for several equations in ONE batch
    (fitness.py:127-174).  Window weights cached on the equation are reused (fitness.py:155-156);
    otherwise the windows are solved on all features (mode 1)."""
    modes, rows, bs, cached = [], [], [], []
    for o in objectives:
        row, b = _given_from_weights(o, dense=True)
        sw = getattr(o, '_cached_sw_weights', None)
        stash = getattr(o, '_b200_stability', None)
        have = sw is not None and stash is not None and stash[0] is sw
        cached.append(stash[1] if have else None)
        if sw is not None and not have:
            raise RuntimeError('_cached_sw_weights was not produced by the B200 sparsity operator; clear it '
                               '(None) to have the windows re-solved on the device')
        modes.append(2 if have else 1)
        rows.append(row); bs.append(b)
    res, _ = _evaluate(objectives, [[o.target_idx] for o in objectives],
                       given=(np.asarray(modes, np.int32), np.stack(rows), np.asarray(bs)))
    return [(float(res.fitness_l1[v]), float(cached[v] if cached[v] is not None else res.stability[v]))
            for v in range(len(objectives))]


class EqRightPartSelector(CompoundOperator):
    """right_part_selection.py:19-152.  The numeric loop over candidate targets (:60-74) is one
    engine batch; the structural steps call the Equation's own methods exactly where the reference
    does (reset_state, restore_property, randomize, remove_zero_terms)."""
    key = 'FitnessCheckingRightPartSelector'

    def apply(self, objective, arguments: dict):
        self.apply_population([objective], arguments)

    def apply_population(self, objectives, arguments: dict = None):
        """Run the selector on many equations, batching every round's candidate targets of all
        unfinished equations into one engine call."""
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
            res, evals = _evaluate(active, cand)
            nxt = []
            for e, o in enumerate(active):
                best, best_fit = None, np.inf
                for v, (ee, t) in enumerate(evals):
                    if ee != e or res.all_zero[v]:
                        continue
                    if res.fitness_l1[v] < best_fit:
                        best, best_fit = v, float(res.fitness_l1[v])
                o.weights_internal_evald = False
                o.weights_final_evald = False
                if best is None:
                    o.randomize()
                    nxt.append(o)
                    continue
                o.target_idx = evals[best][1]
                _write_sparsity(o, res, best)
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


def evaluate_population(objectives, bound: Binding | None = None):
    """In-place fitness + stability objectives of a whole population in one batch (what
    InitialParetoLevelSorting / OffspringUpdater trigger equation by equation through
    SoEq.obj_fun, moeadd_specific.py:409-465,507-540; eq_mo_objectives.py:19-41,106-115)."""
    out = _inplace_l2lr(list(objectives))
    for o, (fit, stab) in zip(objectives, out):
        o.aic = None
        o.aic_calculated = True
        o.fitness_calculated = True
        o.fitness_value = fit
        o.stability_calculated = True
        o.coefficients_stability = stab
    return out
