"""EPDE-side wiring of the B200 evaluator: needs the ``epde`` package importable (it is NOT imported by
``import epde_b200``; import this module from the EPDE process).

What lives here (SURVEY.md section 8b 'injection point', section 8f item 1 'batched callers'):

  * :func:`bind_from_epde_globals` -- engine bound to EPDE's module-global caches (globals.py:26-42): the
    engine is created for ``global_var.grid_cache.inner_shape`` and every token an equation references is
    pulled from ``global_var.tensor_cache`` on first use and kept in HBM from then on.  Installed as the
    operators' auto-binder, so a director-built strategy needs no explicit ``bind`` call; a new
    ``EpdeSearch`` (which re-creates the caches) re-binds automatically.
  * :class:`B200MOEADDDirector` -- ``MOEADDDirector.use_baseline`` (optimizers/moeadd/strategy.py:31-96,
    ``use_solver=False`` branch) with the B200 operator mirrors in place of EPDE's numeric operators, and
    with batched versions of the two callers of the hot path:
      - :class:`B200InitialParetoLevelSorting` (moeadd_specific.py:489-540): right-part selection of all
        ``pop_size`` candidates as one engine batch per round, then all fitness evaluations as one batch;
      - :class:`B200OffspringUpdater` (moeadd_specific.py:406-465): the offspring of one call are mutated,
        right-part-selected and evaluated in rounds (one batch per round) instead of one by one.
    The structural / combinatorial operators (selection, variation, mutation, Pareto updates) are EPDE's
    own, untouched.  Use:  ``EpdeSearch(multiobjective_mode=True, use_default_strategy=False,
    director=make_moeadd_director())``  (interface/interface.py:312-313).

Batching changes the ORDER in which EPDE's global numpy RNG is consumed (mutations of all offspring happen
before the first evaluation), so a run is not draw-for-draw identical to an unbatched one; every evaluated
equation still gets exactly the values the unbatched operators give it (tests/test_reference_binding.py).
"""
from __future__ import annotations

from copy import deepcopy
from functools import partial

import numpy as np

import epde.globals as global_var
from epde.operators.utils.operator_mappers import map_operator_between_levels
from epde.operators.utils.template import CompoundOperator, add_base_param_to_operator
from epde.operators.multiobjective.selections import MOEADDSelection
from epde.operators.multiobjective.variation import get_basic_variation
from epde.operators.multiobjective.mutations import get_basic_mutation
from epde.operators.multiobjective.moeadd_specific import (InitialParetoLevelSorting, OffspringUpdater,
                                                           SimpleNeighborSelector, get_basic_populator_updater,
                                                           is_rps_in_other_equation)
from epde.optimizers.builder import add_sequential_operators, OptimizationPatternDirector, StrategyBuilder
from epde.optimizers.moeadd.strategy_elems import MOEADDSectorProcesser

from . import engine as _engine
from . import operators as ops

# the mirrors must derive from EPDE's CompoundOperator (template.py:72-78 checks issubclass)
if ops.OPERATOR_BASE is not CompoundOperator:
    ops.install_operator_classes(CompoundOperator)


# ------------------------------------------------------------------------------------------
# engine <-> EPDE module globals
# ------------------------------------------------------------------------------------------
_engine_factory = None        # callable(inner_shape) -> Engine-like; default: epde_b200.Engine on cuda:0
_bound_to = None              # the tensor_cache object the current binding was made for


def set_engine_factory(factory):
    """factory(inner_shape) -> engine.  Default creates ``epde_b200.Engine(inner_shape, device=0)``."""
    global _engine_factory
    _engine_factory = factory


def _token_source(label):
    """label -> (N,) float64 vector out of EPDE's caches, or None.  Plain tokens are the power-1 entries
    of the tensor cache (keys (label, (1.0,)), cache.py:37-63); ('grid', d) is coordinate axis d restricted
    to the interior like CustomEvaluator does (evaluators.py:87)."""
    tc, gc = global_var.tensor_cache, global_var.grid_cache
    if isinstance(label, tuple) and len(label) == 2 and label[0] == 'grid':
        grid = np.asarray(gc.get(str(int(label[1]))))
        return grid[gc.g_func != 0].reshape(-1)
    key = (label, (1.0,))
    if (key, False) in tc:
        return np.asarray(tc.get(key)).reshape(-1)
    return None


def bind_from_epde_globals(force=False):
    """(Re)bind the operators to EPDE's current module-global caches; no-op while they are unchanged."""
    global _bound_to
    tc = getattr(global_var, 'tensor_cache', None)
    gc = getattr(global_var, 'grid_cache', None)
    if tc is None or gc is None:
        return None
    if not force and ops._binding is not None and _bound_to is tc:
        return ops._binding
    inner = tuple(int(s) for s in gc.inner_shape)
    old = ops._binding
    factory = _engine_factory or (lambda shape: _engine.Engine(shape, device=0))
    eng = factory(inner)
    b = ops.bind(eng, _token_source)
    g = np.asarray(gc.g_func[gc.g_func_mask], dtype=np.float64).reshape(-1)     # sparsity.py:202
    eng.set_sample_weights(g)
    _bound_to = tc
    if old is not None and hasattr(old.engine, 'close'):
        old.engine.close()
    return b


ops.set_auto_binder(bind_from_epde_globals)


# ------------------------------------------------------------------------------------------
# batched callers
# ------------------------------------------------------------------------------------------
def _unwrap(operator):
    """the gene-level operator inside OperatorMapper / OperatorCondition wrappers (operator_mappers.py)."""
    seen = 0
    while seen < 8:
        seen += 1
        if hasattr(operator, '_conditioned_operator'):
            operator = operator._conditioned_operator
            continue
        try:
            inner = operator.suboperators['to_map']
        except (KeyError, TypeError, AttributeError):
            break
        operator = inner
    return operator


def _select_right_parts(sys_rps, candidates, arguments):
    """right-part selection of many systems.  Single-variable systems whose selector is the B200 mirror go
    through one batched call; everything else (multi-variable systems need the is_rps_in_other_equation
    loop, moeadd_specific.py:511-517) is handed to the mapped operator one by one."""
    gene_rps = _unwrap(sys_rps)
    batched, single = [], []
    for cand in candidates:
        if hasattr(gene_rps, 'apply_population') and len(cand.vars_to_describe) == 1:
            batched.append(cand)
        else:
            single.append(cand)
    if batched:
        eqs = [eq for cand in batched for eq in cand if not eq.right_part_selected]
        gene_rps.apply_population(eqs, arguments)
    for cand in single:
        sys_rps.apply(objective=cand, arguments=arguments)
        if len(cand.vars_to_describe) > 1:
            replaced = is_rps_in_other_equation(cand)
            while any(replaced):
                cand.reset_state(True)
                sys_rps.apply(objective=cand, arguments=arguments)
                replaced = is_rps_in_other_equation(cand)


def _evaluate_fitness(chromosome_fitness, candidates, arguments):
    """in-place fitness of many systems as one batch when the mapped operator is the L2LRFitness mirror."""
    gene_fit = _unwrap(chromosome_fitness)
    if hasattr(gene_fit, 'apply_population'):
        eqs = [eq for cand in candidates for eq in cand if not eq.fitness_calculated]
        gene_fit.apply_population(eqs, arguments)
    else:
        for cand in candidates:
            chromosome_fitness.apply(objective=cand, arguments=arguments)


class B200InitialParetoLevelSorting(InitialParetoLevelSorting):
    """moeadd_specific.py:489-540 with the candidates' right-part selection and fitness batched."""

    def apply(self, objective, arguments: dict):
        self_args, subop_args = self.parse_suboperator_args(arguments=arguments)
        if len(objective.population) == 0:
            sys_rps = self.suboperators['right_part_selector']
            fitness = self.suboperators['chromosome_fitness']
            candidates = list(objective.unplaced_candidates)
            for cand in candidates:
                cand.reset_state(True)
            _select_right_parts(sys_rps, candidates, subop_args['right_part_selector'])
            for cand in candidates:                      # duplicates are re-created one by one (:519-531)
                system = cand.terms_labels
                while system in objective.history:
                    cand.create()
                    cand.reset_state(True)
                    _select_right_parts(sys_rps, [cand], subop_args['right_part_selector'])
                    system = cand.terms_labels
                objective.history.add(system)
            _evaluate_fitness(fitness, candidates, subop_args['chromosome_fitness'])
            if global_var.verbose.candidate_objectives:
                for cand in candidates:
                    print(cand.obj_fun)
            objective.associate_weights()
            objective.initial_placing()
        return objective


class B200OffspringUpdater(OffspringUpdater):
    """moeadd_specific.py:406-465: every unplaced offspring is mutated, right-part-selected and -- when its
    system is new -- evaluated and offered to the Pareto-level updater.  Here the offspring of one call
    advance in rounds: mutate all pending ones, ONE batched right-part selection, then ONE batched fitness
    evaluation of the accepted ones; the Pareto updates run in the original order."""

    def apply(self, objective, arguments: dict):
        self_args, subop_args = self.parse_suboperator_args(arguments=arguments)
        mutation_attempt_limit = self.params['mutation_attempt_limit']
        offspring_attempt_limit = self.params['offspring_attempt_limit']
        mutation = self.suboperators['chromosome_mutation']
        sys_rps = self.suboperators['right_part_selector']
        fitness = self.suboperators['chromosome_fitness']
        pl_updater = self.suboperators['pareto_level_updater']

        pending = []
        while objective.unplaced_candidates:
            offspring = objective.unplaced_candidates.pop()
            offspring.reset_state(True)
            pending.append({'sys': deepcopy(offspring), 'attempt': 0, 'replaced': 0})
        while pending:
            for item in pending:
                item['sys'] = mutation.apply(objective=item['sys'], arguments=subop_args['chromosome_mutation'])
                item['sys'].reset_state(True)
            _select_right_parts(sys_rps, [item['sys'] for item in pending], subop_args['right_part_selector'])
            accepted, nxt = [], []
            for item in pending:
                system = item['sys'].terms_labels
                if system not in objective.history:
                    objective.history.add(system)
                    accepted.append(item['sys'])
                    continue
                if item['replaced'] == offspring_attempt_limit:
                    if global_var.verbose.candidate_objectives:
                        print('Could not generate unique offspring')
                    continue
                if item['attempt'] == mutation_attempt_limit:
                    item['sys'].create()
                    item['replaced'] += 1
                    item['attempt'] = 0
                item['attempt'] += 1
                nxt.append(item)
            _evaluate_fitness(fitness, accepted, subop_args['chromosome_fitness'])
            for cand in accepted:
                pl_updater.apply(objective=(cand, objective), arguments=subop_args['pareto_level_updater'])
                if global_var.verbose.candidate_objectives:
                    print(cand.obj_fun)
            pending = nxt
        return objective


# ------------------------------------------------------------------------------------------
# director
# ------------------------------------------------------------------------------------------
class B200MOEADDDirector(OptimizationPatternDirector):
    """MOEADDDirector (optimizers/moeadd/strategy.py:27-96) with the evaluator path on the B200 engine."""

    def use_baseline(self, use_solver: bool = False, use_pic: bool = True, variation_params: dict = {},
                     mutation_params: dict = {}, sorter_params: dict = {}, pareto_combiner_params: dict = {},
                     pareto_updater_params: dict = {}, batched: bool = True, **kwargs):
        if use_solver:
            raise NotImplementedError('SolverBasedFitness / PIC train a network per candidate: a different path, '
                                      'not served by the B200 evaluator (DESIGN.md, out of scope)')
        add_kwarg_to_operator = partial(add_base_param_to_operator, target_dict=kwargs)

        neighborhood_selector = SimpleNeighborSelector(['number_of_neighbors'])
        add_kwarg_to_operator(operator=neighborhood_selector)
        selection = MOEADDSelection(['delta', 'parents_fraction'])
        add_kwarg_to_operator(operator=selection)
        selection.set_suboperators({'neighborhood_selector': neighborhood_selector})
        variation = get_basic_variation(variation_params)

        right_part_selector = ops.EqRightPartSelector()
        sparsity = ops.LASSOSparsity()
        coeff_calc = ops.LinRegBasedCoeffsEquation()
        fitness = ops.L2LRFitness(['penalty_coeff'])
        add_kwarg_to_operator(operator=fitness)
        fitness.set_suboperators({'sparsity': sparsity, 'coeff_calc': coeff_calc})
        fitness_cond = lambda x: not getattr(x, 'fitness_calculated')          # noqa: E731
        right_part_selector.set_suboperators({'fitness_calculation': fitness})
        fitness_c = map_operator_between_levels(fitness, 'gene level', 'chromosome level',
                                                objective_condition=fitness_cond)
        sparsity_c = map_operator_between_levels(sparsity, 'gene level', 'chromosome level')
        rps_cond = lambda x: any([not elem_eq.right_part_selected for elem_eq in x.vals])   # noqa: E731
        sys_rps = map_operator_between_levels(right_part_selector, 'gene level', 'chromosome level',
                                              objective_condition=rps_cond)

        sorter = (B200InitialParetoLevelSorting if batched else InitialParetoLevelSorting)()
        add_base_param_to_operator(sorter, sorter_params)
        sorter.set_suboperators(operators={'right_part_selector': sys_rps, 'chromosome_fitness': fitness_c})

        updater = (B200OffspringUpdater if batched else OffspringUpdater)()
        add_base_param_to_operator(updater, pareto_combiner_params)
        updater.set_suboperators(operators={'chromosome_mutation': get_basic_mutation(mutation_params),
                                            'pareto_level_updater': get_basic_populator_updater(pareto_updater_params),
                                            'sparsity': sparsity_c,
                                            'right_part_selector': sys_rps,
                                            'chromosome_fitness': fitness_c})

        self.builder = add_sequential_operators(self.builder, [('initial_sorter', sorter),
                                                               ('selection', selection),
                                                               ('variation', variation),
                                                               ('pareto_updater_compl', updater)])
        self.operators = {'fitness': fitness, 'sparsity': sparsity, 'coeff_calc': coeff_calc,
                          'right_part_selector': right_part_selector, 'sys_rps': sys_rps, 'fitness_c': fitness_c}


def make_moeadd_director(batched=True, params_filename=None, **director_params):
    """A ready B200MOEADDDirector.  The operators read their defaults from the EvolutionaryParams singleton
    (default_parameter_loader.py:26), which EpdeSearch.__init__ normally creates AFTER it has been handed
    the director (interface.py:302-304) -- so it is created here first, in multi-objective mode."""
    from epde.operators.utils.default_parameter_loader import EvolutionaryParams
    EvolutionaryParams.reset()
    EvolutionaryParams(parameter_file=params_filename, mode='multi objective')
    director = B200MOEADDDirector()
    director.builder = StrategyBuilder(MOEADDSectorProcesser)
    director.use_baseline(use_solver=False, batched=batched, **director_params)
    return director


# ------------------------------------------------------------------------------------------
# the process-global tensor cache on the device (epde/cache/cache.py:141-524,
# globals.py:26-42) and its persistence (section 8f item 4: loader.py:242-283)
# ------------------------------------------------------------------------------------------
from collections.abc import MutableMapping      # noqa: E402

from epde.cache.cache import Cache as _RefCache  # noqa: E402


class _HbmDict(MutableMapping):
    """The ``memory_default['numpy']`` dictionary of a Cache with its values in HBM: storing an (N,) float64 vector
    uploads it as an engine token and keeps NO host copy; reading downloads a fresh copy (callers get values, never
    references -- the reference's contract is 'do not mutate', cache.py:397-427).  Anything that is not a vector of N
    points (N = points of the inner grid) stays on the host."""

    def __init__(self, owner):
        self._owner = owner
        self._ids = {}          # key -> engine token id
        self._host = {}         # key -> object kept on the host (not an (N,) vector)

    def _bound(self):
        return self._owner.device_binding()

    def __setitem__(self, key, value):
        arr = np.asarray(value)
        b = self._bound()
        if arr.dtype != object and arr.ndim >= 1 and arr.size == b.engine.n_points:
            tid = b.token_id(('cache', key), create=True)
            b.engine.upload(tid, np.ascontiguousarray(arr, dtype=np.float64).reshape(-1))
            self._ids[key] = (tid, arr.shape)
            self._host.pop(key, None)
            # power-1 entries of plain tokens are what equations reference by factor label (supplementary.py:216-220)
            if isinstance(key, tuple) and len(key) == 2 and isinstance(key[0], str) and key[1] == (1.0,):
                b.token_ids[key[0]] = tid
        else:
            self._host[key] = value
            self._ids.pop(key, None)

    def __getitem__(self, key):
        if key in self._ids:
            tid, shape = self._ids[key]
            return self._bound().engine.download(tid).reshape(shape)
        return self._host[key]

    def __delitem__(self, key):
        if key in self._ids:
            tid, _ = self._ids.pop(key)
            b = self._bound()
            b.engine.delete(tid)
            b.token_ids.pop(('cache', key), None)
            if isinstance(key, tuple) and len(key) == 2 and b.token_ids.get(key[0]) == tid:
                b.token_ids.pop(key[0], None)
        else:
            del self._host[key]

    def __iter__(self):
        yield from self._ids
        yield from self._host

    def __len__(self):
        return len(self._ids) + len(self._host)

    def __contains__(self, key):
        try:
            return key in self._ids or key in self._host
        except TypeError:
            return False

    def nbytes(self):
        return 8 * self._bound().engine.n_points * len(self._ids)


class DeviceTensorCache(_RefCache):
    """``epde.cache.cache.Cache`` whose base tensors live in HBM only: a drop-in for ``global_var.tensor_cache``.
    Every inherited method keeps its behaviour (add / get / __contains__ / delete_entry / get_all / prune_tensors /
    the count-based budget of memory_usage_properties); what changes is where ``memory_default['numpy']`` keeps its
    values.  ``use_structural`` with base data marks entries as merged instead of copying the dictionary (the copy
    would bring every tensor back to the host; ``get(structural=True)`` already redirects merged labels to the base
    data, cache.py:412-414).  The engine is bound lazily from EPDE's globals at the first stored tensor (the inner
    shape is known once ``grid_cache.set_boundaries`` has run, cache.py:250-270)."""

    def __init__(self, device='cpu'):
        super().__init__(device=device)
        self.memory_default['numpy'] = _HbmDict(self)
        self._b200_binding = None

    def device_binding(self):
        if self._b200_binding is None:
            global _bound_to
            self._b200_binding = bind_from_epde_globals(force=True)
            _bound_to = self
        return self._b200_binding

    def use_structural(self, use_base_data=True, label=None, replacing_data=None):
        if use_base_data and label is None:
            for key in self.memory_default['numpy'].keys():
                self.structural_and_base_merged[key] = True
            return
        if use_base_data and isinstance(label, tuple):
            self.structural_and_base_merged[label] = True
            return
        super().use_structural(use_base_data=use_base_data, label=label, replacing_data=replacing_data)

    @property
    def consumed_memory(self):
        mem = self.memory_default['numpy'].nbytes()
        mem += int(np.sum([np.asarray(v).nbytes for v in self.memory_normalized['numpy'].values()]))
        for lab, merged in self.structural_and_base_merged.items():
            if not merged:
                mem += self.memory_structural['numpy'][lab].nbytes
        return mem

    # ---- persistence (loader.py:242-283 pickles a Cache through its __dict__): host form and back
    def to_host_cache(self):
        """a plain reference ``Cache`` holding numpy copies of everything: what EPDELoader / pickle can store."""
        host = _RefCache(device=self._device)
        for k, v in self.__dict__.items():
            if k in ('memory_default', '_b200_binding'):
                continue
            setattr(host, k, v)
        host.memory_default = {'torch': dict(self.memory_default['torch']),
                               'numpy': {k: np.array(v) for k, v in self.memory_default['numpy'].items()}}
        return host

    @classmethod
    def from_host_cache(cls, host):
        """re-upload a restored reference ``Cache`` (SURVEY.md section 5 'checkpoint / resume')."""
        dev = cls(device=getattr(host, '_device', 'cpu'))
        for k, v in host.__dict__.items():
            if k != 'memory_default':
                setattr(dev, k, v)
        dev._b200_binding = None
        for k, v in host.memory_default['numpy'].items():
            dev.memory_default['numpy'][k] = v
        dev.memory_default['torch'] = dict(host.memory_default.get('torch', {}))
        return dev


def install_device_tensor_cache():
    """Make every ``EpdeSearch`` created from now on keep its tensor cache in HBM: wraps ``globals.init_caches``
    (globals.py:26-42), which EPDE calls while setting the domain up (interface.py:486-509), so that it installs a
    :class:`DeviceTensorCache` as ``global_var.tensor_cache``.  Returns a function that undoes the wrapping."""
    original = global_var.init_caches
    if getattr(original, '_b200_wrapped', False):
        return lambda: None

    def init_caches(set_grids=False, device='cpu'):
        original(set_grids=set_grids, device=device)
        global_var.tensor_cache = DeviceTensorCache(device=device)
        ops.unbind()

    init_caches._b200_wrapped = True
    global_var.init_caches = init_caches

    def undo():
        global_var.init_caches = original
    return undo
