"""Stand-ins for the reference's Factor / Term / Equation with the SAME attribute names the operator
mirrors read (the reference tree does not exist on the GPU box, so the -m gpu tests drive
epde_b200.operators through these; tests/test_reference_binding.py drives the same code with the
reference's real objects in the build container).

Semantics follow structure/factor.py:150-275 (params / params_description / cache_label),
structure/main_structures.py:308-367 (Term.contains_deriv, term_label_without_power),
:501-513 (Equation.remove_zero_terms), :643-659 (reset_state), :861-874 (terms_labels)."""
import numpy as np


class Factor:
    def __init__(self, label, power=1, is_deriv=False, variable='u', extra=None):
        """extra: ordered dict of further parameters (e.g. {'freq': 2.0, 'dim': 1} for trig tokens)."""
        names = ['power'] + list((extra or {}).keys())       # trig tokens: (power, freq, dim)
        vals = [float(power)] + [float(v) for v in (extra or {}).values()]
        self.label = label
        self.params = np.array(vals)
        self.params_description = {i: {'name': n, 'bounds': (0, 0)} for i, n in enumerate(names)}
        self.is_deriv = is_deriv
        self.variable = variable
        self.status = {'meaningful': True}
        self.ftype = 'trigonometric' if label in ('sin', 'cos') else 'plain'

    @property
    def cache_label(self):
        return (self.label, tuple(self.params))


class Term:
    def __init__(self, factors):
        self.structure = list(factors)

    def contains_deriv(self, variable=None):
        return sum(f.is_deriv and (variable is None or f.variable == variable) for f in self.structure) == 1

    def contains_meaningful(self):
        return any(f.status['meaningful'] for f in self.structure)

    @property
    def term_label_without_power(self):
        return frozenset(f.cache_label[0] if len(f.params) == 1 else (f.cache_label[0], f.cache_label[1][-1])
                         for f in self.structure)

    def reset_saved_state(self):
        pass

    def randomize(self):
        raise RuntimeError('stand-in terms cannot be re-randomised (the test equation degenerated)')


class Equation:
    def __init__(self, terms, main_var='u'):
        self.structure = list(terms)
        self.main_var_to_explain = main_var
        self.target_idx = len(self.structure) - 1
        self.simplified = False
        self.is_correct_right_part = False
        self.right_part_selected = False
        self.weights_internal = None
        self.weights_internal_evald = False
        self.weights_final = None
        self.weights_final_evald = False
        self._cached_sw_weights = None
        self._eval_cache = {}
        self.fitness_calculated = False
        self.fitness_value = None
        self.stability_calculated = False
        self.coefficients_stability = None

    @property
    def terms_labels(self):
        return frozenset(frozenset(f.cache_label for f in t.structure) for t in self.structure)

    def reset_state(self, reset_right_part=True):
        if reset_right_part:
            self.right_part_selected = False
            self.is_correct_right_part = False
            self.simplified = False
            self.weights_internal_evald = False
            self.weights_internal = None
            self.weights_final = None
        self.weights_final_evald = False
        self.fitness_calculated = False
        self.fitness_value = None
        self.stability_calculated = False
        self.coefficients_stability = None

    def restore_property(self, **kw):
        raise RuntimeError('stand-in equations always hold a derivative term')

    def randomize(self):
        raise RuntimeError('stand-in equations cannot be re-randomised (every target gave all-zero weights)')

    def remove_zero_terms(self):
        if self.weights_internal_evald:
            zero, bias = [], 0
            for i in range(len(self.structure)):
                if i == self.target_idx:
                    continue
                idx = i if i < self.target_idx else i - 1
                if self.weights_internal[idx] == 0:
                    bias += 1 if i < self.target_idx else 0
                    zero.append(i)
            self.structure = [t for i, t in enumerate(self.structure) if i not in zero]
            self.target_idx -= bias


def equation_from_encoding(eq, labels, deriv_labels):
    """nested encoding [[(token, power, kind, param)]] -> stand-in Equation (kind-0 factors)."""
    return Equation([Term([Factor(labels[f[0]], f[1], is_deriv=labels[f[0]] in deriv_labels) for f in term])
                     for term in eq])
