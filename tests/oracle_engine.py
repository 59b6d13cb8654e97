"""TEST INFRASTRUCTURE ONLY: a stand-in for ``epde_b200.Engine`` with the same host interface whose numbers
come from the ORACLE (oracle/epde_oracle.py).  Lets the CPU suite drive the operator mirrors, the director and
``EpdeSearch.fit`` in the build container, where no GPU exists.  Never imported by product code."""
import numpy as np

from epde_b200 import engine as E
from oracle import epde_oracle as orc


def _decode(enc):
    eqs = []
    for e in range(enc.n_eq):
        terms = []
        for t in range(enc.term_off[e], enc.term_off[e + 1]):
            terms.append([(int(enc.fac_token[f]), float(enc.fac_power[f]), int(enc.fac_kind[f]),
                           float(enc.fac_param[f])) for f in range(enc.fac_off[t], enc.fac_off[t + 1])])
        eqs.append(terms)
    return eqs


class OracleEngine:
    def __init__(self, inner_shape, device=0, shard=None, stream=None):
        self.inner_shape = tuple(int(s) for s in inner_shape)
        self.n_points = int(np.prod(self.inner_shape))
        self.tokens = {}
        self.sample_weights = None
        self.n_batches = 0
        self.n_evals = 0
        self.plan = orc.WindowPlan(self.inner_shape)

    # token cache
    def upload(self, token_id, values):
        a = np.ascontiguousarray(values, dtype=np.float64).reshape(-1)
        assert a.size == self.n_points
        self.tokens[int(token_id)] = a

    def download(self, token_id):
        return self.tokens[int(token_id)].copy()

    def __contains__(self, token_id):
        return int(token_id) in self.tokens

    def delete(self, token_id):
        del self.tokens[int(token_id)]

    @property
    def consumed_memory(self):
        return 8 * self.n_points * len(self.tokens)

    def set_sample_weights(self, g):
        g = np.asarray(g, dtype=np.float64).reshape(-1)
        assert g.size == self.n_points
        self.sample_weights = None if np.all(g == 1.0) else g

    @property
    def n_windows(self):
        return int(sum(len(ax.starts) for ax in self.plan.axes))

    def close(self):
        pass

    # evaluation
    def _token_list(self):
        n = max(self.tokens) + 1 if self.tokens else 0
        return [self.tokens.get(i) for i in range(n)]

    def evaluate(self, enc, want_sw=True, want_l2=True, given=None):
        self.n_batches += 1
        self.n_evals += enc.n_eval
        eqs = _decode(enc)
        W = self.n_windows
        o = E._alloc_out(enc.n_eval, W, True)
        toks = self._token_list()
        by_eq = {}
        for v in range(enc.n_eval):
            by_eq.setdefault(int(enc.eval_eq[v]), []).append(v)
        ndim = len(self.inner_shape)
        for e, vs in by_eq.items():
            eq = eqs[e]
            cols = orc.term_columns(toks, eq)
            lasso_vs = [v for v in vs if given is None or given[0][v] == 0]
            if lasso_vs:
                rs = orc.evaluate_equation(toks, self.inner_shape, eq, targets=[int(enc.eval_target[v]) for v in lasso_vs],
                                           sample_weights=self.sample_weights, columns=cols)
                for v, r in zip(lasso_vs, rs):
                    F = len(eq) - 1
                    o['coef'][v, :F] = r.coef
                    o['intercept'][v] = r.intercept
                    o['fitness_l1'][v] = r.fitness
                    o['fitness_l2'][v] = r.fitness_l2
                    o['stability'][v] = r.stability
                    na = int(np.sum(r.coef != 0))
                    o['n_active'][v] = na
                    o['all_zero'][v] = int(na == 0)
                    o['sw_weights'][v, :, :r.sw_weights.shape[1]] = r.sw_weights
            for v in vs:
                if given is None or given[0][v] == 0:
                    continue
                mode, wterm, b = int(given[0][v]), given[1][v], float(given[2][v])
                tgt = int(enc.eval_target[v])
                feat = [i for i in range(len(eq)) if i != tgt]
                w = np.asarray([wterm[i] for i in feat])
                nz = w != 0
                fsel = cols[:, np.asarray(feat, dtype=np.int64)[nz]] if np.any(nz) else None
                o['fitness_l1'][v] = orc.l1_fitness(cols[:, tgt], fsel, w[nz], b)
                o['fitness_l2'][v] = orc.l2_fitness(cols[:, tgt], fsel, w[nz], b, None)
                o['coef'][v, :len(feat)] = w
                o['intercept'][v] = b
                o['n_active'][v] = int(np.sum(nz))
                o['all_zero'][v] = int(not np.any(nz))
                if mode == 1:
                    g = np.ones(self.n_points) if self.sample_weights is None else self.sample_weights
                    sw = orc.calculate_weights_direct(cols[:, feat], cols[:, tgt], g, self.inner_shape)
                    o['stability'][v] = orc.stability_from_weights(sw, ndim)
                    o['sw_weights'][v, :, :sw.shape[1]] = sw
                    o['n_active'][v] = len(feat)
                else:
                    o['stability'][v] = np.nan
        if not want_sw:
            o['sw_weights'] = None
        return E._trim(o, enc.n_eval)

    def batch(self, enc):
        return _OracleBatch(self, enc)


class _OracleBatch:
    def __init__(self, eng, enc):
        self.eng, self.enc, self.given = eng, enc, None

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        pass

    def set_given(self, mode, wterm, intercept):
        self.given = (np.asarray(mode), np.asarray(wterm), np.asarray(intercept))

    def set_norms(self, want_l2=True):
        pass

    def gram(self, dev_seg=None):
        pass

    def solve(self, dev_seg=None):
        pass

    def residual(self, dev_sums=None):
        pass

    def fetch(self, dev_sums=None, want_sw=True):
        return self.eng.evaluate(self.enc, want_sw=want_sw, given=self.given)
