"""world_size-2 gloo tests (CPU) of the multi-GPU host logic in epde_b200/sharding.py.

The CUDA engine cannot run here, so the grid-sharded driver (`evaluate_grid_sharded`: gram ->
all-reduce -> solve -> residual -> all-reduce -> fetch) is exercised with a stand-in engine that
implements the same staged interface on host tensors with the ORACLE's slab-partial sums
(test infrastructure only).  What is checked: the partition arithmetic, the collective sequence and
the merge of results -- both ranks must reproduce the unsharded oracle evaluation."""
import ctypes
import os
import socket
import sys
from types import SimpleNamespace

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def _free_port():
    with socket.socket() as s:
        s.bind(('127.0.0.1', 0))
        return s.getsockname()[1]


def _view(ptr, n):
    return np.ctypeslib.as_array((ctypes.c_double * n).from_address(ptr))


def _decode(enc):
    eqs = []
    for e in range(enc.n_eq):
        terms = []
        for t in range(enc.term_off[e], enc.term_off[e + 1]):
            terms.append([(int(enc.fac_token[f]), float(enc.fac_power[f]), int(enc.fac_kind[f]), float(enc.fac_param[f]))
                          for f in range(enc.fac_off[t], enc.fac_off[t + 1])])
        eqs.append(terms)
    return eqs


class SlabEngine:
    """stand-in for epde_b200.Engine(shard=(lo, n0)): same staged interface, oracle arithmetic."""

    def __init__(self, local_tokens, global_shape, lo, hi):
        self.tokens, self.shape, self.lo, self.hi = local_tokens, tuple(global_shape), lo, hi

    def batch(self, enc):
        return SlabBatch(self, enc)


class SlabBatch:
    def __init__(self, eng, enc):
        from oracle import epde_oracle as orc
        self.orc, self.eng, self.enc = orc, eng, enc
        self.eqs = _decode(enc)
        self.plan = orc.WindowPlan(eng.shape)
        self.sizes = [sum(ax.nseg for ax in self.plan.axes) * (len(eq) + 1) ** 2 for eq in self.eqs]
        self.gram_size = int(sum(self.sizes))
        self.residual_size = 3 * enc.n_eval

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        pass

    def _cols(self, eq):
        return np.stack([self.orc.eval_term(self.eng.tokens, term) for term in eq], axis=1)

    def gram(self, ptr):
        out, o = _view(ptr, self.gram_size), 0
        shape, lo, hi = self.eng.shape, self.eng.lo, self.eng.hi
        for eq, size in zip(self.eqs, self.sizes):
            K = len(eq) + 1
            aug = np.concatenate([self._cols(eq), np.ones(((hi - lo) * int(np.prod(shape[1:])), 1))], axis=1)
            grid = aug.reshape((hi - lo,) + shape[1:] + (K,))
            parts = []
            for d, ax in enumerate(self.plan.axes):
                Xd = np.moveaxis(grid, d, 0).reshape(grid.shape[d], -1, K)
                slab = np.matmul(Xd.transpose(0, 2, 1), Xd)
                if d == 0:                       # rows of the GLOBAL axis: this rank fills [lo, hi)
                    full = np.zeros((shape[0], K, K))
                    full[lo:hi] = slab
                    slab = full
                parts.append(np.add.reduceat(slab, ax.bounds[:-1], axis=0).reshape(-1))
            out[o:o + size] = np.concatenate(parts)
            o += size

    def solve(self, ptr):
        seg_all, o = _view(ptr, self.gram_size), 0
        self.fit = []
        per_eq = []
        for eq, size in zip(self.eqs, self.sizes):
            K = len(eq) + 1
            flat, segs, q = seg_all[o:o + size], [], 0
            for ax in self.plan.axes:
                segs.append(flat[q:q + ax.nseg * K * K].reshape(ax.nseg, K, K).copy())
                q += ax.nseg * K * K
            per_eq.append((self.orc.window_grams(segs, self.plan), segs[0].sum(axis=0)))
            o += size
        for e, t in zip(self.enc.eval_eq, self.enc.eval_target):
            T = len(self.eqs[e])
            win, total = per_eq[e]
            feat = [i for i in range(T) if i != t]
            coef, b, cached = self.orc.lasso_fit_gram(win, total, feat, int(t))
            w = np.zeros(T)
            w[feat] = coef
            zero = bool(np.all(coef == 0))
            b1 = total[t, T] / total[T, T] if zero else float(b)
            self.fit.append((int(e), int(t), coef, float(b), w, b1, zero,
                             self.orc.stability_from_weights(cached, len(self.plan.shape))))

    def residual(self, ptr):
        out = _view(ptr, self.residual_size)
        local_plan = SimpleNamespace(shape=(self.eng.hi - self.eng.lo,) + self.eng.shape[1:])
        for v, (e, t, coef, b, w, b1, zero, stab) in enumerate(self.fit):
            out[3 * v:3 * v + 3] = self.orc.residual_sums_streamed(self.eng.tokens, self.eqs[e], local_plan,
                                                                   [(t, w, b1, b)])[0]

    def fetch(self, ptr, want_sw=True):
        s = _view(ptr, self.residual_size).reshape(-1, 3)
        return SimpleNamespace(coef=[f[2] for f in self.fit], intercept=np.array([f[3] for f in self.fit]),
                               fitness_l1=s[:, 0] / s[:, 1], fitness_l2=np.sqrt(s[:, 2]),
                               stability=np.array([f[7] for f in self.fit]),
                               all_zero=np.array([f[6] for f in self.fit]))


def _worker(rank, world, port, mode, q):
    import torch
    import torch.distributed as dist
    import epde_b200
    from epde_b200 import sharding, synthetic as syn
    os.environ['MASTER_ADDR'] = '127.0.0.1'
    os.environ['MASTER_PORT'] = str(port)
    dist.init_process_group('gloo', rank=rank, world_size=world)
    try:
        shape, _, kind = syn.CONFIGS['heat64']
        shape = (24, 12, 10)
        eqs = syn.make_population(len(syn.token_labels(kind)), 6, seed=11)
        evals = syn.make_evals(eqs)
        if mode == 'grid':
            (lo, hi), local_shape, shard = sharding.grid_slab(shape, rank, world)
            assert shard == (lo, shape[0]) and local_shape[0] == hi - lo
            toks = syn.make_tokens(shape, kind, seed=3, axis0_range=(lo, hi))
            eng = SlabEngine(toks, shape, lo, hi)
            enc = epde_b200.encode(eqs, evals)
            b = eng.batch(enc)
            seg = torch.zeros(b.gram_size, dtype=torch.float64)
            sums = torch.zeros(b.residual_size, dtype=torch.float64)
            res = sharding.evaluate_grid_sharded(eng, enc, seg, sums)
            q.put((rank, [np.asarray(c) for c in res.coef], res.intercept, res.fitness_l1, res.fitness_l2, res.stability))
        else:
            mine = sharding.population_slice(len(eqs), rank, world)
            loc, pos = sharding.local_evals(evals, mine)
            assert [evals[p] for p in pos] == [(mine[e], t) for e, t in loc]
            # a rank's "result": fields derived from the GLOBAL eval index so the merge can be checked
            n = len(pos)
            fake = SimpleNamespace(coef=np.tile(np.asarray(pos, float)[:, None], (1, epde_b200.MAX_TERMS)),
                                   intercept=np.asarray(pos, float) + 0.5, fitness_l1=np.asarray(pos, float) * 2,
                                   fitness_l2=np.asarray(pos, float) * 3, stability=np.asarray(pos, float) * 4,
                                   n_active=np.asarray(pos, np.int32), all_zero=np.zeros(n, np.int32),
                                   iters=np.full(n, rank, np.int32))
            merged = sharding.gather_population(fake, pos, len(evals))
            q.put((rank, merged, len(evals), [evals[p][0] % world for p in range(len(evals))]))
    finally:
        dist.destroy_process_group()


def _run(mode, world=2):
    import torch.multiprocessing as mp
    ctx = mp.get_context('spawn')
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, mode, q)) for r in range(world)]
    for p in procs:
        p.start()
    outs = [q.get(timeout=240) for _ in procs]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    return sorted(outs, key=lambda o: o[0])


def test_grid_sharded_driver_matches_unsharded_oracle():
    from oracle import epde_oracle as orc
    from epde_b200 import synthetic as syn
    outs = _run('grid')
    shape = (24, 12, 10)
    toks = syn.make_tokens(shape, 'heat', seed=3)
    eqs = syn.make_population(7, 6, seed=11)
    evals = syn.make_evals(eqs)
    ref = {}
    for e, eq in enumerate(eqs):
        tg = [t for ee, t in evals if ee == e]
        for t, r in zip(tg, orc.evaluate_equation(toks, shape, eq, targets=tg)):
            ref[(e, t)] = r
    for rank, coef, b, f1, f2, stab in outs:
        checked = 0
        for v, (e, t) in enumerate(evals):
            r = ref[(e, t)]
            if r.cond > 1e9:
                continue
            assert np.array_equal(coef[v] != 0, r.coef != 0), f'rank {rank} eval {v}: term set'
            np.testing.assert_allclose(coef[v], r.coef, rtol=1e-6, atol=1e-9 * max(1.0, np.abs(r.coef).max()))
            np.testing.assert_allclose(f1[v], r.fitness, rtol=1e-6)
            np.testing.assert_allclose(f2[v], r.fitness_l2, rtol=1e-6)
            checked += 1
        assert checked >= 3
    # deterministic: both ranks hold bit-identical results after the all-reduces
    for a, bb in zip(outs[0][1], outs[1][1]):
        assert np.array_equal(a, bb)
    assert np.array_equal(outs[0][3], outs[1][3])


def test_population_sharding_merge():
    outs = _run('population')
    for rank, merged, n, owner in outs:
        idx = np.arange(n, dtype=float)
        assert np.array_equal(merged['intercept'], idx + 0.5)
        assert np.array_equal(merged['fitness_l1'], idx * 2)
        assert np.array_equal(merged['coef'][:, 0], idx)
        assert np.array_equal(merged['iters'], np.asarray(owner, np.int32))   # who evaluated what


def test_partitions():
    from epde_b200 import sharding
    for n, world in [(128, 8), (7, 2), (3, 4)]:
        got = sorted(i for r in range(world) for i in sharding.population_slice(n, r, world))
        assert got == list(range(n))
    # cost-balanced partition: a permutation of the population, identical on every rank, better than round-robin
    rng = np.random.default_rng(4)
    T = rng.integers(2, 9, size=256).tolist()
    V = [int(rng.integers(1, t + 1)) for t in T]
    parts = sharding.population_partition(T, V, 8)
    assert sorted(i for part in parts for i in part) == list(range(256))
    assert parts == sharding.population_partition(T, V, 8) and all(part == sorted(part) for part in parts)
    cost = [sharding.equation_cost(t, v) for t, v in zip(T, V)]
    load = [sum(cost[i] for i in part) for part in parts]
    rr = [sum(cost[i] for i in sharding.population_slice(256, r, 8)) for r in range(8)]
    assert max(load) <= max(rr) and max(load) / (sum(load) / 8) < 1.01
    assert sharding.population_partition([3, 5], [1, 2], 4) == [[1], [0], [], []]
    with pytest.raises(ValueError):
        sharding.population_partition([3], [1, 2], 2)
    rows = []
    for r in range(8):
        (lo, hi), local, shard = sharding.grid_slab((512, 512, 512), r, 8)
        assert local == (64, 512, 512) and shard == (lo, 512)
        rows += list(range(lo, hi))
    assert rows == list(range(512))
    (lo, hi), local, _ = sharding.grid_slab((61, 61), 1, 2)
    assert (lo, hi) == (30, 61) and local == (31, 61)
    with pytest.raises(ValueError):
        sharding.grid_slab((3, 8), 0, 2)
    with pytest.raises(ValueError):
        sharding.population_slice(4, 2, 2)
