"""Seeded synthetic workloads of BASELINE.json's configs.

Fields are analytic solutions on uniform grids (boundary 0, so inner shape == grid shape); the
tokens are u and its analytic derivatives up to order 2 per axis, as the reference would hold them
in its tensor cache after ``create_pool(..., derivs=...)`` (interface/interface.py:133-155), plus
Gaussian noise.  Populations follow the reference scripts: 1-2 factors per term with probabilities
(0.8, 0.2) (examples/wave_equation_multiobjective.py:114), powers <= 3, every equation holding at
least one derivative term, a mix of fresh (T = 5..8) and pruned (T = 2..4) equations as observed in
a MOEADD run (SURVEY.md section 8d 'observed workload shape').

Works on numpy arrays (CPU, tests / CPU baseline) or torch tensors (GPU, bench) through the tiny
``xp`` shim; nothing here is on the evaluator's hot path.
"""
from __future__ import annotations

import numpy as np

CONFIGS = {
    # name: (shape, population, kind)
    'wave61': ((61, 61), 8, 'wave'),            # stand-in of config 1's size (real data lives in tests/golden)
    'burgers156': ((156, 156), 16, 'wave'),     # config 2's inner size
    'wave1024': ((1024, 1024), 64, 'wave'),     # config 3
    'heat256': ((256, 256, 256), 256, 'heat'),  # config 4
    'heat512': ((512, 512, 512), 128, 'heat'),  # config 5
    'heat64': ((64, 64, 64), 32, 'heat'),       # small stand-ins for tests / CPU baselines
    'heat128': ((128, 128, 128), 32, 'heat'),
}


def token_labels(kind):
    if kind == 'wave':
        return ['u', 'du/dx0', 'd^2u/dx0^2', 'du/dx1', 'd^2u/dx1^2']
    return ['u', 'du/dx0', 'd^2u/dx0^2', 'du/dx1', 'd^2u/dx1^2', 'du/dx2', 'd^2u/dx2^2']


def _axes(xp, shape, kind, lo_hi, dev):
    axes = []
    for n, (lo, hi) in zip(shape, lo_hi):
        if xp is np:
            axes.append(np.linspace(lo, hi, n))
        else:
            axes.append(xp.linspace(lo, hi, n, dtype=xp.float64, device=dev))
    return axes


def make_tokens(shape, kind, xp=np, device=None, noise=1e-3, seed=0, axis0_range=None, only=None):
    """list of flat float64 token vectors (C order) in token_labels(kind) order.

    axis0_range=(lo, hi): only rows [lo, hi) of the time axis (grid-point sharding); the noise
    stream is keyed by (seed, token, global row) so shards reproduce the unsharded field.
    only: iterable of token indices to build (others are returned as None)."""
    shape = tuple(shape)
    lo0, hi0 = (0, shape[0]) if axis0_range is None else axis0_range
    pi = float(np.pi)
    if kind == 'wave':
        c = 0.2
        t, x = _axes(xp, shape, kind, [(0.0, 1.0), (0.0, 1.0)], device)
        t = t[lo0:hi0]
        T, X = t[:, None], x[None, :]
        s1, c1 = xp.sin(pi * X), xp.cos(pi * X)
        s2, c2 = xp.sin(2 * pi * X), xp.cos(2 * pi * X)
        a1, b1 = xp.cos(c * pi * T), xp.sin(c * pi * T)
        a2, b2 = xp.cos(2 * c * pi * T), xp.sin(2 * c * pi * T)
        toks = [lambda: s1 * a1 + 0.5 * s2 * a2,
                lambda: -c * pi * s1 * b1 - c * pi * s2 * b2,
                lambda: -(c * pi) ** 2 * s1 * a1 - 2 * (c * pi) ** 2 * s2 * a2,
                lambda: pi * c1 * a1 + pi * c2 * a2,
                lambda: -pi ** 2 * s1 * a1 - 2 * pi ** 2 * s2 * a2]
    elif kind == 'heat':
        k = 0.5
        t, x, y = _axes(xp, shape, kind, [(0.0, 1.0), (0.0, pi), (0.0, pi)], device)
        t = t[lo0:hi0]
        T, X, Y = t[:, None, None], x[None, :, None], y[None, None, :]
        e1, e2 = xp.exp(-2 * k * T), 0.3 * xp.exp(-5 * k * T)
        sx, cx, s2x, c2x = xp.sin(X), xp.cos(X), xp.sin(2 * X), xp.cos(2 * X)
        sy, cy = xp.sin(Y), xp.cos(Y)
        toks = [lambda: e1 * sx * sy + e2 * s2x * sy,
                lambda: -2 * k * e1 * sx * sy - 5 * k * e2 * s2x * sy,
                lambda: 4 * k * k * e1 * sx * sy + 25 * k * k * e2 * s2x * sy,
                lambda: e1 * cx * sy + 2 * e2 * c2x * sy,
                lambda: -e1 * sx * sy - 4 * e2 * s2x * sy,
                lambda: e1 * sx * cy + e2 * s2x * cy,
                lambda: -e1 * sx * sy - e2 * s2x * sy]
    else:
        raise ValueError(kind)
    out = []
    row = int(np.prod(shape[1:]))
    if noise:
        # noise amplitude from a coarse copy of the GLOBAL field, so every shard uses the same one
        coarse = make_tokens(tuple(min(n, 33) for n in shape), kind, noise=0.0)
    only = None if only is None else set(int(i) for i in only)
    for i, make in enumerate(toks):
        if only is not None and i not in only:
            out.append(None)
            continue
        tk = make().reshape(-1)
        if noise:
            scale = noise * float(coarse[i].std())
            tk = tk + scale * _counter_normal(xp, lo0 * row, tk.shape[0], seed, i, device)
        out.append(tk.contiguous() if xp is not np else np.ascontiguousarray(tk))
    return out


def _s64(c):
    """a 64-bit constant as the signed value torch's int64 arithmetic wraps to"""
    c &= (1 << 64) - 1
    return c - (1 << 64) if c >= (1 << 63) else c


def _counter_normal(xp, start, n, seed, token, device):
    """standard normal noise as a pure function of (seed, token, GLOBAL point index): splitmix64 finaliser on the
    index, two 53-bit uniforms, Box-Muller.  Unlike a device generator the values do not depend on the launch
    geometry torch picks on a given device, a shard of the grid sees exactly the values of the whole grid, and the
    numpy path (CPU baseline / reference arm) draws the SAME stream as the torch path (GPU arm): both wrap 64-bit
    integer arithmetic the same way."""
    key = _s64(seed * 0x632BE59BD9B4E019 + token * 0xD1342543DE82EF95 + 0x2545F4914F6CDD1D)
    if xp is np:
        out = np.empty(n)
        step = 1 << 22                       # bounded temporaries on the host
        with np.errstate(over='ignore'):
            for lo in range(0, n, step):
                m = min(step, n - lo)
                idx = np.arange(start + lo, start + lo + m, dtype=np.int64)
                out[lo:lo + m] = _box_muller(np, idx, np.int64(key), lambda a: a.astype(np.float64))
        return out
    idx = xp.arange(start, start + n, dtype=xp.int64, device=device)
    return _box_muller(xp, idx, key, lambda a: a.to(xp.float64))


def _box_muller(xp, idx, key, to_f64):
    c1, c2, c3 = (np.int64(_s64(c)) if xp is np else _s64(c)
                  for c in (0x9E3779B97F4A7C15, 0xBF58476D1CE4E5B9, 0x94D049BB133111EB))

    def mix(z):
        z = z * c1 + key
        z = (z ^ ((z >> 30) & ((1 << 34) - 1))) * c2
        z = (z ^ ((z >> 27) & ((1 << 37) - 1))) * c3
        return z ^ ((z >> 31) & ((1 << 33) - 1))

    m53 = (1 << 53) - 1
    u1 = (to_f64((mix(idx * 2) >> 11) & m53) + 0.5) * 2.0 ** -53
    u2 = to_f64((mix(idx * 2 + 1) >> 11) & m53) * 2.0 ** -53
    return xp.sqrt(-2.0 * xp.log(u1)) * xp.cos(2.0 * np.pi * u2)


def data_checksum(token, xp=np):
    """sampled sum of a token vector: the same number on every box / arm <=> the same data."""
    return float(token[::4099].sum())


def _term_pools(n_tokens, max_power):
    """every product of 1-3 distinct tokens with powers 1..max_power, split into (admissible targets, others)."""
    from itertools import combinations, product
    adm, other = [], []
    for nf in (1, 2, 3):
        for toks in combinations(range(n_tokens), nf):
            for pw in product(range(1, max_power + 1), repeat=nf):
                term = [(int(t), float(q), 0, 0.0) for t, q in zip(toks, pw)]
                (adm if sum(1 for f in term if f[0] != 0) == 1 else other).append(term)
    return adm, other


def make_population_distinct(n_tokens, pop, seed=0, t_range=(3, 7), max_power=3):
    """stress population with NO sharing: every term of every equation is distinct (products of 1-3 factors with
    powers <= max_power), so the column dictionary de-duplicates nothing (report it beside the default).

    Terms are dealt without replacement from the enumerated pools (bounded work: no rejection loop).  Every equation
    gets at least one admissible target; there are only 6 * max_power * (1 + max_power) of those for 7 tokens, so
    max_power is raised until the pool holds one per equation (pop 128 with 7 tokens: powers up to 5), and the
    targets left over after that go round the population up to three per equation."""
    rng = np.random.default_rng(seed)
    adm, other = _term_pools(n_tokens, max_power)
    while len(adm) < pop:
        max_power += 1
        if max_power > 8:
            raise ValueError(f'no distinct population of {pop} equations over {n_tokens} tokens')
        adm, other = _term_pools(n_tokens, max_power)
    adm = [adm[i] for i in rng.permutation(len(adm))]
    other = [other[i] for i in rng.permutation(len(other))]
    T = rng.integers(t_range[0], t_range[1] + 1, size=pop)
    n_adm = np.ones(pop, dtype=np.int64)
    spare = len(adm) - pop
    for _ in range(2):
        for e in range(pop):
            if spare > 0 and n_adm[e] < T[e] - 1:
                n_adm[e] += 1
                spare -= 1
    if int(np.sum(T - n_adm)) > len(other):
        raise ValueError(f'no distinct population of {pop} equations over {n_tokens} tokens')
    eqs, ia, io = [], 0, 0
    for e in range(pop):
        terms = adm[ia:ia + n_adm[e]] + other[io:io + T[e] - n_adm[e]]
        ia += int(n_adm[e])
        io += int(T[e] - n_adm[e])
        eqs.append([terms[i] for i in rng.permutation(len(terms))])
    return eqs


def make_population_wide(n_tokens, pop, T, seed=0, max_power=3):
    """stress population of T-term equations (T = 16, 30): distinct terms inside an equation."""
    rng = np.random.default_rng(seed)
    eqs = []
    while len(eqs) < pop:
        terms, seen = [], set()
        while len(terms) < T:
            nf = int(rng.integers(1, 4))
            chosen = rng.choice(n_tokens, size=min(nf, n_tokens), replace=False)
            term = [(int(c), float(rng.integers(1, max_power + 1)), 0, 0.0) for c in chosen]
            key = frozenset(term)
            if key in seen:
                continue
            seen.add(key)
            terms.append(term)
        if admissible_targets(terms):
            eqs.append(terms)
    return eqs


def make_population(n_tokens, pop, seed=0, t_fresh=(5, 8), t_pruned=(2, 4), pruned_fraction=0.45,
                    max_power=3):
    """pop equations as nested lists [[(token, power, 0, 0.0), ...] per term].  Token 0 is u, the
    others derivatives; an admissible target is a term with exactly one derivative factor
    (right_part_selection.py:60-63 'contains a derivative of the main variable')."""
    rng = np.random.default_rng(seed)
    eqs = []
    while len(eqs) < pop:
        lo, hi = t_pruned if rng.random() < pruned_fraction else t_fresh
        T = int(rng.integers(lo, hi + 1))
        terms, seen = [], set()
        tries = 0
        while len(terms) < T and tries < 500:
            tries += 1
            nf = 1 if rng.random() < 0.8 else 2
            chosen = rng.choice(n_tokens, size=nf, replace=False)
            term = []
            for c in chosen:
                pmax = max_power if c == 0 else 2
                p = int(rng.integers(1, pmax + 1)) if rng.random() < 0.3 else 1
                term.append((int(c), float(p), 0, 0.0))
            key = frozenset(term)
            if key in seen:
                continue
            seen.add(key)
            terms.append(term)
        if len(terms) == T and admissible_targets(terms):
            eqs.append(terms)
    return eqs


def admissible_targets(eq):
    return [i for i, term in enumerate(eq) if sum(1 for f in term if f[0] != 0) == 1]


def make_evals(eqs):
    """every admissible (equation, target) pair: the D evals EqRightPartSelector runs per equation
    (right_part_selection.py:60-74)."""
    return [(e, t) for e, eq in enumerate(eqs) for t in admissible_targets(eq)]


def distinct_tokens(eq):
    return len({f[0] for term in eq for f in term})
