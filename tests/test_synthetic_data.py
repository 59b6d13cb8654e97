"""CPU tests of the bench's synthetic token generator (epde_b200/synthetic.py): the torch path draws its noise with
a counter-based generator so that the dataset does not depend on the device and a slab of the grid (grid-point
sharding) holds exactly the values of the whole grid."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

from epde_b200 import synthetic as syn  # noqa: E402


def test_counter_noise_is_standard_normal_and_reproducible():
    import torch
    z = syn._counter_normal(torch, 0, 400_000, 5, 3, None)
    assert torch.isfinite(z).all()
    assert abs(float(z.mean())) < 0.01 and abs(float(z.std()) - 1.0) < 0.01
    kurt = float(((z - z.mean()) ** 4).mean() / z.var() ** 2)
    assert abs(kurt - 3.0) < 0.1
    lag1 = float(torch.corrcoef(torch.stack([z[:-1], z[1:]]))[0, 1])
    assert abs(lag1) < 0.01
    # pure function of (seed, token, index): any window of the sequence can be drawn on its own
    w = syn._counter_normal(torch, 12345, 1000, 5, 3, None)
    assert torch.equal(w, z[12345:13345])
    other = syn._counter_normal(torch, 0, 1000, 5, 4, None)
    assert abs(float(torch.corrcoef(torch.stack([z[:1000], other]))[0, 1])) < 0.15
    assert not torch.equal(syn._counter_normal(torch, 0, 1000, 6, 3, None), z[:1000])


def test_slab_tokens_equal_rows_of_the_whole_grid():
    import torch
    shape = (24, 12, 10)
    row = shape[1] * shape[2]
    for kind in ('heat', 'wave'):
        shp = shape if kind == 'heat' else shape[:2]
        rw = int(np.prod(shp[1:]))
        full = syn.make_tokens(shp, kind, xp=torch, seed=5)
        for lo, hi in [(0, 7), (7, 24)]:
            part = syn.make_tokens(shp, kind, xp=torch, seed=5, axis0_range=(lo, hi))
            for f, p in zip(full, part):
                assert torch.equal(f[lo * rw:hi * rw], p)
    # the numpy path (fixtures, CPU baseline) is slab-consistent as well
    full = syn.make_tokens(shape, 'heat', seed=5)
    part = syn.make_tokens(shape, 'heat', seed=5, axis0_range=(7, 24))
    for f, p in zip(full, part):
        assert np.array_equal(f[7 * row:], p)


def test_distinct_population_is_bounded_and_shares_nothing():
    """the all-distinct stress population (bench --population distinct): generation is a bounded deal from the
    enumerated term pools (the round-1 rejection loop never ended for 128 equations over 7 tokens), no term appears
    twice, every equation has an admissible target."""
    from epde_b200 import synthetic as syn
    for n_tokens, pop, seed in ((7, 128, 5), (7, 48, 2), (5, 96, 1)):
        eqs = syn.make_population_distinct(n_tokens, pop, seed=seed)
        terms = [frozenset(t) for eq in eqs for t in eq]
        assert len(eqs) == pop and len(set(terms)) == len(terms)
        assert all(3 <= len(eq) <= 7 and syn.admissible_targets(eq) for eq in eqs)
        assert eqs == syn.make_population_distinct(n_tokens, pop, seed=seed)
