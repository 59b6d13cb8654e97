"""GPU tests of the device token producer (epde_b200.producers.fd_derivatives) against numpy.gradient --
the arithmetic the reference's AdaptiveFiniteDeriv runs (preprocessing/deriv_calculators.py:46-102) -- followed
by the boundary crop.  Bar: bit-exact (the kernel rounds products and sums like numpy)."""
import numpy as np
import pytest

import epde_b200
from epde_b200 import producers

pytestmark = pytest.mark.gpu


def numpy_chain(data, coords, max_order, boundary):
    """AdaptiveFiniteDeriv.differentiate, non-mixed (deriv_calculators.py:50-88), then the interior crop."""
    crop = tuple(slice(b, n - b) for b, n in zip(boundary, data.shape))
    out = []
    for dim in range(data.ndim):
        cur = data
        for _ in range(max_order[dim]):
            cur = np.gradient(cur, coords[dim], axis=dim)
            out.append(np.ascontiguousarray(cur[crop]).reshape(-1))
    return out


@pytest.mark.parametrize('shape,boundary,max_order,uniform', [
    ((81, 81), (10, 10), (2, 2), False), ((64, 48), (0, 0), (1, 2), True), ((33, 20, 27), (4, 4, 4), (1, 2, 2), False),
    ((40, 36, 30), (0, 3, 5), (2, 0, 3), True), ((2, 5), (0, 0), (1, 1), False)])
def test_fd_derivatives_bit_exact(shape, boundary, max_order, uniform):
    rng = np.random.default_rng(sum(shape))
    if uniform:
        coords = [np.arange(n, dtype=np.float64) * 0.125 for n in shape]          # exactly constant spacing
    else:
        coords = [np.linspace(0.0, 1.0 + d, n) if d % 2 == 0 else np.sort(rng.random(n)) * 3 for d, n in enumerate(shape)]
    grids = np.meshgrid(*coords, indexing='ij')
    data = np.sin(2.0 * grids[0]) * np.cos(grids[-1]) + 0.05 * rng.standard_normal(shape)
    inner = tuple(n - 2 * b for n, b in zip(shape, boundary))
    ref = numpy_chain(data, coords, max_order, boundary)
    with epde_b200.Engine(inner, device=0) as eng:
        ids = producers.fd_derivatives(eng, data, coords, max_order, boundary, first_token_id=1, field_token_id=0)
        assert ids == list(range(1, 1 + len(ref)))
        crop = tuple(slice(b, n - b) for b, n in zip(boundary, shape))
        assert np.array_equal(eng.download(0), data[crop].reshape(-1))
        for i, r in zip(ids, ref):
            got = eng.download(i)
            assert np.array_equal(got, r), f'token {i}: max abs diff {np.abs(got - r).max()}'
        assert eng.launch_count >= 2 * len(ref)


def test_fd_labels_and_errors():
    assert producers.deriv_labels('u', (2, 2)) == ['du/dx0', 'd^2u/dx0^2', 'du/dx1', 'd^2u/dx1^2']
    with epde_b200.Engine((6, 6), device=0) as eng:
        with pytest.raises(ValueError):
            producers.fd_derivatives(eng, np.zeros((8, 8)), [np.arange(8.0)], 1, 1)
        with pytest.raises(epde_b200.EpdeB200Error):       # 8 - 2*2 != 6
            producers.fd_derivatives(eng, np.zeros((8, 8)), [np.arange(8.0), np.arange(8.0)], 1, 2)


def test_produced_tokens_feed_the_evaluator():
    """end to end: field -> device derivatives -> cache -> evaluate == upload of numpy derivatives -> evaluate."""
    shape, b = (60, 50), 5
    t, x = np.linspace(0, 1, shape[0]), np.linspace(0, 2, shape[1])
    T, X = np.meshgrid(t, x, indexing='ij')
    u = np.exp(-T) * np.sin(3 * X) + 1e-3 * np.random.default_rng(0).standard_normal(shape)
    inner = (shape[0] - 2 * b, shape[1] - 2 * b)
    eqs = [[[(1, 1.0, 0, 0.0)], [(4, 1.0, 0, 0.0)], [(0, 1.0, 0, 0.0), (3, 1.0, 0, 0.0)], [(2, 1.0, 0, 0.0)]]]
    enc = epde_b200.encode(eqs, [(0, 0), (0, 3)])
    with epde_b200.Engine(inner, device=0) as eng:
        producers.fd_derivatives(eng, u, [t, x], (2, 2), b, first_token_id=1, field_token_id=0)
        a = eng.evaluate(enc)
    ref = numpy_chain(u, [t, x], (2, 2), (b, b))
    with epde_b200.Engine(inner, device=0) as eng:
        eng.upload(0, u[b:-b, b:-b])
        for i, r in enumerate(ref):
            eng.upload(i + 1, r)
        c = eng.evaluate(enc)
    for k in ('coef', 'intercept', 'fitness_l1', 'fitness_l2', 'stability'):
        assert np.array_equal(getattr(a, k), getattr(c, k)), k


def test_poly_derivatives_match_reference_fixture():
    """device Chebyshev producer (producers.poly_derivatives -> epde_poly_derivative) on the noisy Burgers field ==
    the derivative tokens the reference's 'poly' preprocessor (PolynomialDeriv, deriv_calculators.py:111-134) produced
    for BASELINE config 2 (tests/golden/burgers_poly_b50.npz): 1e-10 relative to the token's scale."""
    from conftest import load_fixture
    fx = load_fixture('burgers_poly_b50')
    u, t, x, b = fx['data_full'], fx['axis0'], fx['axis1'], int(fx['boundary'])
    labels = [str(l) for l in fx['labels']]
    inner = tuple(int(s) for s in fx['shape'])
    with epde_b200.Engine(inner, device=0) as eng:
        ids = producers.poly_derivatives(eng, u, [t, x], (1, 2), b, first_token_id=1, field_token_id=0)
        names = producers.deriv_labels('u', (1, 2))
        assert np.array_equal(eng.download(0), fx['tokens'][labels.index('u')])
        for i, lab in zip(ids, names):
            ref = fx['tokens'][labels.index(lab)]
            got = eng.download(i)
            assert np.abs(got - ref).max() <= 1e-10 * np.abs(ref).max(), (lab, np.abs(got - ref).max(), np.abs(ref).max())


@pytest.mark.parametrize('shape,max_order', [((24, 30), (2, 3)), ((14, 16, 18), (1, 2, 2))])
def test_poly_derivatives_with_boundary_rows(shape, max_order):
    """boundary 0: the rows within window // 2 + 1 of the ends use the one-sided differences of fin_der.py; checked
    against the point-by-point restatement of cheb.py / fin_der.py (oracle.poly_derivative_points), non-uniform axes."""
    from oracle import epde_oracle as orc
    rng = np.random.default_rng(5)
    coords = [np.sort(rng.random(n)) * (2 + d) if d == 1 else np.linspace(0, 1 + d, n) for d, n in enumerate(shape)]
    grids = np.meshgrid(*coords, indexing='ij')
    data = np.sin(2.0 * grids[0]) * np.cos(grids[-1]) + 0.02 * rng.standard_normal(shape)
    with epde_b200.Engine(shape, device=0) as eng:
        ids = producers.poly_derivatives(eng, data, coords, max_order, 0, first_token_id=0)
        k = 0
        for dim in range(len(shape)):
            for m in range(1, max_order[dim] + 1):
                ref = orc.poly_derivative_points(data, coords[dim], dim, m, max_order[dim]).reshape(-1)
                got = eng.download(ids[k])
                assert np.abs(got - ref).max() <= 1e-9 * np.abs(ref).max(), (dim, m, np.abs(got - ref).max())
                k += 1
