"""GPU test of the allocation-free entry point ``epde_poly_derivative_noalloc`` (include/epde_b200.h; no counterpart
in the reference): through the C ABI, on two calls in a row, its output
equals ``epde_poly_derivative``'s bit for bit (same kernel, same weights; only the home of the scratch differs), and a
request beyond the fixed scratch is refused with a message that names the limit.  The file name sorts last on purpose."""
import ctypes as C

import numpy as np
import pytest

import epde_b200
from epde_b200 import _lib, producers

pytestmark = pytest.mark.gpu

NOALLOC_MAX_DOUBLES = 65536          # EPDE_NOALLOC_MAX_DOUBLES


def _call(fn, eng, field, out, shape, axis, order, window, W, x):
    shp = (C.c_int64 * len(shape))(*shape)
    _lib.check(fn(eng._h, C.c_void_p(field.data_ptr()), C.c_void_p(out.data_ptr()), len(shape), shp, int(axis), int(order),
                  int(window), W.ctypes.data_as(_lib.c_f64p), x.ctypes.data_as(_lib.c_f64p)))
    eng.synchronize()
    return out.cpu().numpy().copy()


def test_noalloc_poly_derivative_matches_existing_entry_point_twice():
    import torch
    lib = _lib.load()
    dev = torch.device('cuda', 0)
    shape = (96, 70, 33)
    rng = np.random.default_rng(11)
    coords = [np.linspace(0.0, 2.0, shape[0]), np.sort(rng.random(shape[1])) * 3.0, np.linspace(-1.0, 1.0, shape[2])]
    grids = np.meshgrid(*coords, indexing='ij')
    with epde_b200.Engine(shape, device=0) as eng:
        # two calls in a row on DIFFERENT problems (axis, order, window, data): the second must not see the first's scratch
        for call, (axis, order, window) in enumerate([(1, 2, 9), (0, 1, 7)]):
            data = np.sin(2.0 * grids[0] + call) * np.cos(grids[1]) * (1.0 + grids[2]) + 0.01 * rng.standard_normal(shape)
            field = torch.from_numpy(data).to(dev)
            W = producers.chebyshev_window_weights(coords[axis], order, window, n_der=3)
            W = np.ascontiguousarray(W, dtype=np.float64)
            assert W.size + shape[axis] <= NOALLOC_MAX_DOUBLES
            want = _call(lib.epde_poly_derivative, eng, field, torch.empty_like(field), shape, axis, order, window, W, coords[axis])
            out = torch.full_like(field, np.nan)
            torch.cuda.synchronize()
            mem_before = torch.cuda.mem_get_info(0)[0]
            got = _call(lib.epde_poly_derivative_noalloc, eng, field, out, shape, axis, order, window, W, coords[axis])
            assert torch.cuda.mem_get_info(0)[0] == mem_before, 'the allocation-free entry point changed free device memory'
            assert np.all(np.isfinite(got))
            np.testing.assert_array_equal(got, want)


def test_noalloc_poly_derivative_rejects_sizes_beyond_the_scratch():
    import torch
    lib = _lib.load()
    n, window = 8192, 9                                  # 8192 * 10 doubles > 65536
    assert n * (window + 1) > NOALLOC_MAX_DOUBLES
    shape = (n, 4)
    x = np.linspace(0.0, 1.0, n)
    W = np.zeros((n, window))
    field = torch.zeros(shape, dtype=torch.float64, device='cuda:0')
    out = torch.zeros_like(field)
    with epde_b200.Engine(shape, device=0) as eng:
        shp = (C.c_int64 * 2)(*shape)
        rc = lib.epde_poly_derivative_noalloc(eng._h, C.c_void_p(field.data_ptr()), C.c_void_p(out.data_ptr()), 2, shp, 0, 1,
                                              window, W.ctypes.data_as(_lib.c_f64p), x.ctypes.data_as(_lib.c_f64p))
        assert rc != 0
        msg = lib.epde_last_error().decode()
        assert 'scratch' in msg and str(NOALLOC_MAX_DOUBLES) in msg
        # the growing-buffer entry point takes the same request
        _lib.check(lib.epde_poly_derivative(eng._h, C.c_void_p(field.data_ptr()), C.c_void_p(out.data_ptr()), 2, shp, 0, 1,
                                            window, W.ctypes.data_as(_lib.c_f64p), x.ctypes.data_as(_lib.c_f64p)))
        eng.synchronize()
