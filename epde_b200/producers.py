"""Token producers that write straight into the HBM cache.

``fd_derivatives`` = the reference's default preprocessor ``AdaptiveFiniteDeriv``
(epde/preprocessing/deriv_calculators.py:46-102: repeated ``numpy.gradient`` along each axis, non-mixed),
followed by the boundary crop every tensor gets before it enters the cache
(``InputDataEntry.use_global_cache`` interface/interface.py:133-155, ``Cache.set_boundaries``
cache.py:250-270).  The field is uploaded once; every derivative is computed on the full grid on the
device (bit-identical to numpy.gradient) and only its interior is kept, as token vectors of the engine.
torch is used for device memory only.
"""
from __future__ import annotations

import ctypes as C

import numpy as np

from . import _lib


def _i64(values):
    a = np.ascontiguousarray(values, dtype=np.int64)
    return a, a.ctypes.data_as(_lib.c_i64p)


def deriv_labels(var, max_order):
    """cache labels in the order AdaptiveFiniteDeriv returns the derivatives: for every axis with a
    non-zero order d/dx, d^2/dx^2, ... (names as prepared_tokens / interface.py build them)."""
    out = []
    for dim, order in enumerate(max_order):
        for o in range(1, int(order) + 1):
            out.append(f'd{var}/dx{dim}' if o == 1 else f'd^{o}{var}/dx{dim}^{o}')
    return out


def fd_derivatives(engine, data, coords, max_order, boundary, first_token_id=0, field_token_id=None):
    """Differentiate `data` (full grid, numpy array or CUDA torch tensor, float64) and cache the interiors.

    coords: one 1-D coordinate vector per axis (``np.unique`` of the reference's coordinate tensors,
    deriv_calculators.py:96); max_order: int or per-axis orders; boundary: int or per-axis widths (the
    engine's inner shape must equal shape - 2 * boundary).  Tokens first_token_id, first_token_id + 1, ...
    receive the derivatives in ``deriv_labels`` order; `field_token_id` (optional) receives the field itself.
    Returns the list of derivative token ids."""
    import torch
    lib = engine._lib
    dev = torch.device('cuda', engine.device)
    if isinstance(data, np.ndarray):
        field = torch.from_numpy(np.ascontiguousarray(data, dtype=np.float64)).to(dev)
    else:
        field = data.to(dev, torch.float64).contiguous()
    nd = field.dim()
    shape = tuple(int(s) for s in field.shape)
    if isinstance(max_order, int):
        max_order = [max_order] * nd
    if isinstance(boundary, int):
        boundary = [boundary] * nd
    if len(coords) != nd or len(max_order) != nd or len(boundary) != nd:
        raise ValueError('Data dimensionality does not fit passed grids.')      # deriv_calculators.py:59
    coords = [np.ascontiguousarray(c, dtype=np.float64).reshape(-1) for c in coords]
    for d in range(nd):
        if coords[d].size != shape[d]:
            raise ValueError(f'axis {d}: {coords[d].size} coordinates for {shape[d]} points')
    shp, shp_p = _i64(shape)
    bnd, bnd_p = _i64(boundary)
    torch.cuda.current_stream(dev).synchronize()      # the engine's stream may differ from torch's

    def grad(src, axis):
        dst = torch.empty_like(src)
        _lib.check(lib.epde_fd_gradient(engine._h, C.c_void_p(src.data_ptr()), C.c_void_p(dst.data_ptr()), nd, shp_p,
                                        int(axis), coords[axis].ctypes.data_as(_lib.c_f64p)))
        return dst

    def keep(tensor, token_id):
        _lib.check(lib.epde_cache_upload_interior(engine._h, int(token_id), C.c_void_p(tensor.data_ptr()), nd,
                                                  shp_p, bnd_p))

    if field_token_id is not None:
        keep(field, field_token_id)
    ids, nxt = [], int(first_token_id)
    alive = [field]               # temporaries stay allocated until the engine stream has drained
    for dim in range(nd):
        cur = field
        for _ in range(int(max_order[dim])):
            cur = grad(cur, dim)
            alive.append(cur)
            keep(cur, nxt)
            ids.append(nxt)
            nxt += 1
    engine.synchronize()
    del alive
    return ids


def chebyshev_window_weights(coords, order, window=9, n_der=None, poly_order=None):
    """[n][window] weights W with  d^order/dx^order (Chebyshev LSQ fit of the window)(x_i) = sum_j W[i, j] * f[i - h + j].

    Follows get_cheb_for_point / process_point_cheb (preprocessing/cheb.py:14-85) to the letter: polynomial degree
    n_der + 1 (``poly_order`` if given; one less when it equals the window size), ``Chebyshev.fit`` on the raw window
    coordinates (numpy maps them to [-1, 1] and solves the least squares by SVD), derivative of the fitted series at the
    centre coordinate.  The fit is linear in the window values, so fitting the unit vectors gives the functional; rows
    within window // 2 + 1 of either end are left zero (the kernel uses one-sided differences there, fin_der.py)."""
    from numpy.polynomial.chebyshev import Chebyshev
    x = np.ascontiguousarray(coords, dtype=np.float64).reshape(-1)
    n, h = x.size, (window - 1) // 2
    n_der = int(order if n_der is None else n_der)
    deg = (n_der + 1) if poly_order is None else int(poly_order)
    if window == deg:
        deg -= 1
    bound = window // 2 + 1
    W = np.zeros((n, window))
    eye = np.eye(window)
    cache = {}
    for i in range(bound, n - bound + 1):
        xr = x[i - h:i + h + 1]
        key = tuple(np.round((xr - xr[0]) / (xr[-1] - xr[0]), 14))     # equal relative spacing -> same fit up to scale
        span = float(xr[-1] - xr[0])
        ref = cache.get(key)
        if ref is not None:
            # Chebyshev.fit works on the window mapped to [-1, 1]: windows with the same relative spacing have the same
            # fit, and the order-th derivative scales with (1 / span) ** order (a uniform axis needs ONE fit, not n)
            W[i] = ref[0] if ref[1] == span else ref[0] * (ref[1] / span) ** int(order)
            continue
        row = np.array([Chebyshev.fit(xr, eye[j], deg).deriv(m=int(order))(xr[int(xr.size / 2.)]) for j in range(window)])
        W[i] = row
        cache[key] = (row, span)
    return W


def poly_derivatives(engine, data, coords, max_order, boundary, first_token_id=0, field_token_id=None,
                     polynomial_window=9, poly_order=None):
    """The reference's 'poly' preprocessor on the device: ``PolynomialDeriv`` (deriv_calculators.py:111-134 -- a
    4-process pool over grid points, each fitting Chebyshev polynomials through a 9-point window per axis,
    cheb.py:40-85) followed by the boundary crop into the HBM cache.  Same arguments and token order as
    :func:`fd_derivatives`; derivative orders above 3 are not defined by the reference (fin_der.py:33)."""
    import torch
    lib = engine._lib
    dev = torch.device('cuda', engine.device)
    if isinstance(data, np.ndarray):
        field = torch.from_numpy(np.ascontiguousarray(data, dtype=np.float64)).to(dev)
    else:
        field = data.to(dev, torch.float64).contiguous()
    nd = field.dim()
    shape = tuple(int(s) for s in field.shape)
    if isinstance(max_order, int):
        max_order = [max_order] * nd
    if isinstance(boundary, int):
        boundary = [boundary] * nd
    if len(coords) != nd or len(max_order) != nd or len(boundary) != nd:
        raise ValueError('Given derivative orders do not match the data tensor dimensionality')      # cheb.py:55
    coords = [np.ascontiguousarray(c, dtype=np.float64).reshape(-1) for c in coords]
    shp, shp_p = _i64(shape)
    bnd, bnd_p = _i64(boundary)
    torch.cuda.current_stream(dev).synchronize()
    if field_token_id is not None:
        _lib.check(lib.epde_cache_upload_interior(engine._h, int(field_token_id), C.c_void_p(field.data_ptr()), nd,
                                                  shp_p, bnd_p))
    ids, nxt = [], int(first_token_id)
    out = torch.empty_like(field)
    for dim in range(nd):
        for m in range(1, int(max_order[dim]) + 1):
            W = chebyshev_window_weights(coords[dim], m, polynomial_window, n_der=int(max_order[dim]), poly_order=poly_order)
            _lib.check(lib.epde_poly_derivative(engine._h, C.c_void_p(field.data_ptr()), C.c_void_p(out.data_ptr()), nd, shp_p,
                                                int(dim), int(m), int(polynomial_window), W.ctypes.data_as(_lib.c_f64p),
                                                coords[dim].ctypes.data_as(_lib.c_f64p)))
            _lib.check(lib.epde_cache_upload_interior(engine._h, nxt, C.c_void_p(out.data_ptr()), nd, shp_p, bnd_p))
            engine.synchronize()              # `out` is reused by the next derivative
            ids.append(nxt)
            nxt += 1
    return ids
