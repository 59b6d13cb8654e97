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
