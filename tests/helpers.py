"""Shared comparison helpers for the parity tests.

Tolerances (BASELINE.json north_star): identical selected term sets; coefficients, fitness and
stability within 1e-6 relative in fp64.
"""
import numpy as np

RTOL = 1e-6


def rel_close(a, b, rtol=RTOL, atol=0.0):
    a = np.asarray(a, dtype=np.float64)
    b = np.asarray(b, dtype=np.float64)
    return np.all(np.abs(a - b) <= atol + rtol * np.maximum(np.abs(a), np.abs(b)))


def assert_rel(a, b, what, rtol=RTOL, atol=0.0):
    a = np.asarray(a, dtype=np.float64)
    b = np.asarray(b, dtype=np.float64)
    assert a.shape == b.shape, f'{what}: shape {a.shape} vs {b.shape}'
    err = np.abs(a - b) / np.maximum(np.maximum(np.abs(a), np.abs(b)), 1e-300)
    bad = np.abs(a - b) > atol + rtol * np.maximum(np.abs(a), np.abs(b))
    assert not bad.any(), f'{what}: max rel err {err[bad].max():.3e} (rtol {rtol}) at {np.argwhere(bad)[:4].tolist()}'


def natural_scales(columns, target):
    """Natural magnitude of each regression coefficient: mean|target| / rms(column), and
    mean|target| for the intercept.  A coefficient (or a whole window-weight column) that is ~0 by
    cancellation carries no relative information below 1e-9 of this."""
    columns = np.asarray(columns, dtype=np.float64)
    ts = float(np.abs(target).mean())
    rms = np.sqrt((columns ** 2).mean(axis=0)) if columns.size else np.zeros(0)
    return np.append(ts / np.maximum(rms, 1e-300), ts)


def assert_sw(sw, ref, what, rtol=RTOL, col_scales=None):
    """sliding-window weight matrices, compared per column with an absolute floor of
    1e-9 x max(largest |entry| of the column, natural scale of that coefficient)."""
    sw = np.asarray(sw, dtype=np.float64)
    ref = np.asarray(ref, dtype=np.float64)
    assert sw.shape == ref.shape, f'{what}: shape {sw.shape} vs {ref.shape}'
    for c in range(ref.shape[1]):
        scale = np.abs(ref[:, c]).max()
        if col_scales is not None:
            scale = max(scale, float(col_scales[c]))
        assert_rel(sw[:, c], ref[:, c], f'{what} col {c}', rtol=rtol, atol=1e-9 * scale)
