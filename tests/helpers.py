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


def assert_sw(sw, ref, what, rtol=RTOL):
    """sliding-window weight matrices: per column, absolute floor 1e-9 x the column's largest
    magnitude (entries that are ~0 by cancellation carry no relative information)."""
    sw = np.asarray(sw, dtype=np.float64)
    ref = np.asarray(ref, dtype=np.float64)
    assert sw.shape == ref.shape, f'{what}: shape {sw.shape} vs {ref.shape}'
    for c in range(ref.shape[1]):
        assert_rel(sw[:, c], ref[:, c], f'{what} col {c}', rtol=rtol, atol=1e-9 * np.abs(ref[:, c]).max())
