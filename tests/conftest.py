import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

GOLDEN = os.path.join(ROOT, 'tests', 'golden')
EQ_FIXTURES = ['wave_b5', 'wave_b10', 'wave_b8', 'burgers_b50', 'heat3d_b4', 'burgers_poly_b50', 'heat3d_b3']


def pytest_configure(config):
    config.addinivalue_line('markers', 'gpu: needs a CUDA device (run on the B200 box)')


def load_fixture(name):
    z = np.load(os.path.join(GOLDEN, name + '.npz'))
    return {k: z[k] for k in z.files}


def fixture_equations(fx):
    """golden CSR -> [[[(token, power, 0, 0.0)]]] nested encoding."""
    eqs = []
    for e in range(len(fx['term_off']) - 1):
        terms = []
        for t in range(fx['term_off'][e], fx['term_off'][e + 1]):
            terms.append([(int(fx['fac_tok'][f]), float(fx['fac_pow'][f]), 0, 0.0)
                          for f in range(fx['fac_off'][t], fx['fac_off'][t + 1])])
        eqs.append(terms)
    return eqs


@pytest.fixture(scope='session')
def has_gpu():
    try:
        import torch
        return torch.cuda.is_available()
    except Exception:       # noqa: BLE001
        return False


def fixture_equations_procedural(fx):
    """golden CSR with factor kinds / params -> nested encoding [[[(token, power, kind, param)]]]."""
    eqs = []
    for e in range(len(fx['term_off']) - 1):
        terms = []
        for t in range(fx['term_off'][e], fx['term_off'][e + 1]):
            terms.append([(int(fx['fac_tok'][f]), float(fx['fac_pow'][f]), int(fx['fac_kind'][f]), float(fx['fac_par'][f]))
                          for f in range(fx['fac_off'][t], fx['fac_off'][t + 1])])
        eqs.append(terms)
    return eqs
