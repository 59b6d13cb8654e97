"""CPU run of the operator-mirror tests of tests/test_gpu_operators.py with the CUDA engine replaced by the
oracle-backed stand-in (tests/oracle_engine.py): checks the host logic of epde_b200.operators (selector loop,
Equation mutations, error behaviour, cache mirror) against the reference-generated golden fixtures where no GPU
exists.  The same functions run against the real CUDA engine under ``-m gpu``."""
import pytest

import epde_b200
from oracle_engine import OracleEngine

import test_gpu_operators as T
from test_gpu_operators import bound_fixture  # noqa: F401  (fixture)
from conftest import EQ_FIXTURES


class _CtxOracleEngine(OracleEngine):
    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()


@pytest.fixture(autouse=True)
def _oracle_engine(monkeypatch):
    monkeypatch.setattr(epde_b200, 'Engine', _CtxOracleEngine)
    yield


@pytest.mark.parametrize('bound_fixture', EQ_FIXTURES, indirect=True)
def test_right_part_selector_matches_reference(bound_fixture):
    T.test_right_part_selector_matches_reference(bound_fixture)


@pytest.mark.parametrize('bound_fixture', ['wave_b5', 'heat3d_b4'], indirect=True)
def test_out_of_place_fitness_matches_reference(bound_fixture):
    T.test_out_of_place_fitness_matches_reference(bound_fixture)


@pytest.mark.parametrize('bound_fixture', ['wave_b5'], indirect=True)
def test_population_batch_equals_one_by_one(bound_fixture):
    T.test_population_batch_equals_one_by_one(bound_fixture)


def test_l2_fitness_and_penalty():
    T.test_l2_fitness_and_penalty()


def test_operator_contract_and_errors():
    T.test_operator_contract_and_errors()


def test_cache_mirror_and_trig_family():
    T.test_cache_mirror_and_trig_family()
