"""Reference scripted/essential tests (tests/scripted.py) on the CPU oracle; the CUDA library runs the same
checks in tests/test_gpu_parity.py."""
import pytest

from tests import scripted


def test_seed_known_answer(oracle):
    scripted.check_seed_known_answer(oracle)


@pytest.mark.parametrize("check", scripted.ALL_CHECKS, ids=lambda f: f.__name__)
def test_scripted_on_oracle(oracle, oracle_env, check):
    check(oracle, oracle_env)


@pytest.mark.parametrize("check", scripted.ALL_CHECKS, ids=lambda f: f.__name__)
def test_scripted_on_reference(reference, reference_env, check):
    """The restated checks are themselves validated against the reference's own library."""
    check(reference, reference_env)
