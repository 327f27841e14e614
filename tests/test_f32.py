"""Single-precision build (QuEST_PREC=1: libQuEST_f32.so).  north_star tolerance for single precision is 1e-5;
the gate kernels are additionally expected to agree with the CPU libraries bit for bit, as in double."""
import os
import subprocess

import numpy as np
import pytest

from quest_b200 import capi, circuits
from tests import harness

CASES = harness.load_golden_cases()


@pytest.fixture(scope="module")
def oracle32():
    if not os.path.isfile(capi.ORACLE_LIB_F32):
        subprocess.run(["make", "-C", os.path.join(capi.REPO_ROOT, "oracle"), "oracle", "PREC=1"], check=True,
                       stdout=subprocess.DEVNULL)
    lib = capi.load(capi.ORACLE_LIB_F32)
    assert lib.prec == 1
    return lib


@pytest.fixture(scope="module")
def reference32():
    if not os.path.isfile(capi.REFERENCE_LIB_F32):
        pytest.skip("oracle/_ref/libQuEST_f32.so not built")
    return capi.load(capi.REFERENCE_LIB_F32)


def _run(lib, nq, dens, ops, seed):
    env = lib.createQuESTEnv()
    q = lib.createDensityQureg(nq, env) if dens else lib.createQureg(nq, env)
    circuits.prepare(lib, q, seed)
    scal = circuits.run_circuit(lib, q, ops, seed=seed)
    st = lib.get_state(q)
    lib.destroyQureg(q, env)
    return st, np.array(scal)


def test_oracle_f32_golden_vectors(oracle32):
    env = oracle32.createQuESTEnv()
    for case in CASES[::3]:
        # angles in the reference files reach hundreds of radians: float trigonometry there is good to ~1e-5 relative
        harness.check_golden_case(case, harness.run_golden_case(oracle32, env, case), tol=2e-5, rtol=5e-5)


@pytest.mark.parametrize("nq,dens,seed", [(10, False, 201), (6, True, 202)])
def test_oracle_f32_bit_exact_vs_reference_f32(oracle32, reference32, nq, dens, seed):
    ops = circuits.random_circuit(nq, 150, seed, density=dens, measure=False)
    a, _ = _run(oracle32, nq, dens, ops, seed)
    b, _ = _run(reference32, nq, dens, ops, seed)
    assert np.array_equal(a.view(np.float32), b.view(np.float32))


@pytest.mark.gpu
@pytest.mark.parametrize("nq,dens,seed", [(3, False, 211), (12, False, 212), (20, False, 213), (3, True, 214), (9, True, 215)])
def test_gpu_f32_bit_exact_vs_oracle_f32(oracle32, nq, dens, seed):
    gpu = capi.load_product(prec=1)
    ops = circuits.random_circuit(nq, 200, seed, density=dens, measure=False)
    a, av = _run(gpu, nq, dens, ops, seed)
    b, bv = _run(oracle32, nq, dens, ops, seed)
    assert np.array_equal(a.view(np.float32), b.view(np.float32))
    np.testing.assert_allclose(av, bv, rtol=1e-5, atol=1e-6)


@pytest.mark.gpu
def test_gpu_f32_with_measurement_vs_reference_f32(reference32):
    gpu = capi.load_product(prec=1)
    ops = circuits.random_circuit(16, 150, 221, measure=True)
    a, av = _run(gpu, 16, False, ops, 221)
    b, bv = _run(reference32, 16, False, ops, 221)
    np.testing.assert_allclose(av, bv, rtol=1e-5, atol=1e-6)
    np.testing.assert_allclose(a, b, rtol=0, atol=1e-5 * np.abs(b).max())
