"""Pins the CPU oracle (oracle/quest_oracle.c): against every golden vector of the reference's suite, against the
reference library's full-precision outputs, and against the reference library itself on seeded circuits."""
import os

import numpy as np
import pytest

from quest_b200 import circuits
from tests import harness

CASES = harness.load_golden_cases()
REF_OUT = np.load(os.path.join(harness.GOLDEN_DIR, "ref_unit_outputs.npz"))
REF_CIRC = np.load(os.path.join(harness.GOLDEN_DIR, "ref_circuits.npz"))


def test_golden_inventory():
    funcs = {c["func"] for c in CASES}
    # every gate / reduction / noise entry point the north star names has golden vectors
    for f in ["hadamard", "controlledNot", "compactUnitary", "unitary", "multiControlledUnitary", "rotateX", "rotateY",
              "rotateZ", "phaseShift", "multiControlledPhaseShift", "applyOneQubitDephaseError",
              "applyOneQubitDepolariseError", "applyOneQubitDampingError", "calcProbOfOutcome", "collapseToOutcome",
              "calcTotalProb", "calcPurity"]:
        assert f in funcs, f
    assert len(CASES) == 768


def test_oracle_matches_reference_golden_vectors(oracle, oracle_env):
    """All 768 standard-format cases of the reference suite at its own tolerance (1e-10 absolute)."""
    for case in CASES:
        harness.check_golden_case(case, harness.run_golden_case(oracle, oracle_env, case))


def test_oracle_bit_exact_vs_reference_outputs(oracle, oracle_env):
    """Gate outputs equal the reference library's float64 results bit for bit (no FMA, same operand order).
    Cases whose op contains a reduction feeding the state (collapse) are compared at 1e-14 instead."""
    n_exact = 0
    for i, case in enumerate(CASES):
        if f"s{i}" not in REF_OUT.files:
            continue
        got = harness.run_golden_case(oracle, oracle_env, case)["S"]
        ref = REF_OUT[f"s{i}"]
        if case["func"] == "collapseToOutcome":
            np.testing.assert_allclose(got, ref, atol=1e-14, rtol=1e-14)
        else:
            assert np.array_equal(got.view(np.float64), ref.view(np.float64)), case["file"]
            n_exact += 1
    assert n_exact >= 500


@pytest.mark.parametrize("name,nq,dens,depth,seed", [("sv12", 12, False, 120, 11), ("sv16", 16, False, 160, 12),
                                                     ("dm6", 6, True, 90, 13), ("dm8", 8, True, 120, 14)])
def test_oracle_matches_reference_circuit_fixtures(oracle, oracle_env, name, nq, dens, depth, seed):
    """Seeded random circuits over every gate family: final state, digests and every scalar (probabilities,
    measurement outcomes, purity) against fixtures the reference library produced."""
    ops = circuits.random_circuit(nq, depth, seed, density=dens)
    q = oracle.createDensityQureg(nq, oracle_env) if dens else oracle.createQureg(nq, oracle_env)
    circuits.prepare(oracle, q, seed)
    scal = circuits.run_circuit(oracle, q, ops, seed=seed)
    st = oracle.get_state(q)
    oracle.destroyQureg(q, oracle_env)
    np.testing.assert_allclose(scal, REF_CIRC[name + "_scalars"], rtol=1e-12, atol=1e-13)
    sub = st if st.size <= 1 << 16 else st[:: st.size >> 16]
    scale = np.abs(REF_CIRC[name + "_state"]).max()
    np.testing.assert_allclose(sub, REF_CIRC[name + "_state"], rtol=0, atol=1e-12 * scale)
    np.testing.assert_allclose(circuits.state_digest(st), REF_CIRC[name + "_digest"], rtol=1e-11, atol=1e-12)


@pytest.mark.parametrize("nq,dens,seed", [(10, False, 101), (14, False, 102), (5, True, 103), (7, True, 104)])
def test_oracle_vs_live_reference(oracle, oracle_env, reference, reference_env, nq, dens, seed):
    """Same circuit through oracle and oracle/_ref side by side; measurement-free, so the states must be
    identical to the bit."""
    ops = circuits.random_circuit(nq, 150, seed, density=dens, measure=False)
    states = []
    for lib, env in ((oracle, oracle_env), (reference, reference_env)):
        q = lib.createDensityQureg(nq, env) if dens else lib.createQureg(nq, env)
        circuits.prepare(lib, q, seed)
        circuits.run_circuit(lib, q, ops, seed=seed)
        states.append(lib.get_state(q))
        lib.destroyQureg(q, env)
    assert np.array_equal(states[0].view(np.float64), states[1].view(np.float64))


@pytest.mark.parametrize("nq,seed", [(4, 401), (6, 402)])
def test_api_fuzz_oracle_vs_live_reference(oracle, oracle_env, reference, reference_env, nq, seed):
    """Random multi-register API sequences (gates, noise, copies, observers): oracle and reference agree bit for bit
    (the sequence avoids initPureState, whose reference implementation has a known sign slip)."""
    calls = circuits.random_api_sequence(nq, 300, seed)
    a, sa = circuits.run_api_sequence(oracle, oracle_env, nq, calls, seed)
    b, sb = circuits.run_api_sequence(reference, reference_env, nq, calls, seed)
    np.testing.assert_allclose(a, b, rtol=1e-12, atol=1e-13)
    for k in sa:
        np.testing.assert_allclose(sa[k], sb[k], rtol=0, atol=1e-13 * max(1.0, np.abs(sb[k]).max()))
