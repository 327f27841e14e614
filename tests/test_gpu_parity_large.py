"""Element-wise parity at the sizes the benchmark runs at: every kernel class at 28 qubits (14-qubit density matrices)
against the reference's own CPU/OpenMP build (oracle/_ref), amplitude by amplitude.

SURVEY.md 8(c): the reference's 3-qubit golden vectors pin semantics, not the large-index / high-target code paths
(strides of 2^24..2^27 amplitudes, byte offsets beyond 2^31, the pair loop of QuEST_cpu.c:1590-1625 at high index
bits).  These tests walk exactly those paths.  Measurement-free gates must match BIT FOR BIT (same operand order, no
FMA); reductions and collapses to 1e-12 relative (north_star).  bench.py repeats the comparison at 30 qubits on the
benchmark circuit itself (its `parity` object).
"""
import numpy as np
import pytest

from quest_b200 import circuits
from tests import harness

pytestmark = pytest.mark.gpu

NQ = 28
A, B = circuits.SWEEP_ALPHA, circuits.SWEEP_BETA
U = circuits._unitary(__import__("random").Random(28))


def _pair(product, product_env, reference, reference_env, nq=NQ, density=False, normalised=False):
    regs = []
    for lib, env in ((product, product_env), (reference, reference_env)):
        q = lib.createDensityQureg(nq, env) if density else lib.createQureg(nq, env)
        if normalised or density:
            lib.initPlusState(q)
            for t in (0, 1, nq // 2, nq - 2, nq - 1):
                lib.rotateY(q, t, 0.37 * (t + 1))
                lib.rotateZ(q, t, 0.11 * (t + 2))
        else:
            lib.initStateDebug(q)  # every amplitude different, cheap on the CPU side
        regs.append(q)
    return regs


def _run_both(product, reference, gq, cq, ops, seed=None):
    got = circuits.run_circuit(product, gq, ops, seed=seed)
    want = circuits.run_circuit(reference, cq, ops, seed=seed)
    return np.array(got), np.array(want)


def _finish(product, product_env, reference, reference_env, gq, cq, bit_exact):
    rel, same = harness.compare_with_cpu_register(product, gq, cq)
    product.destroyQureg(gq, product_env)
    reference.destroyQureg(cq, reference_env)
    assert rel <= 1e-12, rel
    if bit_exact:
        assert same


@pytest.mark.parametrize("fusion", [0, 1], ids=["gate_at_a_time", "fused"])
def test_dense_gates_on_high_and_low_targets(product, product_env, reference, reference_env, fusion):
    """k_pairs with the target above the vector at strides 2^24..2^27 and inside the vector (targets 0, 1); with
    fusion on the same list runs as fused passes over {24..27} and {0, 1, 2, 27}."""
    ops = []
    for t in (24, 25, 26, 27, 0, 1, 2, 27):
        ops += [("hadamard", t), ("compactUnitary", t, A, B), ("controlledNot", (t + 1) % NQ, t), ("pauliY", t),
                ("unitary", t, U), ("controlledCompactUnitary", (t + 5) % NQ, t, A, B)]
    gq, cq = _pair(product, product_env, reference, reference_env)
    product.lib.qb200_setFusion(fusion)
    try:
        _run_both(product, reference, gq, cq, ops)
        _finish(product, product_env, reference, reference_env, gq, cq, bit_exact=True)
    finally:
        product.lib.qb200_setFusion(1)


def test_fused_groups_spanning_lowest_and_highest_qubits(product, product_env, reference, reference_env):
    """Fused passes whose qubits are {0, 1, 26, 27} and {24, 25, 26, 27}: element offsets of up to 2^27 amplitudes
    (1 GiB) inside one thread's block, controls and phase masks both inside and outside the fused set."""
    ops = []
    for group in ((0, 1, 26, 27), (24, 25, 26, 27)):
        for t in group:
            ops += [("hadamard", t), ("compactUnitary", t, A, B)]
        ops += [("controlledNot", group[2], group[0]), ("controlledNot", group[1], group[3]),
                ("controlledNot", 13, group[2]), ("multiControlledPhaseShift", [group[0], group[2], group[3]], 0.7),
                ("multiControlledPhaseShift", [group[1], 9, 17], -0.4), ("controlledPhaseFlip", group[3], 5),
                ("multiControlledUnitary", [group[0], 11], group[3], U), ("rotateX", group[1], 0.9), ("tGate", group[2])]
    gq, cq = _pair(product, product_env, reference, reference_env)
    _run_both(product, reference, gq, cq, ops)
    _finish(product, product_env, reference, reference_env, gq, cq, bit_exact=True)


def test_diagonal_gates(product, product_env, reference, reference_env):
    """k_diag: one to three selecting bits, inside the vector and at the top of the index."""
    ops = [("phaseShift", 27, 0.3), ("tGate", 0), ("sGate", 1), ("pauliZ", 26), ("controlledPhaseShift", 27, 1, 0.2),
           ("controlledPhaseFlip", 0, 25), ("multiControlledPhaseFlip", [24, 26, 3]),
           ("multiControlledPhaseShift", [0, 1, 27], 0.7), ("multiControlledPhaseShift", [25, 26, 27], -1.1)]
    gq, cq = _pair(product, product_env, reference, reference_env)
    product.lib.qb200_setFusion(0)
    try:
        _run_both(product, reference, gq, cq, ops)
        _finish(product, product_env, reference, reference_env, gq, cq, bit_exact=True)
    finally:
        product.lib.qb200_setFusion(1)


def test_probabilities_and_collapse(product, product_env, reference, reference_env):
    """k_probZero (first query), k_probAllZero (later queries), k_sumSquares, k_collapse on qubits inside the vector,
    in the middle and at the top; measurement outcomes under the same seed.

    Tolerances.  The state is a product state, so every P(qubit = 0) has a closed form, (1 - sin theta)/2: the product
    must match it to 1e-12.  The reference adds |amp|^2 sequentially per OpenMP thread (QuEST_cpu.c:2880-2887): 2^24
    equal-sized terms per thread round the same way and leave it ~1e-10 off the closed form at this size (measured:
    1.25e-10 on 16 threads), and its collapse renormalises with that value.  Against the reference the scalars are
    therefore held to 1e-12 plus the reference's own defect, which its Kahan-summed calcTotalProb after the collapse
    reports (QuEST_cpu_local.c:123-151); amplitudes likewise."""
    gq, cq = _pair(product, product_env, reference, reference_env, normalised=True)
    probes = [(27, 0), (0, 1), (1, 0), (5, 1), (16, 0), (26, 1), (2, 0)]
    ops = [("calcProbOfOutcome", t, o) for t, o in probes]
    ops += [("calcTotalProb",), ("collapseToOutcome", 27, 1), ("calcTotalProb",), ("collapseToOutcome", 0, 0),
            ("measure", 14), ("measureWithStats", 26), ("calcProbOfOutcome", 27, 1), ("calcTotalProb",)]
    got, want = _run_both(product, reference, gq, cq, ops, seed=28)
    rotated = (0, 1, NQ // 2, NQ - 2, NQ - 1)
    exact = []
    for t, o in probes:
        p0 = (1 - np.sin(0.37 * (t + 1))) / 2 if t in rotated else 0.5
        exact.append(p0 if o == 0 else 1 - p0)
    np.testing.assert_allclose(got[:len(probes)], exact, rtol=1e-12)
    assert abs(got[len(probes)] - 1) < 1e-12 and abs(got[-1] - 1) < 1e-12   # the product's own norm, before and after
    ref_defect = max(abs(want[len(probes) + 2] - 1), abs(want[-1] - 1), float(np.abs(want[:len(probes)] - exact).max()))
    assert ref_defect < 2e-9  # 2^24 terms per thread x 2^-53: the bound on the reference's sequential sums
    assert [int(x) for x in got[-4:-2]] == [int(x) for x in want[-4:-2]]     # measure / measureWithStats outcomes
    np.testing.assert_allclose(got, want, rtol=1e-12 + 4 * ref_defect, atol=1e-13)
    rel, _ = harness.compare_with_cpu_register(product, gq, cq)
    product.destroyQureg(gq, product_env)
    reference.destroyQureg(cq, reference_env)
    assert rel <= 1e-12 + 4 * ref_defect, (rel, ref_defect)


@pytest.mark.parametrize("fusion", [0, 1], ids=["gate_at_a_time", "fused"])
def test_noise_channels_on_a_14_qubit_density_matrix(product, product_env, reference, reference_env, fusion):
    """k_noise1 / the noise tier of the fused pass at 2^28 elements: row bit inside the vector (qubit 0, 1), column bit
    at the top of the index (qubit 13 -> bit 27), plus the unitaries (two passes each) and purity."""
    nq = 14
    ops = []
    for t in (13, 0, 1, 7, 12):
        ops += [("hadamard", t), ("rotateY", t, 0.2 + 0.1 * t), ("controlledNot", t, (t + 1) % nq),
                ("applyOneQubitDephaseError", t, 0.02), ("applyOneQubitDepolariseError", t, 0.03),
                ("applyOneQubitDampingError", t, 0.05)]
    ops += [("calcPurity",), ("calcProbOfOutcome", 13, 1), ("calcTotalProb",)]
    gq, cq = _pair(product, product_env, reference, reference_env, nq=nq, density=True)
    product.lib.qb200_setFusion(fusion)
    try:
        got, want = _run_both(product, reference, gq, cq, ops)
        # The reference adds the 2^28 squares of calcPurity sequentially per OpenMP thread (QuEST_cpu.c:854-880): 2^24
        # terms per thread leave it up to ~2e-9 off (measured 2.4e-11), so the scalars are held to that bound against the
        # reference and the product's purity to 1e-12 against a pairwise (numpy) sum over its own amplitudes; the
        # amplitudes themselves must still agree bit for bit.
        np.testing.assert_allclose(got, want, rtol=2e-9, atol=1e-13)
        np.testing.assert_allclose(got[0], harness.sum_squares_chunked(product, gq), rtol=1e-12)
        _finish(product, product_env, reference, reference_env, gq, cq, bit_exact=True)
    finally:
        product.lib.qb200_setFusion(1)
