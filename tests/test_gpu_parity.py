"""GPU parity: the CUDA library (through its C ABI) against the reference's golden vectors, the CPU oracle and
the reference's own CPU build, on the same seeded inputs.

Bars (BASELINE.json north_star): amplitudes within 1e-12 relative in double, measurement outcomes bit-exact
under the same seed.  Gate kernels are compiled without FMA contraction and written with the reference's
operand order, so measurement-free circuits are additionally required to match BIT FOR BIT.
"""
import ctypes
import os

import numpy as np
import pytest

from quest_b200 import circuits
from tests import harness, scripted

pytestmark = pytest.mark.gpu

CASES = harness.load_golden_cases()
REF_OUT = np.load(os.path.join(harness.GOLDEN_DIR, "ref_unit_outputs.npz"))
REF_CIRC = np.load(os.path.join(harness.GOLDEN_DIR, "ref_circuits.npz"))
REL = 1e-12


def test_native_library_is_the_one_loaded(product, product_env):
    assert product.path.endswith("quest_b200/lib/libQuEST.so")
    before = product.lib.qb200_getLaunchCount()
    q = product.createQureg(10, product_env)
    product.hadamard(q, 3)
    product.destroyQureg(q, product_env)
    product.lib.qb200_getLaunchCount.restype = __import__("ctypes").c_ulonglong
    assert product.lib.qb200_getLaunchCount() > before


def test_reference_golden_vectors(product, product_env):
    """All 768 standard-format cases of the reference suite (3 qubits, every gate x target x start state)."""
    for case in CASES:
        harness.check_golden_case(case, harness.run_golden_case(product, product_env, case))


def test_reference_golden_outputs_bit_exact(product, product_env):
    n = 0
    for i, case in enumerate(CASES):
        if f"s{i}" not in REF_OUT.files:
            continue
        got = harness.run_golden_case(product, product_env, case)["S"]
        ref = REF_OUT[f"s{i}"]
        if case["func"] == "collapseToOutcome":
            np.testing.assert_allclose(got, ref, atol=1e-14, rtol=1e-14)
        else:
            assert np.array_equal(got.view(np.float64), ref.view(np.float64)), case["file"]
            n += 1
    assert n >= 500


def test_seed_known_answer(product):
    scripted.check_seed_known_answer(product)


@pytest.mark.parametrize("check", scripted.ALL_CHECKS, ids=lambda f: f.__name__)
def test_reference_scripted(product, product_env, check):
    check(product, product_env)


@pytest.mark.parametrize("name,nq,dens,depth,seed", [("sv12", 12, False, 120, 11), ("sv16", 16, False, 160, 12),
                                                     ("dm6", 6, True, 90, 13), ("dm8", 8, True, 120, 14)])
def test_reference_circuit_fixtures(product, product_env, name, nq, dens, depth, seed):
    """Committed outputs of the reference library for seeded circuits over every gate family."""
    ops = circuits.random_circuit(nq, depth, seed, density=dens)
    q = product.createDensityQureg(nq, product_env) if dens else product.createQureg(nq, product_env)
    circuits.prepare(product, q, seed)
    scal = circuits.run_circuit(product, q, ops, seed=seed)
    st = product.get_state(q)
    product.destroyQureg(q, product_env)
    np.testing.assert_allclose(scal, REF_CIRC[name + "_scalars"], rtol=REL, atol=1e-13)
    sub = st if st.size <= 1 << 16 else st[:: st.size >> 16]
    np.testing.assert_allclose(sub, REF_CIRC[name + "_state"], rtol=0, atol=REL * np.abs(REF_CIRC[name + "_state"]).max())
    np.testing.assert_allclose(circuits.state_digest(st), REF_CIRC[name + "_digest"], rtol=1e-11, atol=1e-12)


def _side_by_side(gpu, genv, cpu, cenv, nq, dens, ops, seed):
    out = []
    for lib, env in ((gpu, genv), (cpu, cenv)):
        q = lib.createDensityQureg(nq, env) if dens else lib.createQureg(nq, env)
        circuits.prepare(lib, q, seed)
        scal = circuits.run_circuit(lib, q, ops, seed=seed)
        out.append((lib.get_state(q), np.array(scal)))
        lib.destroyQureg(q, env)
    return out


@pytest.mark.parametrize("nq,dens,seed", [(1, False, 1), (2, False, 2), (3, False, 3), (5, False, 4), (9, False, 5),
                                          (13, False, 6), (18, False, 7), (22, False, 8), (1, True, 9), (2, True, 10),
                                          (4, True, 11), (7, True, 12), (10, True, 13), (11, True, 14)])
@pytest.mark.parametrize("fusion", [0, 1], ids=["gate_at_a_time", "fused"])
def test_bit_exact_vs_oracle_measurement_free(product, product_env, oracle, oracle_env, nq, dens, seed, fusion):
    """Every gate family, every target position (inside the vector, mid, top), 1 to 22 qubits: identical bits,
    both with one kernel per gate and through the fusion queue."""
    ops = circuits.random_circuit(nq, 220, seed, density=dens, measure=False)
    product.lib.qb200_setFusion(fusion)
    try:
        (gs, gv), (cs, cv) = _side_by_side(product, product_env, oracle, oracle_env, nq, dens, ops, seed)
    finally:
        product.lib.qb200_setFusion(1)
    assert np.array_equal(gs.view(np.float64), cs.view(np.float64))
    np.testing.assert_allclose(gv, cv, rtol=REL, atol=1e-13)


@pytest.mark.parametrize("nq,dens,seed", [(14, False, 21), (20, False, 22), (24, False, 23), (6, True, 24), (12, True, 25)])
def test_vs_live_reference_with_measurement(product, product_env, reference, reference_env, nq, dens, seed):
    """The reference's own CPU/OpenMP library side by side, measurements included: same outcomes, amplitudes to
    1e-12 relative."""
    ops = circuits.random_circuit(nq, 150, seed, density=dens, measure=True)
    (gs, gv), (cs, cv) = _side_by_side(product, product_env, reference, reference_env, nq, dens, ops, seed)
    np.testing.assert_allclose(gv, cv, rtol=REL, atol=1e-13)  # includes the integer measurement outcomes
    np.testing.assert_allclose(gs, cs, rtol=0, atol=REL * np.abs(cs).max())


@pytest.mark.parametrize("vec", ["1", "2", "4"])
def test_every_vector_width_is_bit_exact(vec, oracle, oracle_env):
    """QUEST_B200_VEC forces 8-, 16- and 32-byte vectors: each instantiation must give the same bits.
    Runs in a subprocess because the width is read once at createQuESTEnv."""
    import subprocess
    import sys

    code = (
        "import sys, numpy as np; sys.path.insert(0, %r)\n"
        "from quest_b200 import capi, circuits\n"
        "gpu = capi.load_product(); cpu = capi.load(capi.ORACLE_LIB)\n"
        "ge = gpu.createQuESTEnv(); ce = cpu.createQuESTEnv()\n"
        "for nq, dens, seed in [(3, False, 1), (11, False, 2), (3, True, 3), (6, True, 4)]:\n"
        "    ops = circuits.random_circuit(nq, 200, seed, density=dens, measure=False)\n"
        "    st = []\n"
        "    for lib, env in ((gpu, ge), (cpu, ce)):\n"
        "        q = lib.createDensityQureg(nq, env) if dens else lib.createQureg(nq, env)\n"
        "        circuits.prepare(lib, q, seed); circuits.run_circuit(lib, q, ops, seed=seed)\n"
        "        st.append(lib.get_state(q)); lib.destroyQureg(q, env)\n"
        "    assert np.array_equal(st[0].view(np.float64), st[1].view(np.float64)), (nq, dens)\n"
        "print('ok')\n" % os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
    env = dict(os.environ, QUEST_B200_VEC=vec, QUEST_B200_FUSE="0")  # one single-gate kernel per call
    r = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, env=env)
    assert r.returncode == 0 and "ok" in r.stdout, r.stdout + r.stderr


def test_many_controls_overflow_the_insertion_plan(product, product_env, oracle, oracle_env):
    """More than 7 controls: the extra ones are checked by predicate instead of index insertion."""
    nq = 14
    ops = [("multiControlledUnitary", list(range(1, 12)), 0, circuits._unitary(__import__("random").Random(3))),
           ("multiControlledPhaseShift", list(range(0, 12)), 0.37),
           ("multiControlledPhaseFlip", list(range(2, 14)))]
    ops = [("hadamard", t) for t in range(nq)] + ops
    for fusion in (0, 1):
        product.lib.qb200_setFusion(fusion)
        try:
            (gs, _), (cs, _) = _side_by_side(product, product_env, oracle, oracle_env, nq, False, ops, 5)
        finally:
            product.lib.qb200_setFusion(1)
        assert np.array_equal(gs.view(np.float64), cs.view(np.float64))


# ---- full-size properties (BASELINE configs are 30 qubits; the oracle cannot follow there) -------------------

def test_config2_properties_at_28_qubits(product, product_env):
    """Size-independent properties at a size the CPU cannot check element-wise in test time: unitarity (norm
    kept), H*H = identity bit pattern on a basis state, CNOT involution, P(0)+P(1)=1 on every qubit."""
    nq = 28
    q = product.createQureg(nq, product_env)
    product.initPlusState(q)
    for t in range(nq):
        product.compactUnitary(q, t, product.complex(circuits.SWEEP_ALPHA), product.complex(circuits.SWEEP_BETA))
    assert abs(product.calcTotalProb(q) - 1.0) < 1e-12
    probs = [product.calcProbOfOutcome(q, t, 0) for t in range(nq)]
    # each qubit was rotated independently from |+>: P(0) is the same closed form for every target
    a, b = circuits.SWEEP_ALPHA, circuits.SWEEP_BETA
    expect = abs((a - b.conjugate()) / np.sqrt(2)) ** 2
    np.testing.assert_allclose(probs, expect, rtol=1e-12)
    # involutions restore the exact state: X X, CNOT CNOT, and H H up to rounding
    before = [product.getAmp(q, i) for i in (0, 1, 12345, (1 << nq) - 1)]
    for t in (0, 1, 2, 13, nq - 1):
        product.pauliX(q, t); product.pauliX(q, t)
        product.controlledNot(q, (t + 1) % nq, t); product.controlledNot(q, (t + 1) % nq, t)
    after = [product.getAmp(q, i) for i in (0, 1, 12345, (1 << nq) - 1)]
    assert [(z.real, z.imag) for z in before] == [(z.real, z.imag) for z in after]
    product.destroyQureg(q, product_env)


def test_qft_of_basis_state_is_uniform(product, product_env):
    """configs[2] at 24 qubits: QFT of a computational basis state has |amp|^2 = 2^-n everywhere, so every
    per-qubit probability is exactly one half."""
    nq = 24
    q = product.createQureg(nq, product_env)
    product.initClassicalState(q, 0x2AAAAA & ((1 << nq) - 1))
    scal = circuits.run_circuit(product, q, circuits.config3_qft(nq))
    np.testing.assert_allclose(scal, 0.5, atol=1e-12)
    assert abs(product.calcTotalProb(q) - 1) < 1e-12
    product.destroyQureg(q, product_env)


# ---- gate fusion (deferred queue + k_tile) ----------------------------------------------------------------------

GATE_FAMILIES = circuits.DENSE_1Q + circuits.DIAG_1Q + circuits.CTRL_DENSE + circuits.CTRL_DIAG + circuits.MULTI


def _fusion_stats(lib):
    import ctypes
    a, b = ctypes.c_ulonglong(), ctypes.c_ulonglong()
    lib.lib.qb200_getFusionStats(ctypes.byref(a), ctypes.byref(b))
    return a.value, b.value


@pytest.mark.parametrize("nq,dens,seed", [(4, False, 31), (5, False, 32), (8, False, 33), (13, False, 34), (21, False, 35),
                                          (2, True, 36), (5, True, 37), (10, True, 38)])
def test_fused_long_gate_runs_bit_exact(product, product_env, oracle, oracle_env, nq, dens, seed):
    """300 gates with no read-back in between: the queue forms long groups; results must equal the oracle's bits."""
    ops = circuits.random_circuit(nq, 300, seed, density=dens, measure=False, families=GATE_FAMILIES)
    assert product.lib.qb200_getFusion() == 1
    before = _fusion_stats(product)
    (gs, _), (cs, _) = _side_by_side(product, product_env, oracle, oracle_env, nq, dens, ops, seed)
    after = _fusion_stats(product)
    assert np.array_equal(gs.view(np.float64), cs.view(np.float64))
    assert after[0] > before[0] and (after[1] - before[1]) > 0.8 * len([o for o in ops if o[0] in GATE_FAMILIES])


@pytest.mark.parametrize("nq,dens,seed", [(12, False, 41), (7, True, 42)])
def test_fusion_off_gives_the_same_bits(product, product_env, nq, dens, seed):
    ops = circuits.random_circuit(nq, 250, seed, density=dens, measure=False)
    states = []
    for mode in (1, 0):
        product.lib.qb200_setFusion(mode)
        q = product.createDensityQureg(nq, product_env) if dens else product.createQureg(nq, product_env)
        circuits.prepare(product, q, seed)
        circuits.run_circuit(product, q, ops, seed=seed)
        states.append(product.get_state(q))
        product.destroyQureg(q, product_env)
    product.lib.qb200_setFusion(1)
    assert np.array_equal(states[0].view(np.float64), states[1].view(np.float64))


@pytest.mark.parametrize("nq,seed", [(13, 51), (16, 52)])
def test_contracted_arithmetic_is_opt_in_and_within_tolerance(product, product_env, oracle, oracle_env, nq, seed):
    """qb200_setArithmetic(1): the fused pass may use fma.  Off by default; when on, amplitudes stay within the 1e-12
    bar of the oracle but are allowed to differ in the last bits; switching it off restores bit-identical results."""
    product.lib.qb200_setArithmetic.argtypes = [ctypes.c_int]
    assert product.lib.qb200_getArithmetic() == 0
    ops = circuits.random_circuit(nq, 300, seed, density=False, measure=False)
    cq = oracle.createQureg(nq, oracle_env)
    circuits.prepare(oracle, cq, seed)
    circuits.run_circuit(oracle, cq, ops, seed=seed)
    want = oracle.get_state(cq)
    oracle.destroyQureg(cq, oracle_env)
    got = {}
    try:
        for mode in (1, 0):
            product.lib.qb200_setArithmetic(mode)
            q = product.createQureg(nq, product_env)
            circuits.prepare(product, q, seed)
            circuits.run_circuit(product, q, ops, seed=seed)
            got[mode] = product.get_state(q)
            product.destroyQureg(q, product_env)
    finally:
        product.lib.qb200_setArithmetic(0)
    assert np.array_equal(got[0].view(np.float64), want.view(np.float64))
    assert np.abs(got[1] - want).max() <= 1e-12 * np.abs(want).max()
    assert not np.array_equal(got[1].view(np.float64), want.view(np.float64))  # the mode really changes the arithmetic


@pytest.mark.parametrize("nq,seed", [(5, 61), (13, 62), (20, 63)])
def test_all_qubit_probabilities_come_from_one_pass_and_follow_the_state(product, product_env, oracle, oracle_env, nq, seed):
    """The second calcProbOfOutcome on an unchanged state computes every qubit's probability in one pass; later queries
    launch nothing; anything that changes the state (gate, collapse, partial setAmps, re-initialisation) is seen."""
    product.lib.qb200_getLaunchCount.restype = ctypes.c_ulonglong
    ops = circuits.random_circuit(nq, 120, seed, density=False, measure=False)
    gq, cq = product.createQureg(nq, product_env), oracle.createQureg(nq, oracle_env)
    for lib, q in ((product, gq), (oracle, cq)):
        circuits.prepare(lib, q, seed)
        circuits.run_circuit(lib, q, ops, seed=seed)

    def sweep():
        for t in range(nq):
            for outcome in (0, 1):
                got, want = product.calcProbOfOutcome(gq, t, outcome), oracle.calcProbOfOutcome(cq, t, outcome)
                assert abs(got - want) <= 1e-12, (t, outcome, got, want)

    sweep()
    before = product.lib.qb200_getLaunchCount()
    sweep()
    assert product.lib.qb200_getLaunchCount() == before  # answered from the cached pass
    angle = float(np.random.default_rng(seed).uniform(0.1, 1.0))
    mutations = [
        lambda lib, q: lib.hadamard(q, nq // 2),
        lambda lib, q: lib.collapseToOutcome(q, 1, 0) if lib.calcProbOfOutcome(q, 1, 0) > 1e-6 else lib.pauliX(q, 1),
        lambda lib, q: lib.set_state(q, np.array([0.25 - 0.5j, 0.125j, -0.375]), start=3),
        lambda lib, q: lib.controlledPhaseShift(q, 0, nq - 1, 0.4),
        lambda lib, q: lib.initPlusState(q),
        lambda lib, q: lib.rotateY(q, nq - 1, angle),
    ]
    for m in mutations:
        m(product, gq)
        m(oracle, cq)
        sweep()
        sweep()
    product.destroyQureg(gq, product_env)
    oracle.destroyQureg(cq, oracle_env)


def test_queue_is_drained_by_every_observer(product, product_env, oracle, oracle_env):
    """Pending gates must be visible to clone, inner product, compare, collapse, noise, setAmps and getAmp."""
    nq = 9
    res = []
    for lib, env in ((product, product_env), (oracle, oracle_env)):
        a, b = lib.createQureg(nq, env), lib.createQureg(nq, env)
        circuits.prepare(lib, a, 5)
        for t in range(nq):
            lib.hadamard(a, t); lib.rotateY(a, t, 0.1 * t); lib.controlledNot(a, t, (t + 1) % nq)
        lib.cloneQureg(b, a)                       # must see a's pending gates
        lib.pauliX(b, 3); lib.tGate(b, 4)
        ip = lib.calcInnerProduct(a, b)            # must see b's pending gates
        lib.hadamard(a, 0)
        amp = lib.getAmp(a, 5)
        lib.rotateX(a, 2, 0.4)
        lib.set_state(a, np.array([0.5 + 0.25j]), start=7)   # ordered after the pending rotateX
        lib.phaseShift(a, 1, 0.2)
        same = lib.compareStates(a, b, 1e-3)
        p = lib.collapseToOutcome(a, 4, 1)
        out = [ip.real, ip.imag, amp.real, amp.imag, same, p]
        res.append((np.array(out), lib.get_state(a), lib.get_state(b)))
        lib.destroyQureg(a, env); lib.destroyQureg(b, env)
    np.testing.assert_allclose(res[0][0], res[1][0], rtol=1e-12, atol=1e-14)
    np.testing.assert_allclose(res[0][1], res[1][1], rtol=0, atol=1e-13)
    assert np.array_equal(res[0][2].view(np.float64), res[1][2].view(np.float64))


def test_32_qubit_register_uses_64_bit_indexing(product, product_env):
    """64 GiB of amplitudes: indices above 2^31 are reached by gates on the top qubits, reads and reductions.
    Closed forms: rotations from |0...0> and |+...+> give product states."""
    nq = 32
    q = product.createQureg(nq, product_env)
    assert q.numAmpsTotal == 1 << 32
    th = 0.7
    product.rotateY(q, nq - 1, th)            # top qubit
    product.rotateY(q, 0, 2 * th)             # lowest qubit
    product.controlledNot(q, nq - 1, nq - 2)  # entangle the two top qubits
    c, s = np.cos(th / 2), np.sin(th / 2)
    c0, s0 = np.cos(th), np.sin(th)
    # |psi> = (c|00> + s|11>)_{31,30} x (c0|0> + s0|1>)_0
    for idx, want in [(0, c * c0), (1, c * s0), ((3 << 30), s * c0), ((3 << 30) | 1, s * s0), (1 << 31, 0.0), (1 << 30, 0.0)]:
        amp = product.getAmp(q, idx)
        assert abs(amp.real - want) < 1e-15 and amp.imag == 0.0, (idx, amp.real, want)
    np.testing.assert_allclose(product.calcProbOfOutcome(q, nq - 1, 1), s * s, rtol=1e-13)
    np.testing.assert_allclose(product.calcProbOfOutcome(q, nq - 2, 1), s * s, rtol=1e-13)
    np.testing.assert_allclose(product.calcProbOfOutcome(q, 0, 1), s0 * s0, rtol=1e-13)
    np.testing.assert_allclose(product.calcTotalProb(q), 1.0, rtol=1e-13)
    # a long fused run over low, middle and top qubits keeps the norm
    product.initPlusState(q)
    for t in (0, 1, 5, 17, 30, 31, 2, 29):
        product.hadamard(q, t)
        product.compactUnitary(q, t, product.complex(circuits.SWEEP_ALPHA), product.complex(circuits.SWEEP_BETA))
        product.controlledPhaseShift(q, t, (t + 7) % nq, 0.3)
    np.testing.assert_allclose(product.calcTotalProb(q), 1.0, rtol=1e-12)
    # collapse of the top qubit
    p = product.collapseToOutcome(q, nq - 1, 1)
    assert 0 < p < 1
    np.testing.assert_allclose(product.calcTotalProb(q), 1.0, rtol=1e-12)
    assert product.calcProbOfOutcome(q, nq - 1, 0) < 1e-15
    product.destroyQureg(q, product_env)


@pytest.mark.parametrize("nq,seed", [(4, 401), (6, 402), (9, 403), (11, 404)])
def test_api_fuzz_vs_oracle(product, product_env, oracle, oracle_env, nq, seed):
    """Random multi-register API sequences - gates interleaved with clone, inner product, fidelity, compare, mixing,
    partial setAmps, re-initialisation, probabilities, getAmp - through the fusion queue, against the CPU oracle."""
    calls = circuits.random_api_sequence(nq, 400, seed, extended=True)
    a, sa = circuits.run_api_sequence(product, product_env, nq, calls, seed)
    b, sb = circuits.run_api_sequence(oracle, oracle_env, nq, calls, seed)
    np.testing.assert_allclose(a, b, rtol=1e-12, atol=1e-13)
    for k in sa:
        np.testing.assert_allclose(sa[k], sb[k], rtol=0, atol=1e-13 * max(1.0, np.abs(sb[k]).max()))


# ---- opt-in gate reordering (qb200_setReordering): commuting gates may change places --------------------------------

@pytest.mark.parametrize("nq,dens,seed", [(13, False, 31), (18, False, 32), (22, False, 33), (9, True, 34)])
def test_reordering_mode_stays_within_tolerance(product, product_env, oracle, oracle_env, nq, dens, seed):
    """With reordering on, gates move forward past gates they commute with (fewer passes).  That changes the order of
    roundings, so the bar is the north_star tolerance (1e-12 relative), not bit-identity; the default mode, checked by
    every other test of this file, stays bit-identical."""
    import ctypes as C

    product.lib.qb200_setReordering.argtypes = [C.c_int]
    ops = circuits.random_circuit(nq, 260, seed, density=dens, measure=False)
    # a layered tail (one kind of gate on every qubit at a time): the shape reordering is meant for
    for l in range(3):
        ops += [("hadamard", i) for i in range(nq)] + [("rotateY", i, 0.1 * (i + 1) + 0.01 * l) for i in range(nq)]
        ops += [("controlledNot", i, (i + 1) % nq) for i in range(nq)]
        if dens:
            ops += [("applyOneQubitDepolariseError", i, 0.03) for i in range(nq)]
    product.lib.qb200_setReordering(1)
    try:
        (gs, gv), (cs, cv) = _side_by_side(product, product_env, oracle, oracle_env, nq, dens, ops, seed)
    finally:
        product.lib.qb200_setReordering(0)
    np.testing.assert_allclose(gv, cv, rtol=REL, atol=1e-13)
    np.testing.assert_allclose(gs, cs, rtol=0, atol=REL * np.abs(cs).max())


def test_pass_mixing_every_gate_family_picks_the_widest_kernel_tier(product, product_env, oracle, oracle_env):
    """One pass that holds gates of all three kernel tiers (common gates; pauliY / phase flip; density-matrix noise):
    it has to run on the tier that understands all of them."""
    nq = 6
    ops = []
    for t in range(nq):
        ops += [("hadamard", t), ("pauliY", t), ("controlledPhaseFlip", t, (t + 1) % nq), ("applyOneQubitDepolariseError", t, 0.1),
                ("rotateX", t, 0.3 + t), ("controlledPauliY", (t + 2) % nq, t), ("applyOneQubitDampingError", t, 0.05),
                ("multiControlledPhaseFlip", [t, (t + 1) % nq, (t + 3) % nq]), ("applyOneQubitDephaseError", t, 0.07)]
    (gs, gv), (cs, cv) = _side_by_side(product, product_env, oracle, oracle_env, nq, True, ops + [("calcPurity",)], 3)
    assert np.array_equal(gs.view(np.float64), cs.view(np.float64))
    np.testing.assert_allclose(gv, cv, rtol=REL, atol=1e-13)
